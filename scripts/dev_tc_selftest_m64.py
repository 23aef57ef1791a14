"""Developer aid: where do the 64 accumulator rows of a tcgen05.mma with M = 64 (cta_group::1) live in TMEM?
Runs sat_tc_selftest with SAT_SELFTEST_M=64 and matches every dumped (lane, column) against the fp64 product."""
import ctypes as C, os, sys
os.environ["SAT_SELFTEST_M"] = "64"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200
from sat_b200._lib import lib, check
K = 64
rng = np.random.default_rng(1)
A = rng.normal(0, 1, size=(128, K)).astype(np.float32)
Wt = rng.normal(0, 1, size=(16, K)).astype(np.float32)
dA, dW = torch.from_numpy(A).cuda(), torch.from_numpy(Wt).cuda()
D = torch.full((128, 16), float("nan"), device="cuda")
check(lib().sat_tc_selftest(C.c_void_p(dA.data_ptr()), C.c_void_p(dW.data_ptr()), K, C.c_void_p(D.data_ptr()),
                            C.c_void_p(torch.cuda.current_stream().cuda_stream)))
torch.cuda.synchronize()
ref = A[:64].astype(np.float64) @ Wt.astype(np.float64).T
got = D.cpu().numpy()
lane_of_row = {}
for lane in range(128):
    d = np.abs(ref - got[lane][None, :]).max(axis=1)
    r = int(d.argmin())
    if d[r] < 2e-2:
        lane_of_row[r] = lane
print("rows found:", len(lane_of_row))
print("lane of row r:", [lane_of_row.get(r, -1) for r in range(64)])
