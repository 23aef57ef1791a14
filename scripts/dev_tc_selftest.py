"""Developer aid: run sat_tc_selftest on the GPU and compare with fp64 numpy (TF32 tolerance)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200
from sat_b200._lib import lib, check
for K in (32, 64, 640):
    rng = np.random.default_rng(K)
    A = rng.normal(0, 1, size=(128, K)).astype(np.float32)
    Wt = rng.normal(0, 1, size=(16, K)).astype(np.float32)
    dA, dW = torch.from_numpy(A).cuda(), torch.from_numpy(Wt).cuda()
    D = torch.full((128, 16), float("nan"), device="cuda")
    check(lib().sat_tc_selftest(C.c_void_p(dA.data_ptr()), C.c_void_p(dW.data_ptr()), K, C.c_void_p(D.data_ptr()),
                                C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    ref = A.astype(np.float64) @ Wt.astype(np.float64).T
    got = D.cpu().numpy()
    err = np.abs(got - ref).max() / np.abs(ref).max()
    print(f"K={K}: rel err {err:.3e}  nan={np.isnan(got).sum()}  got[0,:4]={got[0,:4]} ref[0,:4]={ref[0,:4]}")
    if err > 5e-3:
        # help locate layout bugs: which reference entry does each output match?
        for (i, j) in [(0, 0), (1, 0), (0, 1), (8, 0), (32, 0), (64, 5)]:
            d = np.abs(ref - got[i, j])
            print("  got", (i, j), got[i, j], "closest ref at", np.unravel_index(d.argmin(), d.shape), d.min())
