// gemm.cu -- fp32 SIMT row-major GEMM / implicit conv1d with fused epilogue.
//
// Used for every time-batched dense contraction of the path that is not (yet) on
// the tcgen05 kernel: encoder prenet (with the embedding gather fused on load,
// models/models.py:28,54), CBHG conv bank / projections (tacotron2 Conv1d = conv ->
// BN(inference) -> act, folded into scale/shift; modules/module.py:46-68 with the
// MaxPooling1D(2,1,SAME) of :54 fused on load), highway layers, attention memory
// layers (TF BahdanauAttention keys/values, A.5), post-net convs (A.8).
//
// Tiling: 128x64 output tile per CTA, 256 threads, 8x4 accumulators per thread,
// K chunks of 16 staged through shared memory; coalesced float4 global loads when
// the operand is 16-byte aligned.
#include "common.cuh"

namespace sat {

namespace {
constexpr int BM = 128, BN = 64, BK = 16, NT = 256;

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case ACT_RELU: return fmaxf(v, 0.f);
    case ACT_TANH: return tanhf(v);
    case ACT_SIGMOID: return 1.f / (1.f + expf(-v));
    case ACT_SOFTSIGN: return v / (1.f + fabsf(v));
    default: return v;
  }
}

struct GemmDev {
  const float* X; int Bn, T, Cin, ldx;
  const float* W; int taps, pad_left, N;
  const float* scale; const float* shift; int act;
  const float* R; int ldr;
  float* C; int ldc;
  const int64_t* row_lengths; const int64_t* gather_ids; long long gather_offset;
  long long gather_rows; int* err_flag;
  int maxpool2; int vecA; int vecB; long long M;
  int highway;
};

// source row pointer for global output row g at tap; nullptr when the tap falls in the zero padding
__device__ __forceinline__ const float* src_row(const GemmDev& p, long long g, int tap, const float** next_row) {
  *next_row = nullptr;
  if (g >= p.M) return nullptr;
  if (p.gather_ids) {
    long long r = p.gather_ids[g] - p.gather_offset;
    if (p.gather_rows > 0 && (r < 0 || r >= p.gather_rows)) {  // id outside the table: zero row + sticky error flag
      if (p.err_flag) *p.err_flag = 1;
      return nullptr;
    }
    return p.X + r * p.ldx;
  }
  int b = (int)(g / p.T), t = (int)(g % p.T) + tap - p.pad_left;
  if (t < 0 || t >= p.T) return nullptr;
  if (p.row_lengths && t >= p.row_lengths[b]) return nullptr;
  const float* row = p.X + ((long long)b * p.T + t) * p.ldx;
  if (p.maxpool2 && t + 1 < p.T && !(p.row_lengths && t + 1 >= p.row_lengths[b])) *next_row = row + p.ldx;
  return row;
}

__global__ void __launch_bounds__(NT) gemm_kernel(GemmDev p) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN];
  const int tid = threadIdx.x;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int tx = tid % 16, ty = tid / 16;  // thread tile: rows ty*8..+7, cols tx*4..+3
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int a_r = tid / 4, a_k = (tid % 4) * 4;   // A loader: rows a_r and a_r+64, k a_k..a_k+3
  const int b_k = tid / 16, b_n = (tid % 16) * 4;  // B loader

  for (int tap = 0; tap < p.taps; ++tap) {
    const float* nrow0; const float* nrow1;
    const float* row0 = src_row(p, m0 + a_r, tap, &nrow0);
    const float* row1 = src_row(p, m0 + a_r + 64, tap, &nrow1);
    const float* Wt = p.W + (long long)tap * p.Cin * p.N;
    for (int k0 = 0; k0 < p.Cin; k0 += BK) {
      // ---- stage A (transposed into As[k][m]) ----
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float* row = h ? row1 : row0;
        const float* nrow = h ? nrow1 : nrow0;
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        int k = k0 + a_k;
        if (row) {
          if (p.vecA && k + 3 < p.Cin) {
            float4 q = __ldg((const float4*)(row + k));
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
            if (nrow) {
              float4 r = __ldg((const float4*)(nrow + k));
              v[0] = fmaxf(v[0], r.x); v[1] = fmaxf(v[1], r.y); v[2] = fmaxf(v[2], r.z); v[3] = fmaxf(v[3], r.w);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (k + i < p.Cin) {
                v[i] = __ldg(row + k + i);
                if (nrow) v[i] = fmaxf(v[i], __ldg(nrow + k + i));
              }
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) As[a_k + i][a_r + 64 * h] = v[i];
      }
      // ---- stage B ----
      {
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        int k = k0 + b_k, n = n0 + b_n;
        if (k < p.Cin) {
          const float* w = Wt + (long long)k * p.N + n;
          if (p.vecB && n + 3 < p.N) {
            float4 q = __ldg((const float4*)w);
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (n + i < p.N) v[i] = __ldg(w + i);
          }
        }
        *(float4*)&Bs[b_k][b_n] = make_float4(v[0], v[1], v[2], v[3]);
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        float4 a0 = *(const float4*)&As[k][ty * 8];
        float4 a1 = *(const float4*)&As[k][ty * 8 + 4];
        float4 b = *(const float4*)&Bs[k][tx * 4];
        float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
  // ---- epilogue ----
  if (p.highway) {
    // HighwayNet (tacotron2): columns are interleaved (2u = H_u, 2u+1 = T_u);
    // y_u = relu(H) * sigmoid(T) + x_u * (1 - sigmoid(T)), x = R
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      long long g = m0 + ty * 8 + i;
      if (g >= p.M) continue;
#pragma unroll
      for (int j = 0; j < 4; j += 2) {
        int n = n0 + tx * 4 + j;
        if (n + 1 >= p.N) continue;
        const int u = n >> 1;
        const float hh = fmaxf(acc[i][j] + __ldg(p.shift + n), 0.f);
        const float tt = 1.f / (1.f + expf(-(acc[i][j + 1] + __ldg(p.shift + n + 1))));
        const float x = p.R[g * p.ldr + u];
        p.C[g * p.ldc + u] = hh * tt + x * (1.f - tt);
      }
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    long long g = m0 + ty * 8 + i;
    if (g >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float v = acc[i][j];
      if (p.scale) v *= __ldg(p.scale + n);
      if (p.shift) v += __ldg(p.shift + n);
      v = apply_act(v, p.act);
      if (p.R) v += p.R[g * p.ldr + n];
      p.C[g * p.ldc + n] = v;
    }
  }
}
}  // namespace

void gemm_launch(const GemmArgs& a, cudaStream_t st) {
  GemmDev p;
  p.X = a.X; p.Bn = a.Bn; p.T = a.T; p.Cin = a.Cin; p.ldx = a.ldx;
  p.W = a.W; p.taps = a.taps; p.pad_left = a.pad_left; p.N = a.N;
  p.scale = a.scale; p.shift = a.shift; p.act = a.act;
  p.R = a.R; p.ldr = a.ldr; p.C = a.C; p.ldc = a.ldc;
  p.row_lengths = a.row_lengths; p.gather_ids = a.gather_ids; p.gather_offset = a.gather_offset;
  p.gather_rows = a.gather_rows; p.err_flag = a.err_flag;
  p.maxpool2 = a.maxpool2;
  p.highway = a.highway;
  p.M = (long long)a.Bn * a.T;
  p.vecA = ((a.ldx % 4) == 0) && ((((uintptr_t)a.X) & 15) == 0);
  p.vecB = ((a.N % 4) == 0) && ((((uintptr_t)a.W) & 15) == 0);
  if (p.M == 0 || a.N == 0) return;
  dim3 grid((unsigned)ceil_div64(p.M, BM), (unsigned)ceil_div(a.N, BN));
  gemm_kernel<<<grid, NT, 0, st>>>(p);
  count_launch();
  check_launch("gemm_kernel");
}

}  // namespace sat
