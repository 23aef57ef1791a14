// decoder_dev.cuh -- device-side building blocks of the persistent decoder kernel.
//
// Data layout in HBM (all fp32):
//   * recurrent activations are FEATURE-MAJOR  X[k][ld]  (ld = batch padded to 4): a warp
//     reads 128 consecutive batch rows of one feature with one coalesced LDG.128 per lane;
//   * weights are packed once at create time into column tiles of 8: Wp[tile][k][8]
//     (LSTM tiles hold 2 units x 4 gates i,j,f,o so the cell update is thread-local);
//     LSTM tiles owned by a CTA stay resident in its shared memory for the whole run;
//   * attention memory (keys/values) and per-row attention state are ROW-MAJOR [B][T][U].
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace sat {
namespace dec {

constexpr int NT = 512;        // threads per CTA
constexpr int NW = NT / 32;    // warps per CTA
constexpr int TILE_N = 8;      // output columns per tile
constexpr int ROWS_PER_GROUP = 128;
constexpr int RED_FLOATS = NW * TILE_N * ROWS_PER_GROUP;  // cross-warp reduction scratch (64 KB)
constexpr int MAX_HOPS = 2;
constexpr int MAX_B = 256;

// element (feature n, utterance row b) of a tensor-core activation buffer (B >= 64 form; fp16, `kc` 64-k chunks per row):
// index in halves into [row block of 128][chunk][128 rows][64 k], every [128][64] tile in the 128-byte-swizzled order
// the tcgen05 A operand has in shared memory -- a row block is then one contiguous run that bulk copies move verbatim
__device__ __forceinline__ size_t tc_act_index(int n, int b, int kc) {
  const int r = b & 127, kk = n & 63;
  return ((size_t)((b >> 7) * kc + (n >> 6)) << 13) + (size_t)(r * 64 + ((((kk >> 3) ^ (r & 7)) << 3) | (kk & 7)));
}

struct DenseDesc {
  const float* Wp;    // [n_tiles][K][8] packed (global)
  const float* bias;  // [n_tiles*8] packed (zeros where absent)
  int K, N, n_tiles;
  int smem_off;       // float offset of this layer's resident slots in dynamic smem, -1 = streamed
  int slots;          // resident slots per CTA
  const float* Wrm;   // row-major copy [K][Np] (Np = N padded to 4) for the row-split frame phase; may be null
  int Np;
  const void* Wrm16;  // the same in fp16 [K][Np8] (Np8 = N padded to 8), used in tensor-core mode
  int Np8;
  int reps;           // both row-major copies exist `reps` times back to back (rep_stride / rep_stride16 elements apart):
  size_t rep_stride, rep_stride16;  // CTAs read replica blockIdx % reps so that the broadcast read of a small weight
                                    // set spreads over more L2 slices
  int rpg, R;         // rows per work item (32 / 64 / 128) and row groups = ceil(Bp / rpg): small layers split rows
                      // finer so that n_tiles * R work items cover the grid
};

// ---- small math helpers (accurate versions: parity first) ------------------
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// exp2 / reciprocal on the MUFU pipe
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Gate non-linearities of the LSTM cell updates: 1 / (1 + 2^(-x log2 e)) and 1 - 2 / (2^(2 x log2 e) + 1) on the MUFU
// pipe (abs error ~2e-7, saturate cleanly at +-inf) -- a fifth of the instructions of expf / tanhf / IEEE division on
// the critical path behind every gate GEMM.  Decisions (stop rule) keep the accurate sigmoidf_.
__device__ __forceinline__ float sigmoid_fast(float x) { return rcp_approx(1.f + ex2_approx(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanh_fast(float x) {
  return fmaf(-2.f, rcp_approx(ex2_approx(2.8853900817779268f * x) + 1.f), 1.f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide reductions through a 32-float scratch (all NT threads must call)
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x & 31) < NW ? scratch[threadIdx.x & 31] : 0.f;
  return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x & 31) < NW ? scratch[threadIdx.x & 31] : -INFINITY;
  return warp_max(r);
}

// ---- grid-wide barrier -------------------------------------------------------
// Monotonic arrival counter in global memory; barrier #n completes when the counter
// reaches n * gridDim.x: bar.sync, then one thread does red.release.gpu + ld.acquire.gpu spin, bar.sync
// (release / acquire at gpu scope instead of the membar.gl pair of cooperative-groups grid.sync()).
// Requires all CTAs co-resident (cooperative launch).  Measured on B200: 1.49 us per empty barrier; neither the
// membar.gl -> release/acquire change nor spreading the arrivals over four counters moved that number.
struct GridBarrier {
  unsigned* counter;
  unsigned gen;
  __device__ __forceinline__ void sync() {
    sync([] {});
  }
  // `between` runs on thread 0 after this CTA has arrived and before it starts waiting for the others: work that only
  // needs data from earlier phases (e.g. TMA prefetch for the next phase) hides behind the barrier latency
  template <class F>
  __device__ __forceinline__ void sync(F&& between) {
    __syncthreads();
    if (threadIdx.x == 0) {
      gen += 1;
      const unsigned target = gen * gridDim.x;
      // release at gpu scope: cumulative over the CTA's writes ordered before it by the bar.sync above
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
      between();
      unsigned seen;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
      } while (seen < target);
    }
    __syncthreads();
  }
};

// ---- the dense-tile primitive -------------------------------------------------
// For every work item (tile, row-group) owned by this CTA:
//   g[b][c] = sum_k X[k][b] * W[tile][k][c]      (b in the row group, c in 0..7)
// then epi(b, tile, ul, g4) is called once per (row, half-tile) with the 4 reduced
// pre-bias sums of columns tile*8 + ul*4 .. +3.
// Items are dealt round-robin: item = tile * R + rg -> CTA item % gridDim.x; the j-th
// item of a CTA uses resident slot j when the layer is resident.
template <class Epi>
__device__ void dense_phase(const DenseDesc& d, const float* X, int ld, int Bp, float* smem, float* red, Epi epi) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = d.R;
  const int n_items = d.n_tiles * R;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++slot) {
    const int tile = item / R, rg = item % R;
    const int r0 = rg * d.rpg;
    const int nrows = min(d.rpg, Bp - r0);
    const int nq = nrows >> 2;
    int qpw = 1;
    while (qpw < nq) qpw <<= 1;  // quads per warp (power of two, <= 32)
    const int ksub = 32 / qpw;
    const int q = lane & (qpw - 1), ks = lane / qpw;
    const float* W = (d.smem_off >= 0 && slot < d.slots)
                         ? smem + d.smem_off + (size_t)slot * d.K * TILE_N
                         : d.Wp + (size_t)tile * d.K * TILE_N;
    float acc[4][TILE_N];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < TILE_N; ++c) acc[i][c] = 0.f;
    if (q < nq) {
      const float* xp = X + r0 + 4 * q;
      const int kstride = NW * ksub;
      // batches of PF activation loads in flight (one L2 round trip per batch; K = 672 at 16 rows is one batch)
      constexpr int PF = 8;
      for (int kb = warp * ksub + ks; kb < d.K; kb += PF * kstride) {
        float4 xq[PF];
#pragma unroll
        for (int s = 0; s < PF; ++s) {
          const int k = kb + s * kstride;
          xq[s] = k < d.K ? __ldcg((const float4*)(xp + (size_t)k * ld)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int s = 0; s < PF; ++s) {
          const int k = kb + s * kstride;
          if (k >= d.K) break;
          const float4 w0 = *(const float4*)(W + k * TILE_N);
          const float4 w1 = *(const float4*)(W + k * TILE_N + 4);
          const float xv[4] = {xq[s].x, xq[s].y, xq[s].z, xq[s].w};
          const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int c = 0; c < TILE_N; ++c) acc[i][c] = fmaf(xv[i], wv[c], acc[i][c]);
        }
      }
    }
    // reduce over the k-sub lanes of the warp
    for (int o = qpw; o < 32; o <<= 1) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < TILE_N; ++c) acc[i][c] += __shfl_xor_sync(0xffffffffu, acc[i][c], o);
    }
    if (ks == 0 && q < nq) {
#pragma unroll
      for (int c = 0; c < TILE_N; ++c)
        *(float4*)(red + ((warp * TILE_N + c) * ROWS_PER_GROUP) + 4 * q) =
            make_float4(acc[0][c], acc[1][c], acc[2][c], acc[3][c]);
    }
    __syncthreads();
    // cross-warp reduction + epilogue: thread <-> (row, half tile)
    for (int e = threadIdx.x; e < nrows * 2; e += NT) {
      const int row = e % nrows, ul = e / nrows;
      float g[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int w = 0; w < NW; ++w)
#pragma unroll
        for (int c = 0; c < 4; ++c) g[c] += red[(w * TILE_N + ul * 4 + c) * ROWS_PER_GROUP + row];
      epi(r0 + row, tile, ul, g);
    }
    __syncthreads();
  }
}

// copy this CTA's resident tiles of a layer into shared memory (once, before the loop)
__device__ inline void load_resident(const DenseDesc& d, float* smem) {
  if (d.smem_off < 0) return;
  const int R = d.R;
  const int n_items = d.n_tiles * R;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items && slot < d.slots; item += gridDim.x, ++slot) {
    const int tile = item / R;
    const float4* src = (const float4*)(d.Wp + (size_t)tile * d.K * TILE_N);
    float4* dst = (float4*)(smem + d.smem_off + (size_t)slot * d.K * TILE_N);
    for (int i = threadIdx.x; i < d.K * TILE_N / 4; i += NT) dst[i] = __ldg(src + i);
  }
}

}  // namespace dec
}  // namespace sat
