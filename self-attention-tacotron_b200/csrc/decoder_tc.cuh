// decoder_tc.cuh -- tensor-core form of the ZoneoutLSTM gate GEMMs of the persistent decoder kernel, used once the
// utterance batch makes them a real contraction (B >= 64; BASELINE.json north_star).
//
// Work split: a CTA owns (column tile of 16 gate columns = 4 units x {i,j,f,o}) x (row block of 128 utterances).
// The W^T slice [16][K] of every LSTM stays resident in shared memory for the whole run as a K-major, 128-byte
// swizzled fp16 image; the activations X[128 rows][K] (row-major fp16 in HBM/L2) stream through a 6-stage ring of
// 16 KB TMA boxes {64 k, 128 rows}; one elected thread issues tcgen05.mma kind::f16 (M=128, N=16, K=16) into
// a 16-column TMEM accumulator; all sixteen warps read it back (tcgen05.ld 32x32b.x4 -> one utterance row and one
// complete unit per thread) and apply the zoneout cell update in registers.
// Operands are fp16 with fp32 accumulation: fp16 has the 11-bit significand of tf32 at half the bytes, and the
// phase is bound by the L2 -> shared-memory stream of X.
#pragma once
#include "decoder_dev.cuh"
#include "tc.cuh"

namespace sat {
namespace dec {

constexpr int TC_STAGES = 6;
constexpr int TC_STAGE_BYTES = 128 * tc::CHUNK_K16 * 2;  // 16 KB: 128 rows x 64 fp16
constexpr int TC_N = 16;                                 // gate columns per tile
constexpr int TC_M = 128;                                // utterance rows per tile
constexpr int TC_STATE_BYTES = 2 * 4 * TC_M * 4;         // per resident item: c and zoned h of 4 units x 128 rows (fp32)

// element (feature n, row b) of an activation buffer
struct Act2D {
  void* p;
  int sd, sb;  // strides of the feature / row index (feature-major: sd = ld, sb = 1; row-major: sd = 1, sb = K)
  int f16;     // 1: fp16 buffer feeding a tensor-core GEMM (values rounded to nearest), 0: fp32
  __device__ __forceinline__ void put(int n, int b, float v) const {
    const size_t i = (size_t)n * sd + (size_t)b * sb;
    if (f16) ((__half*)p)[i] = tc::to_f16(v);
    else ((float*)p)[i] = v;
  }
};

struct TcLstmDesc {
  const float* Wp;    // packed tiles [units/2][K][8] (same packing as the SIMT path; global)
  const float* bias;  // [units/2 * 8]
  int K, units;
  int n_tiles;        // units / 4 column tiles of 16
  int img_off;        // byte offset of this layer's resident images inside the TC weight region
  int st_off;         // byte offset of this layer's resident cell state inside the TC weight region
  int slots;          // resident (tile,row block) items per CTA
  int kchunks;        // ceil(K / 64)
  int n_pre;          // leading chunks of `order` whose columns were produced at least two phases ago: their TMA
                      // loads are issued while the CTA waits at the grid barrier in front of the phase
  unsigned char order[48];  // chunk visiting order (a permutation of 0 .. kchunks-1), "old" columns first
};

// per-CTA pipeline state that persists across phases and decoder steps
struct TcState {
  uint8_t* ring;       // TC_STAGES x 16 KB, 1024-aligned (aliases the phase scratch)
  uint8_t* wimg;       // resident weight images + cell state, 1024-aligned
  uint64_t* full;      // [TC_STAGES]
  uint64_t* empty;     // [TC_STAGES]
  uint64_t* done;
  uint32_t tmem;
  uint32_t n_chunks;   // chunks pushed through the ring so far (identical in every thread)
  uint32_t n_items;    // accumulator hand-overs so far
  uint32_t prefilled;  // chunks of the next LSTM phase already requested by tc_prefill (thread 0 only)
  long long* prof;     // optional cycle counters (CTA 0): [0] producer loop, [1] wait for the accumulator, [2] cell update
};

// build this CTA's resident K-major SW128 fp16 image of one LSTM (chunk c of slot s at wimg + img_off +
// (s*kchunks + c)*2 KB) and zero its cell state
__device__ inline void tc_load_resident(const TcLstmDesc& d, int RB, const TcState& st) {
  if (d.slots == 0) return;
  const int n_items = d.n_tiles * RB;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items && slot < d.slots; item += gridDim.x, ++slot) {
    const int tile = item / RB;
    uint8_t* img = st.wimg + d.img_off + (size_t)slot * d.kchunks * 2048;
    // column n of the 16-wide tile = column (n & 7) of packed 8-wide tile 2*tile + (n >> 3)
    for (int i = threadIdx.x; i < d.kchunks * tc::CHUNK_K16 * TC_N; i += NT) {
      const int n = i & (TC_N - 1), k = i >> 4;
      float w = 0.f;
      if (k < d.K) w = __ldg(d.Wp + ((size_t)(2 * tile + (n >> 3)) * d.K + k) * TILE_N + (n & 7));
      *(__half*)(img + (k / tc::CHUNK_K16) * 2048 + tc::sw128_offset16(n, k % tc::CHUNK_K16)) = tc::to_f16(w);
    }
  }
  float* state = (float*)(st.wimg + d.st_off);
  for (int i = threadIdx.x; i < d.slots * (TC_STATE_BYTES / 4); i += NT) state[i] = 0.f;
}

// thread 0, between arriving at and waiting on the grid barrier in front of an LSTM phase: request the first chunks
// of the CTA's first item (columns that do not depend on the phase that is just finishing)
__device__ inline void tc_prefill(const TcLstmDesc& d, const CUtensorMap* xmap, int RB, TcState& st) {
  const int item = blockIdx.x;
  if (item >= d.n_tiles * RB || d.n_pre == 0) return;
  const int row0 = (item % RB) * TC_M;
  tc::fence_proxy_async();
  for (int c = 0; c < d.n_pre; ++c) {
    const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
    if (use > 0) tc::mbar_wait(st.empty + s, (use - 1) & 1);
    tc::mbar_arrive_expect_tx(st.full + s, TC_STAGE_BYTES);
    tc::tma_load_2d(st.ring + s * TC_STAGE_BYTES, xmap, st.full + s, d.order[c] * tc::CHUNK_K16, row0);
  }
  st.prefilled = d.n_pre;
}

struct TcLstmIO {
  const CUtensorMap* xmap;  // activations of this step: row-major fp16 [rows][Kpad]
  Act2D out0, out1;         // un-zoned cell output (written; either .p may be null)
  Act2D hz;                 // zoned h for the next step (written)
  float* rm_out;            // optional row-major fp32 copy of the output: rm_out[b * rm_stride + rm_off + u]
  int rm_stride, rm_off;
  __half* hl;               // optional fp16 hi + lo copy of the output (the input of a dense_phase_tc layer):
  int hl_stride, hl_lo;     // hl[b * hl_stride + u] = f16(h), hl[b * hl_stride + hl_lo + u] = f16(h - f16(h))
};

// ---- a dense layer of the step on the tensor cores (B >= 64): query layers, frame / stop projection, prenet -------
// y[rows, N] = x[rows, K] W[K, N] as tiles of 16 columns x 128 rows, the CTA's W^T slice resident in shared memory
// (K-major SW128 fp16 image, built once from the fp32 row-major weights).  fp32-grade accuracy from fp16 operands:
// the activations arrive as an fp16 pair x = xh + xl ([rows][2 * kc * 64]: hi columns, then lo columns) and the GEMM
// runs the terms xh Wh + xl Wh (+ xh Wl when terms = 3, Wl = f16(W - Wh)) as one K loop of terms * kc chunks.
struct TcDense {
  const float* W[2];   // up to two column groups (the query layers of two mechanisms): fp32 row-major [K][ldw]
  int ldw[2], nt[2];   // row stride and 16-column tiles of each group
  int ncol[2];         // valid columns of each group (columns beyond read as zero)
  const float* bias;   // [n_tiles * 16] or null
  int K, kc;           // kc = ceil(K / 64)
  int n_tiles, terms;
  int img_off, slots;  // resident images: slots x terms' weight chunks (kc for Wh, + kc for Wl) x 2 KB
};

__device__ inline void tc_load_dense_resident(const TcDense& d, int RB, const TcState& st) {
  if (d.slots == 0) return;
  const int n_items = d.n_tiles * RB, wch = (d.terms == 3 ? 2 : 1) * d.kc;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items && slot < d.slots; item += gridDim.x, ++slot) {
    const int tile = item / RB;
    const int grp = tile < d.nt[0] ? 0 : 1, lt = tile - (grp ? d.nt[0] : 0);
    uint8_t* img = st.wimg + d.img_off + (size_t)slot * wch * 2048;
    for (int i = threadIdx.x; i < d.kc * tc::CHUNK_K16 * TC_N; i += NT) {
      const int n = i & (TC_N - 1), k = i >> 4;
      float w = 0.f;
      if (k < d.K && lt * TC_N + n < d.ncol[grp]) w = __ldg(d.W[grp] + (size_t)k * d.ldw[grp] + lt * TC_N + n);
      const __half hi = tc::to_f16(w);
      const uint32_t o = (k / tc::CHUNK_K16) * 2048 + tc::sw128_offset16(n, k % tc::CHUNK_K16);
      *(__half*)(img + o) = hi;
      if (d.terms == 3) *(__half*)(img + d.kc * 2048 + o) = tc::to_f16(w - __half2float(hi));
    }
  }
}

// Every thread of the CTA calls it.  epi(row b, tile, ul, g[4]) receives the pre-bias sums of columns
// tile * 16 + ul * 4 .. + 3 of utterance row b (one call per thread and item).
template <class Epi>
__device__ inline void dense_phase_tc(const TcDense& d, const CUtensorMap* xmap, int RB, TcState& st, Epi epi) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_items = d.n_tiles * RB, wch = (d.terms == 3 ? 2 : 1) * d.kc, nch = d.terms * d.kc;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++slot) {
    const int tile = item / RB, row0 = (item % RB) * TC_M;
    const uint8_t* img = st.wimg + d.img_off + (size_t)slot * wch * 2048;
    if (tid == 0) {
      for (int c = 0; c < nch; ++c) {
        const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
        if (use > 0) tc::mbar_wait(st.empty + s, (use - 1) & 1);
        const int term = c / d.kc, cc = c - term * d.kc;
        tc::mbar_arrive_expect_tx(st.full + s, TC_STAGE_BYTES);
        tc::tma_load_2d(st.ring + s * TC_STAGE_BYTES, xmap, st.full + s, ((term == 1 ? d.kc : 0) + cc) * tc::CHUNK_K16, row0);
      }
    } else if (tid == 32) {
      const uint32_t idesc = tc::idesc_f16(TC_M, TC_N);
      for (int c = 0; c < nch; ++c) {
        const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
        tc::mbar_wait(st.full + s, use & 1);
        tc::tc_fence_after();
        const int term = c / d.kc, cc = c - term * d.kc;
        const uint32_t a0 = tc::smem_u32(st.ring + s * TC_STAGE_BYTES);
        const uint32_t b0 = tc::smem_u32(img + ((term == 2 ? d.kc : 0) + cc) * 2048);
#pragma unroll
        for (int k16 = 0; k16 < tc::CHUNK_K16 / tc::UMMA_K16; ++k16)
          tc::mma_f16(st.tmem, tc::smem_desc_sw128(a0 + k16 * 32), tc::smem_desc_sw128(b0 + k16 * 32), idesc,
                      (c | k16) != 0);
        tc::mma_commit(st.empty + s);
      }
      tc::mma_commit(st.done);
    }
    __syncwarp();
    tc::mbar_wait(st.done, st.n_items & 1);
    tc::tc_fence_after();
    {
      const int q = warp & 3, ul = warp >> 2;
      float g[4];
      tc::tmem_ld4(st.tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(ul * 4), g);
      epi(row0 + q * 32 + lane, tile, ul, g);
    }
    tc::tc_fence_before();
    st.n_chunks += nch;
    st.n_items += 1;
    __syncthreads();  // TMEM accumulator and ring hand-over to the next item / phase
  }
}

// One ZoneoutLSTM phase on the tensor cores.  Every thread of the CTA calls it.
// Roles: thread 0 issues the TMA loads, thread 32 the MMAs; once the accumulator is complete all 16 warps run the
// cell update -- warp w reads TMEM lanes 32*(w%4).. (one utterance row per thread) and the four gate columns of unit
// w/4 of the tile.  The cell state c and the zoned h of the CTA's own (tile, row block) never leave shared memory.
__device__ inline void lstm_phase_tc(const TcLstmDesc& d, const TcLstmIO& io, int B, int Bp, int RB, float zc, float zh,
                                     TcState& st) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_items = d.n_tiles * RB;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++slot) {
    const int tile = item / RB, rb = item % RB;
    const int row0 = rb * TC_M;
    const uint8_t* img = st.wimg + d.img_off + (size_t)slot * d.kchunks * 2048;
    const bool prof_on = st.prof && blockIdx.x == 0;
    long long t0 = 0;
    if (prof_on) t0 = clock64();
    if (tid == 0) {
      // ---- TMA producer (the first st.prefilled chunks of the first item are already in flight) ----
      const int c0 = slot == 0 ? (int)st.prefilled : 0;
      st.prefilled = 0;
      for (int c = c0; c < d.kchunks; ++c) {
        const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
        if (use > 0) tc::mbar_wait(st.empty + s, (use - 1) & 1);
        tc::mbar_arrive_expect_tx(st.full + s, TC_STAGE_BYTES);
        tc::tma_load_2d(st.ring + s * TC_STAGE_BYTES, io.xmap, st.full + s, d.order[c] * tc::CHUNK_K16, row0);
      }
      if (prof_on) st.prof[0] += clock64() - t0;
    } else if (tid == 32) {
      // ---- MMA issuer ----
      const uint32_t idesc = tc::idesc_f16(TC_M, TC_N);
      for (int c = 0; c < d.kchunks; ++c) {
        const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
        tc::mbar_wait(st.full + s, use & 1);
        tc::tc_fence_after();
        const uint32_t a0 = tc::smem_u32(st.ring + s * TC_STAGE_BYTES), b0 = tc::smem_u32(img + d.order[c] * 2048);
#pragma unroll
        for (int k16 = 0; k16 < tc::CHUNK_K16 / tc::UMMA_K16; ++k16)
          tc::mma_f16(st.tmem, tc::smem_desc_sw128(a0 + k16 * 32), tc::smem_desc_sw128(b0 + k16 * 32), idesc,
                      (c | k16) != 0);
        tc::mma_commit(st.empty + s);
      }
      tc::mma_commit(st.done);
    }
    __syncwarp();
    // ---- cell update: TMEM -> registers (one utterance row, one unit per thread) ----
    tc::mbar_wait(st.done, st.n_items & 1);
    tc::tc_fence_after();
    long long t1 = 0;
    if (prof_on && tid == 0) {
      t1 = clock64();
      st.prof[1] += t1 - t0;
    }
    {
      const int q = warp & 3, ul = warp >> 2;
      float g[4];
      tc::tmem_ld4(st.tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(ul * 4), g);
      const int r = q * 32 + lane, b = row0 + r, u = tile * 4 + ul;
      float* cs = (float*)(st.wimg + d.st_off + (size_t)slot * TC_STATE_BYTES) + ul * TC_M + r;
      float* hs = cs + 4 * TC_M;
      const float4 bb = __ldg((const float4*)(d.bias + (size_t)(2 * tile + (ul >> 1)) * TILE_N + (ul & 1) * 4));
      const float gi = g[0] + bb.x, gj = g[1] + bb.y, gf = g[2] + bb.z, go = g[3] + bb.w;
      const float cp = *cs, hp = *hs;
      const float cn = sigmoid_fast(gf + 1.0f) * cp + sigmoid_fast(gi) * tanh_fast(gj);
      const float hn = sigmoid_fast(go) * tanh_fast(cn);
      if (b < Bp) {
        *cs = cn - zc * (cn - cp);
        // the recurrent input of the next step is the fp16 value; the blend uses the same rounded value
        const __half hzh = tc::to_f16(hn - zh * (hn - hp));
        *hs = __half2float(hzh);
        ((__half*)io.hz.p)[(size_t)u * io.hz.sd + (size_t)b * io.hz.sb] = hzh;
        if (io.out0.p) io.out0.put(u, b, hn);
        if (io.out1.p) io.out1.put(u, b, hn);
        if (io.hl) {
          const __half hh = tc::to_f16(hn);
          io.hl[(size_t)b * io.hl_stride + u] = hh;
          io.hl[(size_t)b * io.hl_stride + io.hl_lo + u] = tc::to_f16(hn - __half2float(hh));
        }
        if (io.rm_out && b < B) io.rm_out[(size_t)b * io.rm_stride + io.rm_off + u] = hn;
      }
    }
    tc::tc_fence_before();
    if (prof_on && tid == 0) st.prof[2] += clock64() - t1;
    st.n_chunks += d.kchunks;
    st.n_items += 1;
    __syncthreads();  // TMEM accumulator and ring hand-over to the next item / phase
  }
}

}  // namespace dec
}  // namespace sat
