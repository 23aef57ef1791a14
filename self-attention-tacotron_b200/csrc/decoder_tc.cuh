// decoder_tc.cuh -- tensor-core form of the ZoneoutLSTM gate GEMMs of the persistent decoder kernel, used once the
// utterance batch makes them a real contraction (B >= 64; BASELINE.json north_star).
//
// Work split: a CTA owns (column tile of 16 gate columns = 4 units x {i,j,f,o}) x (row block of 128 utterances).
// The W^T slice [16][K] of every LSTM stays resident in shared memory for the whole run as a K-major, 128-byte
// swizzled fp16 image.  The activations X[128 rows][K] live in HBM/L2 as fp16 in the very byte order of their
// shared-memory operand tiles -- [row block][64-k chunk][128 rows][64 k], 128-byte swizzle applied by the writers
// (tc_act_index) -- so that a row block's K range is ONE contiguous run and streams through a 2-stage ring as a few
// large 1-D bulk copies (cp.async.bulk, up to 64 KB each): the SM's TMA unit retires one copy or tensor box per
// ~0.28 us whatever its size (scripts/dev/tma_bench.cu), so twelve 16 KB boxes cost 3.5 us and three 64 KB copies 1.5.
// One elected thread issues tcgen05.mma kind::f16 (M=128, N=16, K=16) into a 16-column TMEM accumulator; all sixteen
// warps read it back (tcgen05.ld 32x32b.x4 -> one utterance row and one complete unit per thread) and apply the
// zoneout cell update in registers.
// Operands are fp16 with fp32 accumulation: fp16 has the 11-bit significand of tf32 at half the bytes, and the
// phase is bound by the L2 -> shared-memory stream of X.
#pragma once
#include "decoder_dev.cuh"
#include "tc.cuh"

namespace sat {
namespace dec {

constexpr int TC_N = 16;                                 // gate columns per tile
constexpr int TC_M = 128;                                // utterance rows per tile
constexpr int TC_TILE_BYTES = TC_M * tc::CHUNK_K16 * 2;  // one operand tile: 128 rows x 64 fp16 = 16 KB
constexpr int TC_MAX_STAGES = 4;
constexpr int TC_STATE_BYTES = 2 * 4 * TC_M * 4;         // per resident item: c and zoned h of 4 units x 128 rows (fp32)

// element (feature n, row b) of an activation buffer
struct Act2D {
  void* p;
  int sd, sb;  // fp32: strides of the feature / row index (feature-major: sd = ld, sb = 1)
               // fp16 (tensor-core operand buffer, tc_act_index order): sd = feature offset of n = 0, sb = chunks per row
  int f16;     // 1: fp16 buffer feeding a tensor-core GEMM (values rounded to nearest), 0: fp32
  __device__ __forceinline__ void put(int n, int b, float v) const {
    if (f16) ((__half*)p)[tc_act_index(sd + n, b, sb)] = tc::to_f16(v);
    else ((float*)p)[(size_t)n * sd + (size_t)b * sb] = v;
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
  int ncopy;          // bulk copies per item: ceil(kchunks / G)
  int n_pre;          // leading copies of `order` whose columns were produced at least two phases ago: they are requested
                      // while the CTA waits at the grid barrier in front of the phase (at most one per ring stage)
  unsigned char order[16];  // copy visiting order (a permutation of 0 .. ncopy-1), "old" columns first
};

// per-CTA pipeline state that persists across phases and decoder steps
struct TcState {
  int G, stages;       // 64-k chunks per bulk copy, copies in flight (ring = stages x G x 16 KB)
  uint8_t* ring;       // 1024-aligned (aliases the phase scratch)
  uint8_t* wimg;       // resident weight images + cell state, 1024-aligned
  uint64_t* full;      // [TC_MAX_STAGES]
  uint64_t* empty;     // [TC_MAX_STAGES]
  uint64_t* done;
  uint32_t tmem;
  uint32_t n_chunks;   // copies pushed through the ring so far (identical in every thread)
  uint32_t n_items;    // accumulator hand-overs so far
  uint32_t prefilled;  // copies of the coming LSTM phase already requested by tc_prefill (identical in every thread)
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

// thread 0, between arriving at and waiting on the grid barrier in front of an LSTM phase: request the leading copies of
// the CTA's first item (columns that do not depend on the phase that is just finishing; the phase scratch, which the
// ring aliases, is free once the CTA has arrived).  Every thread then sets st.prefilled = d.n_pre.
__device__ inline void tc_prefill(const TcLstmDesc& d, const __half* x, int RB, const TcState& st) {
  const int item = blockIdx.x;
  if (item >= d.n_tiles * RB) return;
  const uint32_t sbytes = st.G * TC_TILE_BYTES;
  const uint8_t* src = (const uint8_t*)x + (size_t)(item % RB) * d.kchunks * TC_TILE_BYTES;
  tc::fence_proxy_async();
  for (int i = 0; i < d.n_pre; ++i) {
    const uint32_t n = st.n_chunks + i, s = n % st.stages, use = n / st.stages;
    const int c = d.order[i];
    const uint32_t bytes = (uint32_t)min(st.G, d.kchunks - c * st.G) * TC_TILE_BYTES;
    if (use > 0) tc::mbar_wait(st.empty + s, (use - 1) & 1);
    tc::mbar_arrive_expect_tx(st.full + s, bytes);
    tc::bulk_load_1d(st.ring + s * sbytes, src + (size_t)c * sbytes, bytes, st.full + s);
  }
}

struct TcLstmIO {
  const __half* x;          // activations of this step: fp16 operand-tile order [row block][kchunks][128][64]
  Act2D out0, out1;         // un-zoned cell output (written; either .p may be null)
  Act2D hz;                 // zoned h for the next step (written)
  float* rm_out;            // optional row-major fp32 copy of the output: rm_out[b * rm_stride + rm_off + u]
  int rm_stride, rm_off;
};

// One ZoneoutLSTM phase on the tensor cores.  Every thread of the CTA calls it.
// Roles: thread 0 issues the bulk copies, thread 32 the MMAs; once the accumulator is complete all 16 warps run the
// cell update -- warp w reads TMEM lanes 32*(w%4).. (one utterance row per thread) and the four gate columns of unit
// w/4 of the tile.  The cell state c and the zoned h of the CTA's own (tile, row block) never leave shared memory.
__device__ inline void lstm_phase_tc(const TcLstmDesc& d, const TcLstmIO& io, int B, int Bp, int RB, float zc, float zh,
                                     TcState& st) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int warp_u = __shfl_sync(0xffffffffu, warp, 0);  // the warp index as a value the compiler knows to be warp-uniform
  const int n_items = d.n_tiles * RB;
  const uint32_t stages = st.stages, G = st.G, sbytes = G * TC_TILE_BYTES;
  const int ncopy = d.ncopy;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++slot) {
    const int tile = item / RB, rb = item % RB;
    const int row0 = rb * TC_M;
    const uint8_t* img = st.wimg + d.img_off + (size_t)slot * d.kchunks * 2048;
    const bool prof_on = st.prof && blockIdx.x == 0;
    long long t0 = 0;
    if (prof_on) t0 = clock64();
    if ((tid & 63) == 0 && (uint32_t)(tid >> 6) < stages) {
      // ---- producers, one thread per ring stage (issuing a bulk copy holds its thread for ~0.3 us, so the copies of a
      // phase are issued side by side): the row block's K range is one contiguous run of operand tiles ----
      const uint8_t* src = (const uint8_t*)io.x + (size_t)rb * d.kchunks * TC_TILE_BYTES;
      const uint32_t s = (uint32_t)(tid >> 6);
      // the i-th copy of the item is the (st.n_chunks + i)-th through the ring: this thread's are those that land in
      // stage s (the first st.prefilled copies of the first item are already in flight)
      uint32_t n = st.n_chunks + (s + stages - st.n_chunks % stages) % stages;
      if (slot == 0)
        while (n < st.n_chunks + st.prefilled) n += stages;
      for (; n < st.n_chunks + (uint32_t)ncopy; n += stages) {
        const int c = d.order[n - st.n_chunks];
        const uint32_t use = n / stages;
        const uint32_t bytes = (uint32_t)min((int)G, d.kchunks - c * (int)G) * TC_TILE_BYTES;
        if (use > 0) tc::mbar_wait(st.empty + s, (use - 1) & 1);
        tc::mbar_arrive_expect_tx(st.full + s, bytes);
        tc::bulk_load_1d(st.ring + s * sbytes, src + (size_t)c * sbytes, bytes, st.full + s);
      }
      if (prof_on && tid == 0) st.prof[0] += clock64() - t0;
    } else if (warp_u == 1) {
      // ---- MMA issuer: all of warp 1 runs the loop on warp-uniform values, an elected lane issues ----
      const uint32_t idesc = tc::idesc_f16(TC_M, TC_N);
      uint32_t s = st.n_chunks % stages, use = st.n_chunks / stages;
      for (int c = 0; c < ncopy; ++c) {
        tc::mbar_wait(st.full + s, use & 1);
        tc::tc_fence_after();
        const int ch0 = d.order[c] * (int)G, nch = min((int)G, d.kchunks - ch0);
        uint32_t alo = tc::desc_lo(tc::smem_u32(st.ring + s * sbytes)), blo = tc::desc_lo(tc::smem_u32(img + ch0 * 2048));
        for (int j = 0; j < nch; ++j, alo += TC_TILE_BYTES >> 4, blo += 2048 >> 4) {
#pragma unroll
          for (int k16 = 0; k16 < tc::CHUNK_K16 / tc::UMMA_K16; ++k16)
            tc::mma_f16_warp(st.tmem, alo + 2 * k16, blo + 2 * k16, idesc, (c | j | k16) != 0);
        }
        tc::mma_commit_warp(st.empty + s);
        if (++s == stages) s = 0, ++use;
      }
      tc::mma_commit_warp(st.done);
      tc::mbar_wait(st.done, st.n_items & 1);
    }
    // everyone else parks at the hardware barrier: 500 threads polling the mbarrier in shared memory would compete with
    // the bulk copies for the shared-memory port
    __syncthreads();
    // ---- cell update: TMEM -> registers (one utterance row, one unit per thread) ----
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
        ((__half*)io.hz.p)[tc_act_index(io.hz.sd + u, b, io.hz.sb)] = hzh;
        if (io.out0.p) io.out0.put(u, b, hn);
        if (io.out1.p) io.out1.put(u, b, hn);
        if (io.rm_out && b < B) io.rm_out[(size_t)b * io.rm_stride + io.rm_off + u] = hn;
      }
    }
    tc::tc_fence_before();
    if (prof_on && tid == 0) st.prof[2] += clock64() - t1;
    st.n_chunks += ncopy;
    st.n_items += 1;
    st.prefilled = 0;
    __syncthreads();  // TMEM accumulator and ring hand-over to the next item / phase
  }
}

}  // namespace dec
}  // namespace sat
