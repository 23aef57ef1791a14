// decoder_tc.cuh -- tensor-core form of the ZoneoutLSTM gate GEMMs of the persistent decoder kernel, used once the
// utterance batch makes them a real contraction (B >= 64; BASELINE.json north_star).
//
// Work split: a CTA owns (column tile of 16 gate columns = 4 units x {i,j,f,o}) x (row block of 128 utterances).
// The W^T slice [16][K] of every LSTM stays resident in shared memory for the whole run as a K-major, 128-byte
// swizzled tf32 image; the activations X[128 rows][K] (row-major in HBM/L2) stream through a 6-stage ring of
// 16 KB TMA boxes {32 k, 128 rows}; one elected thread issues tcgen05.mma kind::tf32 (M=128, N=16, K=8) into
// a 16-column TMEM accumulator; four warps read their 32 TMEM lanes back (tcgen05.ld 32x32b.x16 -> one utterance
// row and 4 complete units per thread) and apply the zoneout cell update in registers.
#pragma once
#include "decoder_dev.cuh"
#include "tc.cuh"

namespace sat {
namespace dec {

constexpr int TC_STAGES = 6;
constexpr int TC_STAGE_BYTES = 128 * tc::CHUNK_K * 4;  // 16 KB: 128 rows x 32 tf32
constexpr int TC_N = 16;                               // gate columns per tile
constexpr int TC_M = 128;                              // utterance rows per tile
constexpr int TC_EPI_WARP0 = 4;                        // warps 4..7 read TMEM lanes 0..127

// element (feature n, row b) of an activation buffer
struct Act2D {
  float* p;
  int sd, sb;  // strides of the feature / row index (feature-major: sd = ld, sb = 1; row-major: sd = 1, sb = K)
  int tf32;    // 1: the buffer feeds a tensor-core GEMM -> values are stored rounded (RN) to tf32
  __device__ __forceinline__ float* at(int n, int b) const { return p + (size_t)n * sd + (size_t)b * sb; }
  __device__ __forceinline__ void put(int n, int b, float v) const { *at(n, b) = tf32 ? tc::round_tf32(v) : v; }
};

struct TcLstmDesc {
  const float* Wp;    // packed tiles [units/2][K][8] (same packing as the SIMT path; global)
  const float* bias;  // [units/2 * 8]
  int K, units;
  int n_tiles;        // units / 4 column tiles of 16
  int img_off;        // byte offset of this layer's resident images inside the TC weight region
  int slots;          // resident (tile,row block) items per CTA
  int kchunks;        // ceil(K / 32)
};

// per-CTA pipeline state that persists across phases and decoder steps
struct TcState {
  uint8_t* ring;       // TC_STAGES x 16 KB, 1024-aligned (aliases the phase scratch)
  uint8_t* wimg;       // resident weight images, 1024-aligned
  uint64_t* full;      // [TC_STAGES]
  uint64_t* empty;     // [TC_STAGES]
  uint64_t* done;
  uint32_t tmem;
  uint32_t n_chunks;   // chunks pushed through the ring so far (identical in every thread)
  uint32_t n_items;    // accumulator hand-overs so far
  long long* prof;     // optional cycle counters (CTA 0): [0] producer loop, [1] wait for the accumulator, [2] cell update
};

// build this CTA's resident K-major SW128 image of one LSTM: chunk c of slot s at wimg + img_off + (s*kchunks + c)*2 KB
__device__ inline void tc_load_resident(const TcLstmDesc& d, int RB, const TcState& st) {
  if (d.slots == 0) return;
  const int n_items = d.n_tiles * RB;
  int slot = 0;
  for (int item = blockIdx.x; item < n_items && slot < d.slots; item += gridDim.x, ++slot) {
    const int tile = item / RB;
    uint8_t* img = st.wimg + d.img_off + (size_t)slot * d.kchunks * 2048;
    // column n of the 16-wide tile = column (n & 7) of packed 8-wide tile 2*tile + (n >> 3)
    for (int i = threadIdx.x; i < d.kchunks * tc::CHUNK_K * TC_N; i += NT) {
      const int n = i & (TC_N - 1), k = i >> 4;
      float w = 0.f;
      if (k < d.K) w = __ldg(d.Wp + ((size_t)(2 * tile + (n >> 3)) * d.K + k) * TILE_N + (n & 7));
      *(float*)(img + (k / tc::CHUNK_K) * 2048 + tc::sw128_offset(n, k % tc::CHUNK_K)) = tc::round_tf32(w);
    }
  }
}

struct TcLstmIO {
  const CUtensorMap* xmap;  // activations of this step: row-major [rows][Kpad]
  Act2D hprev;              // previous zoned h (read), feature index = unit
  Act2D out0, out1;         // un-zoned cell output (written; out1.p may be null)
  Act2D hz;                 // zoned h for the next step (written)
  float* c;                 // cell state [units][ld]
  float* rm_out;            // optional row-major copy of the output: rm_out[b * rm_stride + rm_off + u]
  int rm_stride, rm_off;
};

// One ZoneoutLSTM phase on the tensor cores.  Every thread of the CTA calls it.
__device__ inline void lstm_phase_tc(const TcLstmDesc& d, const TcLstmIO& io, int B, int Bp, int ld, int RB, float zc,
                                     float zh, TcState& st) {
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
      // ---- TMA producer ----
      for (int c = 0; c < d.kchunks; ++c) {
        const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
        if (use > 0) tc::mbar_wait(st.empty + s, (use - 1) & 1);
        tc::mbar_arrive_expect_tx(st.full + s, TC_STAGE_BYTES);
        tc::tma_load_2d(st.ring + s * TC_STAGE_BYTES, io.xmap, st.full + s, c * tc::CHUNK_K, row0);
      }
      if (prof_on) st.prof[0] += clock64() - t0;
    } else if (tid == 32) {
      // ---- MMA issuer ----
      const uint32_t idesc = tc::idesc_tf32(TC_M, TC_N);
      for (int c = 0; c < d.kchunks; ++c) {
        const uint32_t n = st.n_chunks + c, s = n % TC_STAGES, use = n / TC_STAGES;
        tc::mbar_wait(st.full + s, use & 1);
        tc::tc_fence_after();
        const uint32_t a0 = tc::smem_u32(st.ring + s * TC_STAGE_BYTES), b0 = tc::smem_u32(img + c * 2048);
#pragma unroll
        for (int k8 = 0; k8 < tc::CHUNK_K / tc::UMMA_K; ++k8)
          tc::mma_tf32(st.tmem, tc::smem_desc_sw128(a0 + k8 * 32), tc::smem_desc_sw128(b0 + k8 * 32), idesc,
                       (c | k8) != 0);
        tc::mma_commit(st.empty + s);
      }
      tc::mma_commit(st.done);
    } else if (warp >= TC_EPI_WARP0 && warp < TC_EPI_WARP0 + 4) {
      // ---- epilogue: TMEM -> registers -> zoneout cell update (one utterance row, 4 units per thread) ----
      tc::mbar_wait(st.done, st.n_items & 1);
      tc::tc_fence_after();
      const int q = warp - TC_EPI_WARP0;
      float g[TC_N];
      tc::tmem_ld16(st.tmem + ((uint32_t)(q * 32) << 16), g);
      long long t1 = 0;
      if (prof_on && tid == TC_EPI_WARP0 * 32) {
        t1 = clock64();
        st.prof[1] += t1 - t0;
      }
      const int b = row0 + q * 32 + lane;
      if (b < Bp) {
        // 4 consecutive units of one row: the row-major buffers take one 16-byte access each
        const int u0 = tile * 4;
        const float4 hp4 = __ldcg((const float4*)io.hprev.at(u0, b));  // hprev / hz are always row-major here
        const float hpv[4] = {hp4.x, hp4.y, hp4.z, hp4.w};
        float hzv[4], hnv[4];
#pragma unroll
        for (int ul = 0; ul < 4; ++ul) {
          const float4 bb = __ldg((const float4*)(d.bias + (size_t)(2 * tile + (ul >> 1)) * TILE_N + (ul & 1) * 4));
          const float gi = g[ul * 4 + 0] + bb.x, gj = g[ul * 4 + 1] + bb.y, gf = g[ul * 4 + 2] + bb.z, go = g[ul * 4 + 3] + bb.w;
          const size_t idx = (size_t)(u0 + ul) * ld + b;
          const float cp = io.c[idx];
          const float cn = sigmoidf_(gf + 1.0f) * cp + sigmoidf_(gi) * tanhf(gj);
          const float hn = sigmoidf_(go) * tanhf(cn);
          io.c[idx] = cn - zc * (cn - cp);
          hnv[ul] = hn;
          hzv[ul] = tc::round_tf32(hn - zh * (hn - hpv[ul]));
        }
        *(float4*)io.hz.at(u0, b) = make_float4(hzv[0], hzv[1], hzv[2], hzv[3]);
        if (io.out0.sd == 1) {
          *(float4*)io.out0.at(u0, b) = make_float4(tc::round_tf32(hnv[0]), tc::round_tf32(hnv[1]),
                                                    tc::round_tf32(hnv[2]), tc::round_tf32(hnv[3]));
        } else {
#pragma unroll
          for (int ul = 0; ul < 4; ++ul) io.out0.put(u0 + ul, b, hnv[ul]);
        }
        if (io.out1.p) {
#pragma unroll
          for (int ul = 0; ul < 4; ++ul) io.out1.put(u0 + ul, b, hnv[ul]);
        }
        if (io.rm_out && b < B) {
#pragma unroll
          for (int ul = 0; ul < 4; ++ul) io.rm_out[(size_t)b * io.rm_stride + io.rm_off + u0 + ul] = hnv[ul];
        }
      }
      tc::tc_fence_before();
      if (prof_on && tid == TC_EPI_WARP0 * 32) st.prof[2] += clock64() - t1;
    }
    st.n_chunks += d.kchunks;
    st.n_items += 1;
    __syncthreads();  // TMEM accumulator and ring hand-over to the next item / phase
  }
}

}  // namespace dec
}  // namespace sat
