// mha.cu -- multi-head scaled-dot self-attention, fp32 SIMT form.
//
// Restates modules/self_attention.py (citations into /root/reference):
//   MultiHeadAttention.call :108-128   K/V/Q Dense (bias) -> heads -> attention -> merge -> output Dense
//   ScaledDotProductAttentionMechanism.__call__ :45-65   softmax(mask(Q K^T / sqrt(d_head))) V
//   padding mask :67-78 (query rows AND key columns past the length -> -inf), subsequent mask :80-86
//   SelfAttentionTransformer.call modules/module.py:363-371   x + tanh(Dense(mha(x)))
//
// K/V/Q projections run as ONE fused [Din, 3D] GEMM; the attention kernel gives one warp per query
// row (lanes split the keys for the scores, then (key group, 4 head dims) for the context).
#include "common.cuh"
#include "tc.cuh"
#include <cstdlib>

namespace sat {
namespace {

constexpr int MHA_WARPS = 8;

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// qkv [B][T][3D] (key | value | query); ctx [B][T][D]; probs [B][H][T][T] or null.
// grid = (ceil(T / MHA_WARPS), B*H); dynamic smem = MHA_WARPS * (Tpad + dh) floats.
__global__ void __launch_bounds__(MHA_WARPS * 32) mha_attn_kernel(const float* __restrict__ qkv, int T, int D, int H,
                                                                  int causal, const long long* __restrict__ lengths,
                                                                  float* __restrict__ ctx, float* __restrict__ probs,
                                                                  int head_major) {
  extern __shared__ __align__(16) float sm[];
  const int dh = D / H, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y / H, h = blockIdx.y % H;
  const int i = blockIdx.x * MHA_WARPS + warp;
  if (i >= T) return;
  const int Tpad = (T + 3) & ~3;
  float* sc = sm + (size_t)warp * (Tpad + dh);
  float* q = sc + Tpad;
  const size_t ld = 3 * (size_t)D;
  const float* base = qkv + (size_t)b * T * ld + (size_t)h * dh;
  for (int d = lane; d < dh; d += 32) q[d] = base[(size_t)i * ld + 2 * D + d];
  __syncwarp();
  const int len = lengths ? (int)lengths[b] : T;
  const int nk = causal ? i + 1 : T;
  const bool row_ok = i < len;
  const float scale = 1.f / sqrtf((float)dh);
  const bool vec = (dh & 3) == 0 && (D & 3) == 0;
  float mx = -INFINITY;
  for (int j = lane; j < nk; j += 32) {
    const float* kr = base + (size_t)j * ld;
    float s = 0.f;
    if (vec) {
      for (int d = 0; d < dh; d += 4) {
        const float4 kv = __ldg((const float4*)(kr + d));
        s = fmaf(q[d], kv.x, s); s = fmaf(q[d + 1], kv.y, s); s = fmaf(q[d + 2], kv.z, s); s = fmaf(q[d + 3], kv.w, s);
      }
    } else {
      for (int d = 0; d < dh; ++d) s = fmaf(q[d], __ldg(kr + d), s);
    }
    s *= scale;
    if (!row_ok || j >= len) s = -INFINITY;
    sc[j] = s;
    mx = fmaxf(mx, s);
  }
  mx = wmax(mx);
  if (mx == -INFINITY) mx = 0.f;  // fully masked row: exp(-inf) = 0 everywhere -> 0/0 = NaN like the reference
  float sum = 0.f;
  for (int j = lane; j < nk; j += 32) {
    const float e = expf(sc[j] - mx);
    sc[j] = e;
    sum += e;
  }
  sum = wsum(sum);
  const float inv = 1.f / sum;
  __syncwarp();
  if (probs) {
    // [B][H][T][T], or [H][B][T][T] when head_major (one tensor per head, models/models.py:104)
    float* pr = probs + ((head_major ? (size_t)h * gridDim.y / H + b : (size_t)b * H + h) * T + i) * T;
    for (int j = lane; j < T; j += 32) pr[j] = j < nk ? sc[j] * inv : 0.f;
  }
  float* out = ctx + ((size_t)b * T + i) * D + (size_t)h * dh;
  const float* vbase = base + D;
  const int q4 = dh >> 2;
  if (vec && q4 <= 32 && (q4 & (q4 - 1)) == 0) {
    const int G = 32 / q4, d0 = (lane % q4) * 4, grp = lane / q4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = grp; j < nk; j += G) {
      const float p = sc[j];
      const float4 v = __ldg((const float4*)(vbase + (size_t)j * ld + d0));
      acc.x = fmaf(p, v.x, acc.x); acc.y = fmaf(p, v.y, acc.y); acc.z = fmaf(p, v.z, acc.z); acc.w = fmaf(p, v.w, acc.w);
    }
    for (int o = q4; o < 32; o <<= 1) {
      acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o);
      acc.z += __shfl_xor_sync(0xffffffffu, acc.z, o); acc.w += __shfl_xor_sync(0xffffffffu, acc.w, o);
    }
    if (grp == 0) *(float4*)(out + d0) = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
  } else {
    for (int d = lane; d < dh; d += 32) {
      float acc = 0.f;
      for (int j = 0; j < nk; ++j) acc = fmaf(sc[j], __ldg(vbase + (size_t)j * ld + d), acc);
      out[d] = acc * inv;
    }
  }
}


// ---- tensor-core form (tcgen05 / TMEM) for short sequences and small heads: T <= 256, dh in {16, 32, 48, 64} ------
// One CTA = one (batch row, head, 128-query tile).  Q (pre-scaled), K and V^T are converted to fp16 hi + lo pairs
// into K-major 128-byte-swizzled shared memory; S = Q K^T (3 MMAs per K-step: lo*hi + hi*lo + hi*hi, i.e. fp32-grade
// products) accumulates in TMEM columns [0, Tp); every thread owns one query row (= one TMEM lane), applies the
// masks and the softmax straight from TMEM, writes the probabilities (optional) and the fp16 P operand; O = P V
// (V as hi + lo) accumulates in TMEM columns [256, 256 + dh) and is written back as fp32.
constexpr int MT_THREADS = 128;

// The fp32 Q tile and the K / V blocks of the head arrive by TMA (three un-swizzled boxes of the [B*T, 3D] projection
// output into a staging area, one mbarrier); the threads then split them into the swizzled fp16 hi / lo operands.
// use_tma = 0 (staging does not fit next to the operands: wide heads) reads them with plain loads instead.
__global__ void __launch_bounds__(MT_THREADS, 1)
mha_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_kv, int use_tma,
              const float* __restrict__ qkv, int T, int Tp, int D, int H, int causal, const long long* __restrict__ lengths,
              float* __restrict__ ctx, float* __restrict__ probs, int head_major) {
  extern __shared__ __align__(1024) uint8_t mt_smem[];
  uint8_t* base = (uint8_t*)(((uintptr_t)mt_smem + 1023) & ~(uintptr_t)1023);
  const int dh = D / H, nchunk = (Tp + 63) / 64;
  uint8_t* Qh = base;
  uint8_t* Ql = Qh + 16384;
  uint8_t* Kh = Ql + 16384;
  uint8_t* Kl = Kh + (size_t)256 * 128;
  uint8_t* Vh = Kl + (size_t)256 * 128;              // nchunk x [dh rows][128 B]
  uint8_t* Vl = Vh + (size_t)nchunk * dh * 128;
  uint8_t* P = (uint8_t*)(((uintptr_t)(Vl + (size_t)nchunk * dh * 128) + 1023) & ~(uintptr_t)1023);  // nchunk x 16 KB
  float* Qst = (float*)(P + (size_t)nchunk * 16384);  // TMA staging: [128][dh], [Tp][dh], [Tp][dh] fp32
  float* Kst = Qst + 128 * dh;
  float* Vst = Kst + (size_t)Tp * dh;
  __shared__ __align__(8) uint64_t bars[3];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int b = blockIdx.y / H, h = blockIdx.y % H, q0 = blockIdx.x * 128;
  const size_t ld = 3 * (size_t)D;
  const float* src = qkv + (size_t)b * T * ld + (size_t)h * dh;
  const float scale = 1.f / sqrtf((float)dh);

  if (tid == 0) {
    tc::mbar_init(bars, 1);
    tc::mbar_init(bars + 1, 1);
    tc::mbar_init(bars + 2, 1);
    tc::fence_mbar_init();
    if (use_tma) {
      // rows past this utterance (the next one's, or out of bounds = zeros) are masked below like in the plain form
      tc::mbar_arrive_expect_tx(bars + 2, (uint32_t)((128 + 2 * Tp) * dh * 4));
      tc::tma_load_2d(Qst, &map_q, bars + 2, 2 * D + h * dh, b * T + q0);
      tc::tma_load_2d(Kst, &map_kv, bars + 2, h * dh, b * T);
      tc::tma_load_2d(Vst, &map_kv, bars + 2, D + h * dh, b * T);
    }
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  auto split_store = [](uint8_t* hi, uint8_t* lo, uint32_t off, float x) {
    const __half xh = tc::to_f16(x);
    *(__half*)(hi + off) = xh;
    *(__half*)(lo + off) = tc::to_f16(x - __half2float(xh));
  };
  if (use_tma) {
    __syncthreads();  // the barrier initialisation is visible to the waiting threads
    tc::mbar_wait(bars + 2, 0);
  }
  for (int i = tid; i < 128 * dh; i += MT_THREADS) {
    const int r = i / dh, d = i - r * dh;
    const float x = (q0 + r < T) ? (use_tma ? Qst[i] : __ldg(src + (size_t)(q0 + r) * ld + 2 * D + d)) * scale : 0.f;
    split_store(Qh, Ql, tc::sw128_offset16(r, d), x);
  }
  for (int i = tid; i < Tp * dh; i += MT_THREADS) {
    const int j = i / dh, d = i - j * dh;
    const bool ok = j < T;
    split_store(Kh, Kl, tc::sw128_offset16(j, d), ok ? (use_tma ? Kst[i] : __ldg(src + (size_t)j * ld + d)) : 0.f);
    // V^T: row d, key j -> chunk j / 64
    split_store(Vh, Vl, (uint32_t)((j >> 6) * dh * 128) + tc::sw128_offset16(d, j & 63),
                ok ? (use_tma ? Vst[i] : __ldg(src + (size_t)j * ld + D + d)) : 0.f);
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (tid == 0) {
    const uint32_t idesc = tc::idesc_f16(128, Tp);
    bool first = true;
    for (int ks = 0; ks < dh / 16; ++ks) {
      const uint32_t o = ks * 32;
      tc::mma_f16(tmem, tc::smem_desc_sw128(tc::smem_u32(Ql) + o), tc::smem_desc_sw128(tc::smem_u32(Kh) + o), idesc, !first);
      first = false;
      tc::mma_f16(tmem, tc::smem_desc_sw128(tc::smem_u32(Qh) + o), tc::smem_desc_sw128(tc::smem_u32(Kl) + o), idesc, 1);
      tc::mma_f16(tmem, tc::smem_desc_sw128(tc::smem_u32(Qh) + o), tc::smem_desc_sw128(tc::smem_u32(Kh) + o), idesc, 1);
    }
    tc::mma_commit(bars);
  }
  __syncwarp();
  tc::mbar_wait(bars, 0);
  tc::tc_fence_after();

  // ---- masks + softmax on this thread's query row (TMEM lane tid) ----
  const int qi = q0 + tid;
  const int len = lengths ? (int)lengths[b] : T;
  const int nk = causal ? min(qi + 1, T) : T;
  const int nvalid = (qi < len || !lengths) ? min(nk, len) : 0;  // keys j < nvalid take part
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
  float mx = -INFINITY;
  for (int c0 = 0; c0 < Tp; c0 += 16) {
    float v[16];
    tc::tmem_ld16(lane_addr + c0, v);
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (c0 + k < nvalid) mx = fmaxf(mx, v[k]);
  }
  if (mx == -INFINITY) mx = 0.f;  // fully masked row: 0/0 = NaN like the reference
  float sum = 0.f;
  for (int c0 = 0; c0 < Tp; c0 += 16) {
    float v[16];
    tc::tmem_ld16(lane_addr + c0, v);
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (c0 + k < nvalid) sum += expf(v[k] - mx);
  }
  const float inv = 1.f / sum;
  float* pr = nullptr;
  if (probs && qi < T)
    pr = probs + ((head_major ? (size_t)h * (gridDim.y / H) + b : (size_t)b * H + h) * T + qi) * T;
  for (int c0 = 0; c0 < Tp; c0 += 16) {
    float v[16];
    tc::tmem_ld16(lane_addr + c0, v);
    __half2 ph[8];
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
      const float p0 = c0 + k < nvalid ? expf(v[k] - mx) * inv : (c0 + k < nk ? 0.f * inv : 0.f);
      const float p1 = c0 + k + 1 < nvalid ? expf(v[k + 1] - mx) * inv : (c0 + k + 1 < nk ? 0.f * inv : 0.f);
      v[k] = p0;
      v[k + 1] = p1;
      // fully masked row: NaN probabilities (reference: 0/0), zero P operand, NaN context row written at the end
      ph[k >> 1] = __halves2half2(tc::to_f16(c0 + k < nvalid ? p0 : 0.f), tc::to_f16(c0 + k + 1 < nvalid ? p1 : 0.f));
    }
    uint8_t* pc = P + (size_t)(c0 >> 6) * 16384;
    *(uint4*)(pc + tc::sw128_offset16(tid, c0 & 63)) = *(const uint4*)&ph[0];
    *(uint4*)(pc + tc::sw128_offset16(tid, (c0 & 63) + 8)) = *(const uint4*)&ph[4];
    if (pr)
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (c0 + k < T) pr[c0 + k] = v[k];
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  if (tid == 0) {
    const uint32_t idesc = tc::idesc_f16(128, dh);
    bool first = true;
    for (int c = 0; c < nchunk; ++c)
      for (int ks = 0; ks < 4 && c * 64 + ks * 16 < Tp; ++ks) {
        const uint32_t pa = tc::smem_u32(P + (size_t)c * 16384) + ks * 32, vo = (uint32_t)(c * dh * 128 + ks * 32);
        tc::mma_f16(tmem + 256, tc::smem_desc_sw128(pa), tc::smem_desc_sw128(tc::smem_u32(Vl) + vo), idesc, !first);
        first = false;
        tc::mma_f16(tmem + 256, tc::smem_desc_sw128(pa), tc::smem_desc_sw128(tc::smem_u32(Vh) + vo), idesc, 1);
      }
    tc::mma_commit(bars + 1);
  }
  __syncwarp();
  tc::mbar_wait(bars + 1, 0);
  tc::tc_fence_after();
  for (int d0 = 0; d0 < dh; d0 += 16) {
    float v[16];
    tc::tmem_ld16(lane_addr + 256 + d0, v);
    if (nvalid == 0)
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = __int_as_float(0x7fc00000);
    if (qi < T) {
      float4* out = (float4*)(ctx + ((size_t)b * T + qi) * D + (size_t)h * dh + d0);
#pragma unroll
      for (int k = 0; k < 4; ++k) out[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ---- tensor-core form for long sequences and wide heads (decoder self-attention of the teacher-forced form:
// S up to max_iters, dh = 128): any T, dh a multiple of 16 up to 128.  One CTA = one (batch row, head, 128-query tile),
// one thread = one query row = one TMEM lane.  The keys are visited in chunks of 64, twice (the probabilities are an
// output of the layer, so they must be normalised when they are written):
//   pass 1  S = Q K_c^T (hi/lo fp16 split, 3 MMAs per K-step) -> running row maximum and sum (online softmax);
//   pass 2  S again, p = exp(s - max) / sum: written to `probs` (optional), stored as the fp16 A operand, and
//           O += P V_c (V as hi + lo) accumulates in TMEM without any rescaling.
// With the subsequent mask only the chunks up to the tile's last query are visited.
constexpr int MF_THREADS = 128;
constexpr int MF_KEYS = 64;

// use_tma: the fp32 Q tile and every 64-key chunk of K (and V in pass 2) arrive by TMA (un-swizzled boxes of the
// [B*T, 3D] projection output) in a staging area; the box of the NEXT chunk is requested as soon as the threads have
// split the current one into the fp16 hi / lo operands, so the load runs under the chunk's MMAs and softmax.
__global__ void __launch_bounds__(MF_THREADS, 1)
mha_flash_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_kv, int use_tma,
                    const float* __restrict__ qkv, int T, int D, int H, int causal, const long long* __restrict__ lengths,
                    float* __restrict__ ctx, float* __restrict__ probs, int head_major) {
  extern __shared__ __align__(1024) uint8_t mf_smem[];
  uint8_t* base = (uint8_t*)(((uintptr_t)mf_smem + 1023) & ~(uintptr_t)1023);
  const int dh = D / H, nkc = (dh + 63) / 64;  // 64-wide chunks of the head dimension
  uint8_t* Qh = base;                          // nkc x [128 rows][128 B]
  uint8_t* Ql = Qh + (size_t)nkc * 16384;
  uint8_t* Kh = Ql + (size_t)nkc * 16384;      // nkc x [64 keys][128 B]
  uint8_t* Kl = Kh + (size_t)nkc * 8192;
  uint8_t* Vh = Kl + (size_t)nkc * 8192;       // V^T: [dh rows][64 keys] (16 KB reserved)
  uint8_t* Vl = Vh + 16384;
  uint8_t* P = Vl + 16384;                     // [128 rows][64 keys]
  float* stg = (float*)(P + 16384);            // TMA staging: Q [128][dh], then K [64][dh] | V [64][dh] per chunk (fp32)
  __shared__ __align__(8) uint64_t bars[3];
  uint32_t phs = 0;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int b = blockIdx.y / H, h = blockIdx.y % H, q0 = blockIdx.x * 128;
  const size_t ld = 3 * (size_t)D;
  const float* src = qkv + (size_t)b * T * ld + (size_t)h * dh;
  const float scale = 1.f / sqrtf((float)dh);

  if (tid == 0) {
    tc::mbar_init(bars, 1);
    tc::mbar_init(bars + 1, 1);
    tc::mbar_init(bars + 2, 1);
    tc::fence_mbar_init();
    if (use_tma) {
      tc::mbar_arrive_expect_tx(bars + 2, (uint32_t)(128 * dh * 4));
      tc::tma_load_2d(stg, &map_q, bars + 2, 2 * D + h * dh, b * T + q0);
    }
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
  auto split_store = [](uint8_t* hi, uint8_t* lo, uint32_t off, float x) {
    const __half xh = tc::to_f16(x);
    *(__half*)(hi + off) = xh;
    *(__half*)(lo + off) = tc::to_f16(x - __half2float(xh));
  };
  if (use_tma) {
    __syncthreads();  // the barrier initialisation is visible to the waiting threads
    tc::mbar_wait(bars + 2, phs);
    phs ^= 1;
  }
  for (int i = tid; i < 128 * dh; i += MF_THREADS) {
    const int r = i / dh, d = i - r * dh;
    const float x = (q0 + r < T) ? (use_tma ? stg[i] : __ldg(src + (size_t)(q0 + r) * ld + 2 * D + d)) * scale : 0.f;
    split_store(Qh, Ql, (uint32_t)((d >> 6) * 16384) + tc::sw128_offset16(r, d & 63), x);
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot, tmem_o = tmem + 64;
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);

  const int qi = q0 + tid;
  const int len = lengths ? (int)lengths[b] : T;
  const int nk = causal ? min(qi + 1, T) : T;
  const int nvalid = (qi < len || !lengths) ? min(nk, len) : 0;  // keys j < nvalid take part
  const int kmax = causal ? min(q0 + 128, T) : T;                // keys any row of this tile can see
  const int nchunks = (kmax + MF_KEYS - 1) / MF_KEYS;
  uint32_t ph0 = 0, ph1 = 0;

  // thread 0: request the staging boxes of chunk c (rows past the utterance are masked when they are split)
  auto request = [&](int c, bool with_values) {
    tc::mbar_arrive_expect_tx(bars + 2, (uint32_t)((with_values ? 2 : 1) * MF_KEYS * dh * 4));
    tc::tma_load_2d(stg, &map_kv, bars + 2, h * dh, b * T + c * MF_KEYS);
    if (with_values) tc::tma_load_2d(stg + MF_KEYS * dh, &map_kv, bars + 2, D + h * dh, b * T + c * MF_KEYS);
  };
  if (use_tma && tid == 0) request(0, false);  // (the Q tile has been split: the staging area is free)
  auto load_keys = [&](int c, bool with_values) {
    if (use_tma) {
      tc::mbar_wait(bars + 2, phs);
      phs ^= 1;
    }
    for (int i = tid; i < MF_KEYS * dh; i += MF_THREADS) {
      const int j = i / dh, d = i - j * dh, kj = c * MF_KEYS + j;
      const bool ok = kj < T;
      split_store(Kh, Kl, (uint32_t)((d >> 6) * 8192) + tc::sw128_offset16(j, d & 63),
                  ok ? (use_tma ? stg[i] : __ldg(src + (size_t)kj * ld + d)) : 0.f);
      if (with_values)
        split_store(Vh, Vl, tc::sw128_offset16(d, j), ok ? (use_tma ? stg[MF_KEYS * dh + i] : __ldg(src + (size_t)kj * ld + D + d)) : 0.f);
    }
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    if (use_tma && tid == 0) {
      // the staging area has been read by every thread: the next chunk's boxes travel under this chunk's MMAs / softmax
      if (c + 1 < nchunks) request(c + 1, with_values);
      else if (!with_values) request(0, true);
    }
  };
  auto scores = [&]() {  // S[128 x 64] = Q K_c^T into TMEM columns [0, 64)
    if (tid == 0) {
      const uint32_t idesc = tc::idesc_f16(128, MF_KEYS);
      bool first = true;
      for (int ks = 0; ks < dh / 16; ++ks) {
        const uint32_t qo = (uint32_t)((ks >> 2) * 16384 + (ks & 3) * 32), ko = (uint32_t)((ks >> 2) * 8192 + (ks & 3) * 32);
        tc::mma_f16(tmem, tc::smem_desc_sw128(tc::smem_u32(Ql) + qo), tc::smem_desc_sw128(tc::smem_u32(Kh) + ko), idesc, !first);
        first = false;
        tc::mma_f16(tmem, tc::smem_desc_sw128(tc::smem_u32(Qh) + qo), tc::smem_desc_sw128(tc::smem_u32(Kl) + ko), idesc, 1);
        tc::mma_f16(tmem, tc::smem_desc_sw128(tc::smem_u32(Qh) + qo), tc::smem_desc_sw128(tc::smem_u32(Kh) + ko), idesc, 1);
      }
      tc::mma_commit(bars);
    }
    __syncwarp();
    tc::mbar_wait(bars, ph0);
    ph0 ^= 1;
    tc::tc_fence_after();
  };

  // ---- pass 1: running maximum and sum of this thread's row ----
  float mx = -INFINITY, sum = 0.f;
  for (int c = 0; c < nchunks; ++c) {
    load_keys(c, false);
    scores();
    const int k0 = c * MF_KEYS;
    float cm = -INFINITY;
    float v[MF_KEYS];
#pragma unroll
    for (int c0 = 0; c0 < MF_KEYS; c0 += 16) tc::tmem_ld16(lane_addr + c0, v + c0);
#pragma unroll
    for (int k = 0; k < MF_KEYS; ++k)
      if (k0 + k < nvalid) cm = fmaxf(cm, v[k]);
    if (cm > -INFINITY) {
      const float mn = fmaxf(mx, cm);
      float part = 0.f;
#pragma unroll
      for (int k = 0; k < MF_KEYS; ++k)
        if (k0 + k < nvalid) part += expf(v[k] - mn);
      sum = sum * expf(mx - mn) + part;  // expf(-inf) = 0 for the first valid chunk
      mx = mn;
    }
    tc::tc_fence_before();  // this thread's TMEM reads precede the next chunk's MMAs (ordered by the barrier in load_keys)
  }
  if (mx == -INFINITY) mx = 0.f;  // fully masked row: 0/0 = NaN like the reference
  const float inv = 1.f / sum;
  float* pr = nullptr;
  if (probs && qi < T)
    pr = probs + ((head_major ? (size_t)h * (gridDim.y / H) + b : (size_t)b * H + h) * T + qi) * T;
  const bool vec4 = (T & 3) == 0;

  // ---- pass 2: normalised probabilities and O += P V ----
  for (int c = 0; c < nchunks; ++c) {
    load_keys(c, true);
    scores();
    const int k0 = c * MF_KEYS;
#pragma unroll
    for (int c0 = 0; c0 < MF_KEYS; c0 += 16) {
      float v[16];
      tc::tmem_ld16(lane_addr + c0, v);
      __half2 ph[8];
#pragma unroll
      for (int k = 0; k < 16; k += 2) {
        const int j = k0 + c0 + k;
        const float p0 = j < nvalid ? expf(v[k] - mx) * inv : (j < nk ? 0.f * inv : 0.f);
        const float p1 = j + 1 < nvalid ? expf(v[k + 1] - mx) * inv : (j + 1 < nk ? 0.f * inv : 0.f);
        v[k] = p0;
        v[k + 1] = p1;
        // a fully masked row is NaN in the reference (0/0): its probabilities are written as such, its P operand is
        // zero and its context row is set to NaN at the end (fp16 conversion would turn the NaN into a finite value)
        ph[k >> 1] = __halves2half2(tc::to_f16(j < nvalid ? p0 : 0.f), tc::to_f16(j + 1 < nvalid ? p1 : 0.f));
      }
      *(uint4*)(P + tc::sw128_offset16(tid, c0)) = *(const uint4*)&ph[0];
      *(uint4*)(P + tc::sw128_offset16(tid, c0 + 8)) = *(const uint4*)&ph[4];
      if (pr) {
        if (vec4) {
#pragma unroll
          for (int k = 0; k < 16; k += 4)
            if (k0 + c0 + k < T) *(float4*)(pr + k0 + c0 + k) = make_float4(v[k], v[k + 1], v[k + 2], v[k + 3]);
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (k0 + c0 + k < T) pr[k0 + c0 + k] = v[k];
        }
      }
    }
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    if (tid == 0) {
      const uint32_t idesc = tc::idesc_f16(128, dh);
      for (int ks = 0; ks < MF_KEYS / 16; ++ks) {
        const uint32_t pa = tc::smem_u32(P) + ks * 32, vo = (uint32_t)(ks * 32);
        tc::mma_f16(tmem_o, tc::smem_desc_sw128(pa), tc::smem_desc_sw128(tc::smem_u32(Vl) + vo), idesc, (c > 0 || ks > 0) ? 1 : 0);
        tc::mma_f16(tmem_o, tc::smem_desc_sw128(pa), tc::smem_desc_sw128(tc::smem_u32(Vh) + vo), idesc, 1);
      }
      tc::mma_commit(bars + 1);
    }
    __syncwarp();
    tc::mbar_wait(bars + 1, ph1);  // P, V (and K) may be overwritten by the next chunk
    ph1 ^= 1;
    tc::tc_fence_after();
  }
  // causal tiles never look past kmax: the probabilities of the remaining keys are zero
  if (pr)
    for (int j = nchunks * MF_KEYS; j < T; ++j) pr[j] = 0.f;
  const uint32_t lane_o = tmem_o + ((uint32_t)(warp * 32) << 16);
  for (int d0 = 0; d0 < dh; d0 += 16) {
    float v[16];
    tc::tmem_ld16(lane_o + d0, v);
    if (nvalid == 0)
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = __int_as_float(0x7fc00000);
    if (qi < T) {
      float4* out = (float4*)(ctx + ((size_t)b * T + qi) * D + (size_t)h * dh + d0);
#pragma unroll
      for (int k = 0; k < 4; ++k) out[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

size_t mha_flash_smem_bytes(int dh) {
  const size_t nkc = (dh + 63) / 64;
  return 1024 + 2 * nkc * 16384 + 2 * nkc * 8192 + 2 * 16384 + 16384;
}
bool mha_flash_ok(int T, int D, int H) {
  const int dh = D / H;
  if (const char* e = getenv("SAT_MHA_TC"))
    if (atoi(e) == 0) return false;
  return T >= 1 && dh % 16 == 0 && dh <= 128 && D % 4 == 0;
}

size_t mha_tc_smem_bytes(int Tp, int dh) {
  const size_t nchunk = (Tp + 63) / 64;
  return 1024 + 2 * 16384 + 2 * 256 * 128 + 2 * nchunk * dh * 128 + 1024 + nchunk * 16384;
}
bool mha_tc_ok(int T, int D, int H) {
  const int dh = D / H;
  if (const char* e = getenv("SAT_MHA_TC"))
    if (atoi(e) == 0) return false;
  return T <= 256 && dh % 16 == 0 && dh <= 64 && D % 4 == 0 && mha_tc_smem_bytes((T + 15) & ~15, dh) <= 200 * 1024;
}

}  // namespace

MhaDev mha_upload(DevPool& pool, const Blob& W, const std::string& P, const std::string& TP, int Din, int D, int H) {
  SAT_REQUIRE(H >= 1 && D % H == 0, "model_dim %d is not a multiple of num_heads %d", D, H);
  MhaDev m;
  m.Din = Din; m.D = D; m.H = H;
  std::vector<float> w((size_t)Din * 3 * D), b((size_t)3 * D);
  const char* names[3] = {"key", "value", "query"};
  for (int j = 0; j < 3; ++j) {
    const float* k = W.get(P + "/" + names[j] + "/kernel", (int64_t)Din * D);
    const float* bb = W.get(P + "/" + names[j] + "/bias", D);
    for (int kk = 0; kk < Din; ++kk)
      for (int n = 0; n < D; ++n) w[(size_t)kk * 3 * D + (size_t)j * D + n] = k[(size_t)kk * D + n];
    for (int n = 0; n < D; ++n) b[(size_t)j * D + n] = bb[n];
  }
  m.Wqkv = pool.upload(w);
  m.bqkv = pool.upload(b);
  m.Wo = pool.upload(W.get(P + "/output/kernel", (int64_t)D * D), (size_t)D * D);
  m.bo = pool.upload(W.get(P + "/output/bias", D), D);
  if (!TP.empty()) {
    m.Wt = pool.upload(W.get(TP + "/kernel", (int64_t)D * D), (size_t)D * D);
    m.bt = pool.upload(W.get(TP + "/bias", D), D);
  }
  return m;
}

size_t mha_scratch_floats(const MhaDev& m, int64_t rows) { return (size_t)rows * (size_t)(3 * m.D + 2 * m.D) + 64; }

static void scaled_dot_core(const MhaDev& m, const float* qkv, int B, int T, int causal, const int64_t* pad_lengths,
                            float* ctx, float* probs, int head_major, cudaStream_t st);

static void attention_core(const MhaDev& m, const float* x, int B, int T, int causal, const int64_t* pad_lengths,
                           float* ctx, float* qkv, float* probs, int head_major, cudaStream_t st) {
  if (m.tq.img && m.xpack && m.tq.Npad == 3 * m.D) {
    dense_tc3_launch(m.tq, x, m.Din, (size_t)B * T, ACT_NONE, nullptr, 0, qkv, 3 * m.D, m.xpack, st);
  } else {
    GemmArgs g;
    g.X = x; g.Bn = 1; g.T = B * T; g.Cin = m.Din; g.ldx = m.Din;
    g.W = m.Wqkv; g.N = 3 * m.D; g.shift = m.bqkv; g.C = qkv; g.ldc = 3 * m.D;
    gemm_launch(g, st);
  }
  scaled_dot_core(m, qkv, B, T, causal, pad_lengths, ctx, probs, head_major, st);
}

// ScaledDotProductAttentionMechanism.__call__ (modules/self_attention.py:45-65) on projected heads: qkv [B][T][3D]
// (key | value | query, head h = columns h * dh ..) -> ctx [B][T][D], probs [B][H][T][T]; only m.D and m.H are read
static void scaled_dot_core(const MhaDev& m, const float* qkv, int B, int T, int causal, const int64_t* pad_lengths,
                            float* ctx, float* probs, int head_major, cudaStream_t st) {
  const int dh = m.D / m.H, Tpad = (T + 3) & ~3;
  if (mha_tc_ok(T, m.D, m.H)) {
    const int Tp = (T + 15) & ~15;
    size_t smem = mha_tc_smem_bytes(Tp, dh);
    const size_t staging = (size_t)(128 + 2 * Tp) * dh * sizeof(float);
    int use_tma = smem + staging <= 200 * 1024 && (((uintptr_t)qkv) & 15) == 0;
    if (const char* e = getenv("SAT_MHA_TMA")) use_tma = use_tma && atoi(e) != 0;
    CUtensorMap map_q, map_kv;
    memset(&map_q, 0, sizeof(map_q));
    memset(&map_kv, 0, sizeof(map_kv));
    if (use_tma) {
      smem += staging;
      make_tensor_map_2d_plain(&map_q, qkv, (uint64_t)B * T, (uint64_t)3 * m.D, (uint64_t)3 * m.D, (uint32_t)dh, 128);
      make_tensor_map_2d_plain(&map_kv, qkv, (uint64_t)B * T, (uint64_t)3 * m.D, (uint64_t)3 * m.D, (uint32_t)dh, (uint32_t)Tp);
    }
    SAT_CUDA(cudaFuncSetAttribute(mha_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div(T, 128), (unsigned)(B * m.H));
    mha_tc_kernel<<<grid, MT_THREADS, smem, st>>>(map_q, map_kv, use_tma, qkv, T, Tp, m.D, m.H, causal,
                                                  (const long long*)pad_lengths, ctx, probs, head_major);
    count_launch();
    check_launch("mha_tc_kernel");
    return;
  }
  if (mha_flash_ok(T, m.D, m.H)) {
    size_t smem = mha_flash_smem_bytes(dh);
    const size_t staging = (size_t)128 * dh * sizeof(float);  // Q tile = K + V chunk
    int use_tma = smem + staging <= 220 * 1024 && (((uintptr_t)qkv) & 15) == 0;
    if (const char* e = getenv("SAT_MHA_TMA")) use_tma = use_tma && atoi(e) != 0;
    CUtensorMap map_q, map_kv;
    memset(&map_q, 0, sizeof(map_q));
    memset(&map_kv, 0, sizeof(map_kv));
    if (use_tma) {
      smem += staging;
      make_tensor_map_2d_plain(&map_q, qkv, (uint64_t)B * T, (uint64_t)3 * m.D, (uint64_t)3 * m.D, (uint32_t)dh, 128);
      make_tensor_map_2d_plain(&map_kv, qkv, (uint64_t)B * T, (uint64_t)3 * m.D, (uint64_t)3 * m.D, (uint32_t)dh, MF_KEYS);
    }
    SAT_CUDA(cudaFuncSetAttribute(mha_flash_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div(T, 128), (unsigned)(B * m.H));
    mha_flash_tc_kernel<<<grid, MF_THREADS, smem, st>>>(map_q, map_kv, use_tma, qkv, T, m.D, m.H, causal,
                                                        (const long long*)pad_lengths, ctx, probs, head_major);
    count_launch();
    check_launch("mha_flash_tc_kernel");
    return;
  }
  const size_t smem = (size_t)MHA_WARPS * (Tpad + dh) * sizeof(float);
  SAT_REQUIRE(smem <= 200 * 1024, "sequence too long for the SIMT attention kernel (T=%d)", T);
  SAT_CUDA(cudaFuncSetAttribute(mha_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)ceil_div(T, MHA_WARPS), (unsigned)(B * m.H));
  mha_attn_kernel<<<grid, MHA_WARPS * 32, smem, st>>>(qkv, T, m.D, m.H, causal, (const long long*)pad_lengths, ctx, probs,
                                                       head_major);
  count_launch();
  check_launch("mha_attn_kernel");
}

void mha_forward(const MhaDev& m, const float* x, int B, int T, int causal, const int64_t* pad_lengths,
                 float* out, float* probs, float* scratch, cudaStream_t st, int head_major) {
  const size_t rows = (size_t)B * T;
  float* qkv = scratch;
  float* ctx = qkv + rows * 3 * m.D;
  attention_core(m, x, B, T, causal, pad_lengths, ctx, qkv, probs, head_major, st);
  if (m.to.img && m.xpack && m.to.Npad == m.D) {
    dense_tc3_launch(m.to, ctx, m.D, rows, ACT_NONE, nullptr, 0, out, m.D, m.xpack, st);
    return;
  }
  GemmArgs g;
  g.X = ctx; g.Bn = 1; g.T = (int)rows; g.Cin = m.D; g.ldx = m.D;
  g.W = m.Wo; g.N = m.D; g.shift = m.bo; g.C = out; g.ldc = m.D;
  gemm_launch(g, st);
}

void sat_layer_forward(const MhaDev& m, const float* x, int B, int T, int causal, const int64_t* pad_lengths,
                       float* out, float* probs, float* scratch, cudaStream_t st, int head_major) {
  SAT_REQUIRE(m.Wt && m.Din == m.D, "SelfAttentionTransformer needs a transform layer and Din == D");
  const size_t rows = (size_t)B * T;
  float* a = scratch + rows * 4 * m.D;
  mha_forward(m, x, B, T, causal, pad_lengths, a, probs, scratch, st, head_major);
  if (m.tt.img && m.xpack && m.tt.Npad == m.D) {
    dense_tc3_launch(m.tt, a, m.D, rows, ACT_TANH, x, m.D, out, m.D, m.xpack, st);
    return;
  }
  GemmArgs g;
  g.X = a; g.Bn = 1; g.T = (int)rows; g.Cin = m.D; g.ldx = m.D;
  g.W = m.Wt; g.N = m.D; g.shift = m.bt; g.act = ACT_TANH; g.R = x; g.ldr = m.D; g.C = out; g.ldc = m.D;
  gemm_launch(g, st);
}

}  // namespace sat

using namespace sat;

extern "C" int sat_mha_forward(const sat_tensor* weights, int32_t n_weights, const char* prefix, const float* x,
                               int32_t B, int32_t T, int32_t Din, int32_t D, int32_t H, int32_t causal,
                               const int64_t* lengths, float* out, float* probs, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(weights && prefix && x && out && B >= 1 && T >= 1, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    Blob W(weights, n_weights);
    DevPool pool;
    MhaDev m = mha_upload(pool, W, prefix, "", Din, D, H);
    float* scratch = pool.alloc<float>(mha_scratch_floats(m, (int64_t)B * T), false);
    mha_forward(m, x, B, T, causal, lengths, out, probs, scratch, st);
    SAT_CUDA(cudaStreamSynchronize(st));
  });
}

extern "C" int sat_sdpa_forward(const float* qkv, int32_t B, int32_t T, int32_t D, int32_t H, int32_t causal,
                                const int64_t* lengths, float* ctx, float* probs, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(qkv && ctx && B >= 1 && T >= 1 && H >= 1 && D >= H && D % H == 0, "bad argument");
    MhaDev m;
    m.D = D; m.H = H; m.Din = D;
    scaled_dot_core(m, qkv, B, T, causal, lengths, ctx, probs, 0, (cudaStream_t)stream);
  });
}

extern "C" int sat_sat_forward(const sat_tensor* weights, int32_t n_weights, const char* prefix, const float* x,
                               int32_t B, int32_t T, int32_t D, int32_t H, int32_t causal, float* out, float* probs,
                               void* stream) {
  return guarded([&] {
    SAT_REQUIRE(weights && prefix && x && out && B >= 1 && T >= 1, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    Blob W(weights, n_weights);
    DevPool pool;
    const std::string P(prefix);
    MhaDev m = mha_upload(pool, W, P + "/mha", P + "/transform", D, D, H);
    float* scratch = pool.alloc<float>(mha_scratch_floats(m, (int64_t)B * T), false);
    sat_layer_forward(m, x, B, T, causal, nullptr, out, probs, scratch, st);
    SAT_CUDA(cudaStreamSynchronize(st));
  });
}

// ---- allocation-free variant of sat_mha_forward (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the
// qkv / context scratch of the attention lives in a file-scope device array of fixed size, so the entry point calls no
// cudaMalloc / cudaFree / allocating library routine.  Every GPU that loads libsat.so holds the array (8 MiB) until
// the process exits.  All calls share it and it keeps the last call's contents: the region a call uses is cleared
// first, and calls that overlap on the GPU (different streams, or different host threads) race.
static __device__ __align__(256) float g_mha_noalloc_scratch[SAT_MHA_NOALLOC_MAX_FLOATS];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
float* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// device address of the scratch array on the CURRENT device, resolved once per device.  In host code the symbol name
// is only a stand-in for the array, never a device address: it must go through cudaGetSymbolAddress.
float* noalloc_scratch_ptr() {
  int dev = 0;
  SAT_CUDA(cudaGetDevice(&dev));
  SAT_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "device index %d out of range", dev);
  if (!g_noalloc_ptr[dev]) {
    float* p = nullptr;
    SAT_CUDA(cudaGetSymbolAddress((void**)&p, g_mha_noalloc_scratch));
    SAT_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    g_noalloc_ptr[dev] = p;
  }
  return g_noalloc_ptr[dev];
}
}  // namespace

extern "C" int64_t sat_mha_noalloc_max_floats(void) { return (int64_t)SAT_MHA_NOALLOC_MAX_FLOATS; }

extern "C" int sat_mha_forward_noalloc(const float* wqkv, const float* bqkv, const float* wo, const float* bo,
                                       const float* x, int32_t B, int32_t T, int32_t Din, int32_t D, int32_t H,
                                       int32_t causal, const int64_t* lengths, float* out, float* probs, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(wqkv && bqkv && wo && bo && x && out && B >= 1 && T >= 1 && Din >= 1 && D >= 1, "bad argument");
    SAT_REQUIRE(H >= 1 && D % H == 0, "model_dim %d is not a multiple of num_heads %d", D, H);
    const int64_t need = (int64_t)B * T * 4 * D;  // qkv [B*T, 3D] + context [B*T, D]
    SAT_REQUIRE(need <= (int64_t)SAT_MHA_NOALLOC_MAX_FLOATS,
                "sat_mha_forward_noalloc: B*T*4*D = %lld floats of scratch exceed the fixed maximum of %lld "
                "(SAT_MHA_NOALLOC_MAX_FLOATS); use sat_mha_forward for larger problems",
                (long long)need, (long long)SAT_MHA_NOALLOC_MAX_FLOATS);
    cudaStream_t st = (cudaStream_t)stream;
    float* scratch = noalloc_scratch_ptr();
    // the array keeps the previous call's contents: start this call's region from zero
    SAT_CUDA(cudaMemsetAsync(scratch, 0, (size_t)need * sizeof(float), st));
    MhaDev m;
    m.Wqkv = const_cast<float*>(wqkv); m.bqkv = const_cast<float*>(bqkv);
    m.Wo = const_cast<float*>(wo); m.bo = const_cast<float*>(bo);
    m.Din = Din; m.D = D; m.H = H;
    mha_forward(m, x, B, T, causal, lengths, out, probs, scratch, st);
  });
}
