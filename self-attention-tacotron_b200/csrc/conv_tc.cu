// conv_tc.cu -- tensor-core implicit-GEMM conv1d for the dense contractions of the encoder (conv bank, proj1) and the
// post-net: tcgen05.mma kind::f16 (fp16 operands, fp32 accumulation in TMEM), both operands fed by TMA.
//
// Activations live in a zero-haloed fp16 buffer XP[B][Tp][C], Tp = T + 2*HALO (flat rows R = B*Tp): the input window
// of tap tau of a 128-row output tile is the same 128 x 64 box shifted by (tau - pad_left) rows, so 'SAME' padding and
// the utterance boundaries need no gather -- the halo rows are zero and TMA zero-fills outside the buffer.
// Weights are a K-major fp16 image Wimg[N][taps*Cin].  One CTA computes a 128 (rows) x 128 (channels) tile:
//   thread 0   : TMA producer over taps x (Cin/64) chunks, 5-stage ring of {A 16 KB, B 16 KB}
//   thread 32  : MMA issuer, 4 x (M=128, N=128, K=16) per chunk into a 128-column TMEM accumulator
//   4 warps    : epilogue -- TMEM -> registers -> folded BatchNorm scale/shift -> activation -> fp16 (haloed layout,
//                halo rows untouched) or fp32 (dense [B*T][ldc]) stores, 16 bytes per access.
// Restates the Conv1d of tacotron2 (conv -> BN(inference) -> activation) used by
// modules/module.py:77-110 (ZoneoutCBHG) and by PostNetV2.
#include "common.cuh"
#include "tc.cuh"

namespace sat {
namespace {

constexpr int CT_STAGES = 2;  // 2 x 32 KB: three CTAs per SM -- the tiles are latency-bound (TMA fill, epilogue), measured 2.48 (5 stages, 1 CTA) -> 2.14 (3, 2) -> 1.84 ms (2, 3) per encoder forward at B=256
constexpr int CT_STAGE_BYTES = 2 * 16384;
constexpr int CT_THREADS = 128;

struct ConvTcArgs {
  int R, Tp, T, halo;       // flat rows, rows per utterance incl. halos, valid rows, halo size
  int taps, pad_left, cin;  // cin % 64 == 0
  const float* scale;       // [N] folded BN (null = 1)
  const float* shift;       // [N] (null = 0)
  int act;
  __half* y16;              // haloed fp16 output [R][ld16] at channel offset (or null)
  int ld16;
  float* y32;               // dense fp32 output [B*T][ld32] (or null)
  int ld32;
  const float* res;         // optional fp32 residual [B*T][ldr] added before the store (fp32 output only)
  int ldr;
  int highway;              // 1: HighwayNet epilogue over interleaved (H_u, T_u) column pairs, x = res; writes N/2 columns
  int bank;                 // > 0: conv-bank form (ConvTcLaunch::bank): grid = (widths, row tiles), widest first
  const long long* lengths; // optional [B]: rows t >= lengths[b] are stored as zeros
};

__global__ void __launch_bounds__(CT_THREADS, 3)
conv_tc_kernel(const __grid_constant__ CUtensorMap xmap, const __grid_constant__ CUtensorMap wmap, const ConvTcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* ring = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t bars[2 * CT_STAGES + 1];
  __shared__ uint32_t tmem_slot;
  uint64_t* full = bars;
  uint64_t* empty = bars + CT_STAGES;
  uint64_t* done = bars + 2 * CT_STAGES;
  const int tid = threadIdx.x, warp = tid >> 5;
  // conv-bank form: blockIdx.x walks the widths (widest first, so the long tiles of a row block start first and the
  // 16 CTAs that read the same input rows run together), blockIdx.y the row tiles
  const int width = a.bank ? a.bank - (int)blockIdx.x : 0;
  const int r0 = (a.bank ? blockIdx.y : blockIdx.x) * 128, n0 = (a.bank ? width - 1 : (int)blockIdx.y) * 128;
  const int taps = a.bank ? width : a.taps;
  const int pad_left = a.bank ? (a.bank - 1) / 2 : a.pad_left;          // input offset of tap slot 0
  const int tap0 = a.bank ? (a.bank - 1) / 2 - (width - 1) / 2 : 0;      // first tap slot of this width
  const int cpc = a.cin / tc::CHUNK_K16, nchunks = taps * cpc;

  if (tid == 0) {
    for (int i = 0; i < 2 * CT_STAGES + 1; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_mbar_init();
    tc::tma_prefetch_desc(&xmap);
    tc::tma_prefetch_desc(&wmap);
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 128);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (tid == 0) {
    for (int c = 0; c < nchunks; ++c) {
      const int s = c % CT_STAGES, use = c / CT_STAGES;
      if (use > 0) tc::mbar_wait(empty + s, (use - 1) & 1);
      const int tl = c / cpc, c0 = (c - tl * cpc) * tc::CHUNK_K16, tau = tap0 + tl;
      tc::mbar_arrive_expect_tx(full + s, CT_STAGE_BYTES);
      tc::tma_load_2d(ring + s * CT_STAGE_BYTES, &xmap, full + s, c0, r0 + tau - pad_left);
      tc::tma_load_2d(ring + s * CT_STAGE_BYTES + 16384, &wmap, full + s, tau * a.cin + c0, n0);
    }
  } else if (tid == 32) {
    const uint32_t idesc = tc::idesc_f16(128, 128);
    for (int c = 0; c < nchunks; ++c) {
      const int s = c % CT_STAGES, use = c / CT_STAGES;
      tc::mbar_wait(full + s, use & 1);
      tc::tc_fence_after();
      const uint32_t a0 = tc::smem_u32(ring + s * CT_STAGE_BYTES), b0 = a0 + 16384;
#pragma unroll
      for (int k16 = 0; k16 < tc::CHUNK_K16 / tc::UMMA_K16; ++k16)
        tc::mma_f16(tmem, tc::smem_desc_sw128(a0 + k16 * 32), tc::smem_desc_sw128(b0 + k16 * 32), idesc, (c | k16) != 0);
      tc::mma_commit(empty + s);
    }
    tc::mma_commit(done);
  }
  __syncwarp();
  tc::mbar_wait(done, 0);
  tc::tc_fence_after();

  // ---- epilogue: one flat row per thread ----
  const int r = r0 + tid;
  const int b = r / a.Tp, tt = r - b * a.Tp - a.halo;
  const bool valid = r < a.R && tt >= 0 && tt < a.T;
  const bool dead = valid && a.lengths && tt >= (int)a.lengths[b];  // padded frame of a shorter utterance
  const size_t drow = (size_t)b * a.T + tt;
#pragma unroll 1
  for (int cb = 0; cb < 128; cb += 16) {
    float v[16];
    tc::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)cb, v);
    if (!valid) continue;
    if (a.highway) {
      // HighwayNet (tacotron2): y_u = relu(H_u) * sigmoid(T_u) + x_u * (1 - sigmoid(T_u)); 8 units here
      const int u0 = (n0 + cb) >> 1;
      const float4* xp = (const float4*)(a.res + drow * a.ldr + u0);
      const float4 x0 = __ldg(xp), x1 = __ldg(xp + 1);
      const float x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
      float y[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int n = n0 + cb + 2 * k;
        const float hh = fmaxf(v[2 * k] + __ldg(a.shift + n), 0.f);
        const float tt = 1.f / (1.f + expf(-(v[2 * k + 1] + __ldg(a.shift + n + 1))));
        y[k] = dead ? 0.f : hh * tt + x[k] * (1.f - tt);
      }
      if (a.y16) {
        __half2 h[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) h[k] = __halves2half2(tc::to_f16(y[2 * k]), tc::to_f16(y[2 * k + 1]));
        *(uint4*)(a.y16 + (size_t)r * a.ld16 + u0) = *(const uint4*)&h[0];
      }
      if (a.y32) {
        float4* dst = (float4*)(a.y32 + drow * a.ld32 + u0);
        dst[0] = make_float4(y[0], y[1], y[2], y[3]);
        dst[1] = make_float4(y[4], y[5], y[6], y[7]);
      }
      continue;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int n = n0 + cb + i;
      float x = v[i];
      if (a.scale) x *= __ldg(a.scale + n);
      if (a.shift) x += __ldg(a.shift + n);
      if (a.act == ACT_RELU) x = fmaxf(x, 0.f);
      else if (a.act == ACT_TANH) x = tanhf(x);
      v[i] = x;
    }
    if (a.y32 && a.res) {  // residual joins before both stores (proj2 + input)
      const float4* rp = (const float4*)(a.res + drow * a.ldr + n0 + cb);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 q = __ldg(rp + i);
        v[4 * i] += q.x; v[4 * i + 1] += q.y; v[4 * i + 2] += q.z; v[4 * i + 3] += q.w;
      }
    }
    if (dead) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = 0.f;
    }
    if (a.y16) {
      __half2 h[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) h[i] = __halves2half2(tc::to_f16(v[2 * i]), tc::to_f16(v[2 * i + 1]));
      uint4* dst = (uint4*)(a.y16 + (size_t)r * a.ld16 + n0 + cb);
      dst[0] = *(const uint4*)&h[0];
      dst[1] = *(const uint4*)&h[4];
    }
    if (a.y32) {
      float4* dst = (float4*)(a.y32 + drow * a.ld32 + n0 + cb);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 128);
}

// fp32 [B*T][C] (row stride ldx) -> interior of the haloed fp16 buffer [B][Tp][ldp] at channel offset 0
__global__ void pack_f16_haloed_kernel(const float* __restrict__ x, int ldx, int B, int T, int C, int Tp, int halo,
                                       __half* __restrict__ xp, int ldp) {
  const size_t n = (size_t)B * T * (C / 4);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % (C / 4));
    const size_t row = i / (C / 4);
    const int b = (int)(row / T), t = (int)(row % T);
    const float4 v = __ldg((const float4*)(x + row * ldx) + c4);
    __half2* dst = (__half2*)(xp + ((size_t)b * Tp + halo + t) * ldp + 4 * c4);
    dst[0] = __halves2half2(tc::to_f16(v.x), tc::to_f16(v.y));
    dst[1] = __halves2half2(tc::to_f16(v.z), tc::to_f16(v.w));
  }
}

// MaxPooling1D(pool 2, stride 1, 'same') over time on the haloed fp16 layout: y[t] = max(x[t], x[t+1]), y[T-1] = x[T-1]
__global__ void maxpool2_f16_haloed_kernel(const __half* __restrict__ x, int B, int T, int C, int Tp, int halo,
                                           __half* __restrict__ y) {
  const size_t n = (size_t)B * T * (C / 8);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % (C / 8));
    const size_t row = i / (C / 8);
    const int b = (int)(row / T), t = (int)(row % T);
    const size_t o = ((size_t)b * Tp + halo + t) * C + 8 * c8;
    uint4 u = __ldg((const uint4*)(x + o));
    if (t + 1 < T) {
      const uint4 w = __ldg((const uint4*)(x + o + C));
      __half2* p = (__half2*)&u;
      const __half2* q = (const __half2*)&w;
#pragma unroll
      for (int k = 0; k < 4; ++k) p[k] = __hmax2(p[k], q[k]);
    }
    *(uint4*)(y + o) = u;
  }
}

// zero the halo rows (HALO before and after every utterance) of a haloed fp16 buffer [B][Tp][C]
__global__ void zero_halo_kernel(__half* __restrict__ x, int B, int T, int Tp, int halo, int C) {
  const size_t per = (size_t)2 * halo * (C / 8), n = (size_t)B * per;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / per);
    const size_t q = i % per;
    const int hr = (int)(q / (C / 8)), c8 = (int)(q % (C / 8));
    const int row = hr < halo ? hr : T + hr;  // rows [0, halo) and [halo + T, Tp)
    *(uint4*)(x + ((size_t)b * Tp + row) * C + 8 * c8) = make_uint4(0u, 0u, 0u, 0u);
  }
}

}  // namespace

void conv_tc_zero_halo(__half* x, int B, int T, int C, cudaStream_t st) {
  const int Tp = T + 2 * CONV_TC_HALO;
  const size_t n = (size_t)B * 2 * CONV_TC_HALO * (C / 8);
  zero_halo_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 8), 256, 0, st>>>(x, B, T, Tp, CONV_TC_HALO, C);
  count_launch();
}

// K-major fp16 image [N][taps*Cin] of a TF conv kernel [taps][Cin][N]
__half* conv_tc_weight_image(DevPool& pool, const float* w, int taps, int cin, int n) {
  std::vector<__half> img((size_t)n * taps * cin);
  for (int t = 0; t < taps; ++t)
    for (int c = 0; c < cin; ++c)
      for (int o = 0; o < n; ++o) img[(size_t)o * taps * cin + (size_t)t * cin + c] = __float2half_rn(w[((size_t)t * cin + c) * n + o]);
  std::vector<float> packed((img.size() + 1) / 2);
  memcpy(packed.data(), img.data(), img.size() * sizeof(__half));
  return (__half*)pool.upload(packed);
}

// conv-bank image (ConvTcLaunch::bank): rows = 128 channels of width 1, then of width 2, ...; columns = tap slots x cin,
// slot s = input offset s - (KW - 1) / 2; a width's unused slots stay zero (and are never read)
__half* conv_tc_bank_image(DevPool& pool, const std::vector<const float*>& w, int cin) {
  const int KW = (int)w.size(), n = 128, PL = (KW - 1) / 2;
  std::vector<__half> img((size_t)KW * n * KW * cin, __float2half_rn(0.f));
  for (int k = 1; k <= KW; ++k)
    for (int t = 0; t < k; ++t) {
      const int slot = PL - (k - 1) / 2 + t;
      for (int c = 0; c < cin; ++c)
        for (int o = 0; o < n; ++o)
          img[((size_t)(k - 1) * n + o) * KW * cin + (size_t)slot * cin + c] = __float2half_rn(w[k - 1][((size_t)t * cin + c) * n + o]);
    }
  std::vector<float> packed((img.size() + 1) / 2);
  memcpy(packed.data(), img.data(), img.size() * sizeof(__half));
  return (__half*)pool.upload(packed);
}

bool conv_tc_supported(int cin, int cout) { return cin % 64 == 0 && cout % 128 == 0; }

void conv_tc_pack(const float* x, int ldx, int B, int T, int C, __half* xp, int ldp, cudaStream_t st) {
  const int Tp = T + 2 * CONV_TC_HALO;
  const size_t n = (size_t)B * T * (C / 4);
  pack_f16_haloed_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, st>>>(x, ldx, B, T, C, Tp, CONV_TC_HALO, xp, ldp);
  count_launch();
}

void conv_tc_maxpool2(const __half* x, int B, int T, int C, __half* y, cudaStream_t st) {
  const int Tp = T + 2 * CONV_TC_HALO;
  const size_t n = (size_t)B * T * (C / 8);
  maxpool2_f16_haloed_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, st>>>(x, B, T, C, Tp, CONV_TC_HALO, y);
  count_launch();
}

namespace {
// fp32 [rows][ldx] -> fp16 [halo + rows][3K] = [xh | xl | xh], xh = f16(x), xl = f16(x - xh); optional length mask
__global__ void pack3_f16_kernel(const float* __restrict__ x, int ldx, size_t rows, int K, int halo, __half* __restrict__ xp,
                                 const long long* __restrict__ lengths, int T) {
  const int K4 = K / 4;
  const size_t n = rows * K4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % K4);
    const size_t row = i / K4;
    float4 v = __ldg((const float4*)(x + row * ldx) + c4);
    if (lengths && (long long)(row % T) >= lengths[row / T]) v = make_float4(0.f, 0.f, 0.f, 0.f);
    const float f[4] = {v.x, v.y, v.z, v.w};
    __half h[4], l[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      h[k] = tc::to_f16(f[k]);
      l[k] = tc::to_f16(f[k] - __half2float(h[k]));
    }
    __half* dst = xp + (row + halo) * (size_t)(3 * K) + 4 * c4;
    const uint2 hv = *(const uint2*)h, lv = *(const uint2*)l;
    *(uint2*)dst = hv;
    *(uint2*)(dst + K) = lv;
    *(uint2*)(dst + 2 * K) = hv;
  }
}
}  // namespace

DenseTc3 dense_tc3_upload(DevPool& pool, const float* w, int ldw, const float* bias, int K, int N) {
  DenseTc3 d;
  if (K % 64 != 0 || K <= 0 || N <= 0) return d;
  d.K = K; d.N = N; d.Npad = (N + 127) / 128 * 128;
  std::vector<__half> img((size_t)d.Npad * 3 * K, __float2half_rn(0.f));
  for (int k = 0; k < K; ++k)
    for (int n = 0; n < N; ++n) {
      const float v = w[(size_t)k * ldw + n];
      const __half hi = __float2half_rn(v);
      __half* row = img.data() + (size_t)n * 3 * K;
      row[k] = hi;
      row[K + k] = hi;
      row[2 * K + k] = __float2half_rn(v - __half2float(hi));
    }
  std::vector<float> packed((img.size() + 1) / 2);
  memcpy(packed.data(), img.data(), img.size() * sizeof(__half));
  d.img = (const __half*)pool.upload(packed);
  std::vector<float> b(d.Npad, 0.f);
  if (bias)
    for (int n = 0; n < N; ++n) b[n] = bias[n];
  d.bias = pool.upload(b);
  return d;
}

size_t dense_tc3_pack_halves(size_t rows, int K) { return (rows + 2 * CONV_TC_HALO) * (size_t)(3 * K) + 64; }

void dense_tc3_launch(const DenseTc3& d, const float* x, int ldx, size_t rows, int act, const float* res, int ldr, float* y,
                      int ldy, __half* pack, cudaStream_t st, const int64_t* lengths, int T) {
  SAT_REQUIRE(d.img && ldy >= d.Npad && rows > 0 && rows < ((size_t)1 << 30), "dense_tc3_launch: bad arguments");
  const size_t n = rows * (d.K / 4);
  pack3_f16_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, st>>>(
      x, ldx, rows, d.K, CONV_TC_HALO, pack, (const long long*)lengths, T);
  count_launch();
  ConvTcLaunch L;
  L.xp = pack; L.ldxp = 3 * d.K; L.B = 1; L.T = (int)rows; L.cin = 3 * d.K; L.cout = d.Npad; L.taps = 1; L.pad_left = 0;
  L.wimg = d.img; L.shift = d.bias; L.act = act; L.y32 = y; L.ld32 = ldy; L.res = res; L.ldr = ldr;
  conv_tc_launch(L, st);
}

void conv_tc_launch(const ConvTcLaunch& L, cudaStream_t st) {
  SAT_REQUIRE(conv_tc_supported(L.cin, L.cout), "tensor-core conv needs Cin %% 64 == 0 and Cout %% 128 == 0 (got %d, %d)", L.cin, L.cout);
  const int taps_total = L.bank ? L.bank : L.taps, pad_left = L.bank ? (L.bank - 1) / 2 : L.pad_left;
  SAT_REQUIRE(taps_total >= 1 && taps_total <= 2 * CONV_TC_HALO && pad_left <= CONV_TC_HALO && taps_total - 1 - pad_left <= CONV_TC_HALO,
              "conv taps %d exceed the activation halo", taps_total);
  SAT_REQUIRE(!L.bank || (L.cout == L.bank * 128 && !L.highway && !L.y32), "conv-bank form: cout = widths * 128, fp16 output");
  const int Tp = L.T + 2 * CONV_TC_HALO;
  const uint64_t R = (uint64_t)L.B * Tp;
  CUtensorMap xmap, wmap;
  make_tensor_map_2d_f16(&xmap, L.xp, R, (uint64_t)L.cin, (uint64_t)L.ldxp, 128);
  make_tensor_map_2d_f16(&wmap, L.wimg, (uint64_t)L.cout, (uint64_t)taps_total * L.cin, (uint64_t)taps_total * L.cin, 128);
  ConvTcArgs a;
  a.R = (int)R; a.Tp = Tp; a.T = L.T; a.halo = CONV_TC_HALO;
  a.taps = L.taps; a.pad_left = L.pad_left; a.cin = L.cin;
  a.scale = L.scale; a.shift = L.shift; a.act = L.act;
  a.y16 = L.y16; a.ld16 = L.ld16; a.y32 = L.y32; a.ld32 = L.ld32; a.res = L.res; a.ldr = L.ldr; a.highway = L.highway;
  a.bank = L.bank;
  a.lengths = (const long long*)L.lengths;
  SAT_REQUIRE(!L.highway || (L.res && L.shift), "highway epilogue needs the layer input (res) and the bias (shift)");
  static bool attr_set = false;
  const int smem = CT_STAGES * CT_STAGE_BYTES + 1024;
  if (!attr_set) {
    SAT_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  dim3 grid((unsigned)((R + 127) / 128), (unsigned)(L.cout / 128));
  if (L.bank) grid = dim3((unsigned)L.bank, (unsigned)((R + 127) / 128));
  conv_tc_kernel<<<grid, CT_THREADS, smem, st>>>(xmap, wmap, a);
  count_launch();
  check_launch("conv_tc_kernel");
}

}  // namespace sat
