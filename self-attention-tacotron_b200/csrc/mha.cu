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

static void attention_core(const MhaDev& m, const float* x, int B, int T, int causal, const int64_t* pad_lengths,
                           float* ctx, float* qkv, float* probs, int head_major, cudaStream_t st) {
  GemmArgs g;
  g.X = x; g.Bn = 1; g.T = B * T; g.Cin = m.Din; g.ldx = m.Din;
  g.W = m.Wqkv; g.N = 3 * m.D; g.shift = m.bqkv; g.C = qkv; g.ldc = 3 * m.D;
  gemm_launch(g, st);
  const int dh = m.D / m.H, Tpad = (T + 3) & ~3;
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
