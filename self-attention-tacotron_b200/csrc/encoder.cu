// encoder.cu -- sat_encoder_*: embedding lookup + encoder prenet + ZoneoutCBHG (+ self-attention branch).
//
// Restates (citations into /root/reference):
//   Embedding lookup                  models/models.py:28,54 (tacotron2 Embedding)
//   ZoneoutEncoderV1.call             modules/module.py:336-339
//   SelfAttentionCBHGEncoder.call     modules/module.py:425-438
//   ZoneoutCBHG.call                  modules/module.py:77-110
//     conv bank K=1..max (Conv1d = conv -> BN(inference) -> relu, A.2) -> concat -> MaxPooling1D(2,1,SAME)
//     -> proj1 (k3, relu) -> proj2 (k3, linear) -> + input -> [adjustment Dense] -> highway x n (A.3)
//     -> bidirectional ZoneoutLSTM with sequence_length (A.4, A.12)
//
// Dense / conv stages are implicit-GEMM launches (gemm.cu) with BN folded into scale/shift, the
// embedding gather and the max-pool fused on load and the highway gate fused into the epilogue.
// The BiLSTM input projection is one time-batched GEMM; the recurrence runs in a kernel where each
// CTA owns a few (row, direction) sequences for all T steps (rows are independent -> no grid sync).
#include "common.cuh"
#include "tc.cuh"
#include <cstdlib>

namespace sat {
namespace {

constexpr int LSTM_NT = 512;

__device__ __forceinline__ float sigm(float x) { return 1.f / (1.f + expf(-x)); }

// xproj [B][T][2][U][4] (bias included); Wh [2][U(k)][U][4]; out [B][T][2U] (fw | bw), zero past the length.
// grid = (ceil(B / ROWS), 2).  Thread (u, kpart): partial gate sums of unit u over k in its quarter.
// The recurrent weights of the direction (U x U x 4 fp32 = 256 KB at U = 128 -- the whole register file, or more
// than the shared memory of an SM) stay ON the SM for all T steps, split between the two: every thread keeps the
// float4 gate weights of half of its k values in registers (16 x 4 = 64 registers) and the other half in shared
// memory (W_s[kk][thread], 128 KB, conflict-free 16-byte reads), so a step reads nothing from L2 but the input
// projection of its position.  ROWS = sequences per CTA (1 / 2 / 4: small batches spread over more SMs).
template <int U, int ROWS>
__global__ void __launch_bounds__(LSTM_NT, 1) bilstm_kernel(const float* __restrict__ xproj, const float* __restrict__ Wh,
                                                            const long long* __restrict__ lengths, int B, int T,
                                                            float zc, float zh, float* __restrict__ out) {
  constexpr int KP = LSTM_NT / U;  // k partitions
  constexpr int KQ = U / KP;       // k values per partition
  constexpr int KR = KQ / 2;       // ... of which in registers; the rest in shared memory
  extern __shared__ __align__(16) float lstm_smem[];
  float4* W_s = (float4*)lstm_smem;                                     // [KQ - KR][LSTM_NT]
  float (*part)[ROWS][U][4] = (float (*)[ROWS][U][4])(lstm_smem + (size_t)(KQ - KR) * LSTM_NT * 4);  // [KP]
  float (*h_s)[U] = (float (*)[U])(lstm_smem + (size_t)(KQ - KR) * LSTM_NT * 4 + (size_t)KP * ROWS * U * 4);  // [ROWS]
  const int tid = threadIdx.x, u = tid % U, kp = tid / U;
  const int dir = blockIdx.y, b0 = blockIdx.x * ROWS;
  const float4* W4 = (const float4*)(Wh + (size_t)dir * U * U * 4);
  float4 w_reg[KR];
#pragma unroll
  for (int kk = 0; kk < KR; ++kk) w_reg[kk] = __ldg(W4 + (size_t)(kp * KQ + kk) * U + u);
#pragma unroll
  for (int kk = KR; kk < KQ; ++kk) W_s[(size_t)(kk - KR) * LSTM_NT + tid] = __ldg(W4 + (size_t)(kp * KQ + kk) * U + u);
  int maxlen = 0;
#pragma unroll
  for (int r = 0; r < ROWS; ++r)
    maxlen = max(maxlen, (b0 + r < B) ? min((int)lengths[b0 + r], T) : 0);
  for (int i = tid; i < ROWS * U; i += LSTM_NT) (&h_s[0][0])[i] = 0.f;
  // the thread that finalises (row r, unit u) keeps that cell's c and zoned h in registers
  float c_reg = 0.f, h_reg = 0.f;
  const int fr = kp;  // finalising thread of row r has kp == r (needs KP >= ROWS)
  static_assert(KP >= ROWS, "one finalising k-partition per row");
  const int my_len = (fr < ROWS && b0 + fr < B) ? min((int)lengths[b0 + fr], T) : 0;
  __syncthreads();
  for (int s = 0; s < maxlen; ++s) {
    // the input projection of this step's position does not depend on the recurrence: request it first
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    const int pos = dir == 0 ? s : my_len - 1 - s;  // backward direction walks the row reversed by its own length
    if (s < my_len) g = __ldg((const float4*)(xproj + ((((size_t)(b0 + fr) * T + pos) * 2 + dir) * U + u) * 4));
    float acc[ROWS][4];
#pragma unroll
    for (int r = 0; r < ROWS; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[r][q] = 0.f;
#pragma unroll
    for (int kk = 0; kk < KQ; ++kk) {
      const int k = kp * KQ + kk;
      const float4 w = kk < KR ? w_reg[kk < KR ? kk : 0] : W_s[(size_t)(kk < KR ? 0 : kk - KR) * LSTM_NT + tid];
#pragma unroll
      for (int r = 0; r < ROWS; ++r) {
        const float hv = h_s[r][k];
        acc[r][0] = fmaf(hv, w.x, acc[r][0]); acc[r][1] = fmaf(hv, w.y, acc[r][1]);
        acc[r][2] = fmaf(hv, w.z, acc[r][2]); acc[r][3] = fmaf(hv, w.w, acc[r][3]);
      }
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) *(float4*)part[kp][r][u] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
    __syncthreads();
    if (s < my_len) {
      const int r = fr, b = b0 + r;
#pragma unroll
      for (int q = 0; q < KP; ++q) {
        const float4 pp = *(const float4*)part[q][r][u];
        g.x += pp.x; g.y += pp.y; g.z += pp.z; g.w += pp.w;
      }
      const float cn = sigm(g.z + 1.0f) * c_reg + sigm(g.x) * tanhf(g.y);
      const float hn = sigm(g.w) * tanhf(cn);
      c_reg = cn - zc * (cn - c_reg);
      h_reg = hn - zh * (hn - h_reg);
      out[((size_t)b * T + pos) * (2 * U) + dir * U + u] = hn;  // cell output is the un-zoned h (A.4)
      h_s[r][u] = h_reg;  // every thread is past this step's h_s reads (barrier above)
    }
    __syncthreads();
  }
}

template <int U, int ROWS>
void bilstm_launch(const float* xproj, const float* Wh, const long long* lengths, int B, int T, float zc, float zh, float* out,
                   cudaStream_t st) {
  constexpr int KP = LSTM_NT / U, KQ = U / KP, KR = KQ / 2;
  const size_t smem = ((size_t)(KQ - KR) * LSTM_NT * 4 + (size_t)KP * ROWS * U * 4 + (size_t)ROWS * U) * sizeof(float);
  SAT_CUDA(cudaFuncSetAttribute(bilstm_kernel<U, ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)ceil_div(B, ROWS), 2);
  bilstm_kernel<U, ROWS><<<grid, LSTM_NT, smem, st>>>(xproj, Wh, lengths, B, T, zc, zh, out);
}

// The encoder prenet (two relu Dense layers, dropout off at inference: SURVEY.md A.1) is applied to the EMBEDDING of a
// symbol, so for id inputs it is a function of the id alone: prenet(embedding[id]) is tabulated per symbol when the
// weights are loaded (double accumulation) and the two GEMMs of the hot path become one gather.
// y32 [B*T][P] fp32 (residual of proj2); y16 (optional) the zero-haloed fp16 conv input [B][T + 2 halo][P].
// ids outside [0, rows) read as a zero row and set *err (tacotron2 Embedding asserts the range); with `lengths` the
// frames t >= lengths[b] are zeros (sat_encoder::mask_padding).
__global__ void prenet_table_gather_kernel(const long long* __restrict__ ids, const float* __restrict__ tab, int n_sym, int B,
                                           int T, int P, int Tp, int halo, const long long* __restrict__ lengths,
                                           float* __restrict__ y32, __half* __restrict__ y16, int* __restrict__ err) {
  const int P4 = P / 4;
  const size_t n = (size_t)B * T * P4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % P4);
    const size_t row = i / P4;
    const int b = (int)(row / T), t = (int)(row % T);
    const long long id = ids[row];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (id < 0 || id >= n_sym) {
      if (c4 == 0) *err = 1;
    } else if (!lengths || t < lengths[b]) {
      v = __ldg((const float4*)(tab + (size_t)id * P) + c4);
    }
    *((float4*)(y32 + row * P) + c4) = v;
    if (y16) {
      __half2* dst = (__half2*)(y16 + ((size_t)b * Tp + halo + t) * P + 4 * c4);
      dst[0] = __halves2half2(tc::to_f16(v.x), tc::to_f16(v.y));
      dst[1] = __halves2half2(tc::to_f16(v.z), tc::to_f16(v.w));
    }
  }
}

struct Conv {
  __half* wimg = nullptr;  // K-major fp16 image for the tensor-core path (null when the shape is not supported)
  float* w = nullptr;      // [k][cin][cout]
  float* scale = nullptr;  // folded BN
  float* shift = nullptr;
  int k = 1, cin = 0, cout = 0;
};

Conv upload_conv_bn(DevPool& pool, const Blob& W, const std::string& P, int k, int cin, int cout) {
  Conv c;
  c.k = k; c.cin = cin; c.cout = cout;
  c.w = pool.upload(W.get(P + "/kernel", (int64_t)k * cin * cout), (size_t)k * cin * cout);
  if (conv_tc_supported(cin, cout) && k <= 2 * CONV_TC_HALO)
    c.wimg = conv_tc_weight_image(pool, W.get(P + "/kernel", (int64_t)k * cin * cout), k, cin, cout);
  const float* gamma = W.get(P + "/bn/gamma", cout);
  const float* beta = W.get(P + "/bn/beta", cout);
  const float* mean = W.get(P + "/bn/mean", cout);
  const float* var = W.get(P + "/bn/var", cout);
  std::vector<float> sc(cout), sh(cout);
  for (int i = 0; i < cout; ++i) {
    // tf.layers.batch_normalization(training=False), epsilon 1e-3
    const float s = gamma[i] / sqrtf(var[i] + 1e-3f);
    sc[i] = s;
    sh[i] = beta[i] - mean[i] * s;
  }
  c.scale = pool.upload(sc);
  c.shift = pool.upload(sh);
  return c;
}

}  // namespace

// shared with postnet.cu
void conv_bn_act_launch(const float* x, int B, int T, int ldx, const float* w, const float* scale, const float* shift,
                        int k, int cin, int cout, int act, int maxpool2, const float* R, int ldr, float* y, int ldy,
                        cudaStream_t st, const int64_t* row_lengths) {
  GemmArgs g;
  g.row_lengths = row_lengths;  // optional: input frames t >= length read as zeros
  g.X = x; g.Bn = B; g.T = T; g.Cin = cin; g.ldx = ldx;
  g.W = w; g.taps = k; g.pad_left = (k - 1) / 2; g.N = cout;
  g.scale = scale; g.shift = shift; g.act = act; g.maxpool2 = maxpool2;
  g.R = R; g.ldr = ldr; g.C = y; g.ldc = ldy;
  gemm_launch(g, st);
}

}  // namespace sat

using namespace sat;

struct sat_encoder {
  sat_hparams hp;
  DevPool pool;
  Scratch ws;
  float* emb = nullptr;
  float *pre_w[2], *pre_b[2];
  std::vector<Conv> bank;
  Conv proj1, proj2;
  float *adj_w = nullptr, *adj_b = nullptr;
  std::vector<float*> hw_w, hw_b;  // [half][2*half] interleaved (H_u, T_u), bias [2*half]
  float *lstm_wx = nullptr, *lstm_bx = nullptr, *lstm_wh = nullptr;
  float *sap_w = nullptr, *sap_b = nullptr;
  std::vector<MhaDev> hops;
  int half = 0, cin = 0, sm_count = 148;
  bool tc = false;  // conv bank + proj1 on the tensor cores (fp16 operands, fp32 accumulation)
  bool tc_full = false;  // ... and prenet layer 1, proj2, the highway layers and the BiLSTM input projection as well
  float* pre_tab = nullptr;    // [num_symbols][P1]: prenet(embedding[id]) per symbol (prenet_table_gather_kernel)
  __half* pre1_img = nullptr;
  __half* bank_img = nullptr;  // all widths in one image (conv_tc_bank_image): the bank is ONE launch
  float *bank_scale = nullptr, *bank_shift = nullptr;
  std::vector<__half*> hw_img;
  __half* lstm_img = nullptr;
  Scratch ws16;     // haloed fp16 activations of that path
  // padded frames: 0 = the reference's batched graph (padding flows through the conv layers and the encoder
  // self-attention unmasked); 1 = every utterance is computed as if it were alone in the batch (frames
  // past its length are zero for every multi-tap convolution, keys past its length are masked in the self-attention)
  int mask_padding = 0;
  int* d_err = nullptr;  // sticky device flag: a symbol id outside [0, num_symbols) was seen
  int* h_err = nullptr;  // pinned host mirror, refreshed asynchronously at the end of every forward
  ~sat_encoder() {
    if (h_err) cudaFreeHost(h_err);
  }
};

extern "C" int sat_encoder_create(const sat_hparams* hpp, const sat_tensor* weights, int32_t n_weights, sat_encoder** out) {
  return guarded([&] {
    SAT_REQUIRE(hpp && weights && out, "null argument");
    const sat_hparams& hp = *hpp;
    SAT_REQUIRE(hp.encoder_kind == SAT_ENC_ZONEOUT_V1 || hp.encoder_kind == SAT_ENC_SELF_ATTENTION_CBHG, "Unknown encoder");
    Blob W(weights, n_weights);
    auto* h = new sat_encoder();
    std::unique_ptr<sat_encoder> guard(h);
    h->hp = hp;
    {
      int dev = 0;
      SAT_CUDA(cudaGetDevice(&dev));
      SAT_CUDA(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DevPool& pool = h->pool;
    const int E = hp.embedding_dim;
    SAT_REQUIRE(hp.encoder_prenet_units[0] > 0 && hp.encoder_prenet_units[1] > 0, "encoder prenet needs two layers");
    h->emb = pool.upload(W.get("embedding", (int64_t)hp.num_symbols * E), (size_t)hp.num_symbols * E);
    int d = E;
    for (int i = 0; i < 2; ++i) {
      const int u = hp.encoder_prenet_units[i];
      const std::string P = "encoder/prenet/dense_" + std::to_string(i);
      h->pre_w[i] = pool.upload(W.get(P + "/kernel", (int64_t)d * u), (size_t)d * u);
      h->pre_b[i] = pool.upload(W.get(P + "/bias", u), u);
      d = u;
    }
    h->cin = d;
    {
      const int P0 = hp.encoder_prenet_units[0], P1 = hp.encoder_prenet_units[1], S = hp.num_symbols;
      const float* emb = W.get("embedding", (int64_t)S * E);
      const float *w0 = W.get("encoder/prenet/dense_0/kernel", (int64_t)E * P0), *b0 = W.get("encoder/prenet/dense_0/bias", P0);
      const float *w1 = W.get("encoder/prenet/dense_1/kernel", (int64_t)P0 * P1), *b1 = W.get("encoder/prenet/dense_1/bias", P1);
      std::vector<float> tab((size_t)S * P1);
      std::vector<double> a0(P0), a1(P1);
      for (int sym = 0; sym < S; ++sym) {
        for (int j = 0; j < P0; ++j) a0[j] = b0[j];
        for (int k = 0; k < E; ++k) {
          const double x = emb[(size_t)sym * E + k];
          for (int j = 0; j < P0; ++j) a0[j] += x * (double)w0[(size_t)k * P0 + j];
        }
        for (int j = 0; j < P1; ++j) a1[j] = b1[j];
        for (int k = 0; k < P0; ++k) {
          const double x = a0[k] > 0.0 ? a0[k] : 0.0;
          for (int j = 0; j < P1; ++j) a1[j] += x * (double)w1[(size_t)k * P1 + j];
        }
        for (int j = 0; j < P1; ++j) tab[(size_t)sym * P1 + j] = (float)(a1[j] > 0.0 ? a1[j] : 0.0);
      }
      if (P1 % 4 == 0) h->pre_tab = pool.upload(tab);
      if (const char* e = getenv("SAT_ENC_PRENET_TABLE")) if (atoi(e) == 0) h->pre_tab = nullptr;
    }
    h->half = hp.cbhg_out_units / 2;
    SAT_REQUIRE(hp.cbhg_out_units % 2 == 0, "cbhg_out_units must be even");
    SAT_REQUIRE(h->half == 128, "the BiLSTM kernel is built for 128 units per direction (cbhg_out_units=256), got %d", h->half);
    const std::string C = "encoder/cbhg/";
    for (int k = 1; k <= hp.max_filter_width; ++k)
      h->bank.push_back(upload_conv_bn(pool, W, C + "conv1d_K" + std::to_string(k), k, d, hp.conv_channels));
    h->proj1 = upload_conv_bn(pool, W, C + "proj1", 3, hp.max_filter_width * hp.conv_channels, hp.projection1_out_channels);
    h->proj2 = upload_conv_bn(pool, W, C + "proj2", 3, hp.projection1_out_channels, hp.projection2_out_channels);
    SAT_REQUIRE(hp.projection2_out_channels == d, "CBHG residual needs projection2_out_channels == prenet output units");
    const int half = h->half;
    if (hp.projection2_out_channels != half) {
      h->adj_w = pool.upload(W.get(C + "adjustment/kernel", (int64_t)d * half), (size_t)d * half);
      h->adj_b = pool.upload(W.get(C + "adjustment/bias", half), half);
    }
    for (int i = 0; i < hp.num_highway; ++i) {
      const std::string P = C + "highway_" + std::to_string(i);
      const float* wh = W.get(P + "/H/kernel", (int64_t)half * half);
      const float* bh = W.get(P + "/H/bias", half);
      const float* wt = W.get(P + "/T/kernel", (int64_t)half * half);
      const float* bt = W.get(P + "/T/bias", half);
      std::vector<float> w((size_t)half * 2 * half), b((size_t)2 * half);
      for (int k = 0; k < half; ++k)
        for (int u = 0; u < half; ++u) {
          w[(size_t)k * 2 * half + 2 * u] = wh[(size_t)k * half + u];
          w[(size_t)k * 2 * half + 2 * u + 1] = wt[(size_t)k * half + u];
        }
      for (int u = 0; u < half; ++u) { b[2 * u] = bh[u]; b[2 * u + 1] = bt[u]; }
      h->hw_w.push_back(pool.upload(w));
      h->hw_b.push_back(pool.upload(b));
      if (conv_tc_supported(half, 2 * half)) h->hw_img.push_back(conv_tc_weight_image(pool, w.data(), 1, half, 2 * half));
    }
    // BiLSTM: TF kernel [in + U][4U] gate blocks i,j,f,o -> input part [in][dir][u][g], recurrent part [dir][k][u][g]
    {
      const int U = half, in = half;
      std::vector<float> wx((size_t)in * 2 * U * 4), bx((size_t)2 * U * 4), wh((size_t)2 * U * U * 4);
      const char* names[2] = {"lstm_fw", "lstm_bw"};
      for (int dir = 0; dir < 2; ++dir) {
        const float* k = W.get(C + names[dir] + "/kernel", (int64_t)(in + U) * 4 * U);
        const float* b = W.get(C + names[dir] + "/bias", 4 * U);
        for (int u = 0; u < U; ++u)
          for (int g = 0; g < 4; ++g) {
            const int src = g * U + u;
            bx[((size_t)dir * U + u) * 4 + g] = b[src];
            for (int kk = 0; kk < in; ++kk) wx[(size_t)kk * 2 * U * 4 + ((size_t)dir * U + u) * 4 + g] = k[(size_t)kk * 4 * U + src];
            for (int kk = 0; kk < U; ++kk) wh[(((size_t)dir * U + kk) * U + u) * 4 + g] = k[(size_t)(in + kk) * 4 * U + src];
          }
      }
      h->lstm_wx = pool.upload(wx);
      if (conv_tc_supported(in, 8 * U)) h->lstm_img = conv_tc_weight_image(pool, wx.data(), 1, in, 8 * U);
      h->lstm_bx = pool.upload(bx);
      h->lstm_wh = pool.upload(wh);
    }
    h->tc = h->proj1.wimg != nullptr;
    for (const Conv& c : h->bank) h->tc = h->tc && c.wimg != nullptr;
    if (h->tc && hp.conv_channels == 128 && hp.max_filter_width <= 2 * CONV_TC_HALO) {
      std::vector<const float*> ws;
      std::vector<float> sc, sh;
      for (int k = 1; k <= hp.max_filter_width; ++k) {
        const std::string P = C + "conv1d_K" + std::to_string(k);
        ws.push_back(W.get(P + "/kernel", (int64_t)k * d * 128));
        const float *gamma = W.get(P + "/bn/gamma", 128), *beta = W.get(P + "/bn/beta", 128);
        const float *mean = W.get(P + "/bn/mean", 128), *var = W.get(P + "/bn/var", 128);
        for (int i = 0; i < 128; ++i) {
          const float s = gamma[i] / sqrtf(var[i] + 1e-3f);
          sc.push_back(s);
          sh.push_back(beta[i] - mean[i] * s);
        }
      }
      h->bank_img = conv_tc_bank_image(pool, ws, d);
      h->bank_scale = pool.upload(sc);
      h->bank_shift = pool.upload(sh);
      if (const char* e = getenv("SAT_ENC_BANK_FUSED")) if (atoi(e) == 0) h->bank_img = nullptr;
    }
    if (const char* e = getenv("SAT_ENC_TC")) h->tc = h->tc && atoi(e) != 0;
    {
      const int P0 = hp.encoder_prenet_units[0], P1 = hp.encoder_prenet_units[1];
      h->tc_full = h->tc && !h->adj_w && h->proj2.wimg && conv_tc_supported(P0, P1) && conv_tc_supported(half, 2 * half) &&
                   conv_tc_supported(half, 8 * half) && hp.projection1_out_channels % 64 == 0 &&
                   hp.projection2_out_channels == half;
      if (const char* e = getenv("SAT_ENC_TC_FULL")) h->tc_full = h->tc_full && atoi(e) != 0;
      if (h->tc_full) {
        h->pre1_img = conv_tc_weight_image(pool, W.get("encoder/prenet/dense_1/kernel", (int64_t)P0 * P1), 1, P0, P1);
      }
    }
    if (hp.encoder_kind == SAT_ENC_SELF_ATTENTION_CBHG) {
      const int Ds = hp.self_attention_out_units;
      h->sap_w = pool.upload(W.get("encoder/sa_projection/kernel", (int64_t)hp.cbhg_out_units * Ds), (size_t)hp.cbhg_out_units * Ds);
      h->sap_b = pool.upload(W.get("encoder/sa_projection/bias", Ds), Ds);
      for (int i = 0; i < hp.self_attention_num_hop; ++i) {
        const std::string P = "encoder/sa_" + std::to_string(i);
        h->hops.push_back(mha_upload(pool, W, P + "/mha", P + "/transform", Ds, Ds, hp.self_attention_num_heads));
      }
    }
    h->d_err = pool.alloc<int>(1);
    SAT_CUDA(cudaMemset(h->d_err, 0, sizeof(int)));
    SAT_CUDA(cudaHostAlloc((void**)&h->h_err, sizeof(int), cudaHostAllocDefault));
    *h->h_err = 0;
    guard.release();
    *out = h;
  });
}

extern "C" void sat_encoder_destroy(sat_encoder* h) { delete h; }

extern "C" int sat_encoder_set_mask_padding(sat_encoder* h, int32_t on) {
  return guarded([&] {
    SAT_REQUIRE(h, "null argument");
    h->mask_padding = on ? 1 : 0;
  });
}

// no synchronisation: reports (and clears) the flag of the forwards that have FINISHED so far (their last operation
// copies the device flag into pinned host memory)
extern "C" int sat_encoder_poll_error(sat_encoder* h) {
  return guarded([&] {
    SAT_REQUIRE(h, "null argument");
    if (*h->h_err) {
      *h->h_err = 0;
      SAT_CUDA(cudaMemsetAsync(h->d_err, 0, sizeof(int), 0));
      fail(SAT_EINVAL, "a symbol id outside [0, %d) reached the encoder (its embedding was read as zeros)", h->hp.num_symbols);
    }
  });
}

// waits for the stream, then reports (and clears) the out-of-range-id flag of the forwards issued so far
extern "C" int sat_encoder_check(sat_encoder* h, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(h, "null argument");
    SAT_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    int flag = 0;
    SAT_CUDA(cudaMemcpy(&flag, h->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (flag) {
      SAT_CUDA(cudaMemset(h->d_err, 0, sizeof(int)));
      *h->h_err = 0;
      fail(SAT_EINVAL, "a symbol id outside [0, %d) reached the encoder (its embedding was read as zeros)", h->hp.num_symbols);
    }
  });
}

extern "C" int sat_encoder_uses_tensor_cores(sat_encoder* h, int32_t* used) {
  return guarded([&] {
    SAT_REQUIRE(h && used, "null argument");
    *used = h->tc ? 1 : 0;
  });
}

static int encoder_forward_impl(sat_encoder* h, const int64_t* ids, const float* embedded, const int64_t* lengths,
                                int32_t B, int32_t T, float* lstm_out, float* sa_out, float* sa_probs, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(h && (ids || embedded) && lengths && lstm_out && B >= 1 && T >= 1, "bad argument");
    const sat_hparams& hp = h->hp;
    const bool sa = hp.encoder_kind == SAT_ENC_SELF_ATTENTION_CBHG;
    SAT_REQUIRE(sa || (!sa_out && !sa_probs), "ZoneoutEncoderV1 has no self-attention outputs");
    SAT_REQUIRE(!sa || sa_out, "SelfAttentionCBHGEncoder needs sa_out");
    cudaStream_t st = (cudaStream_t)stream;
    if (*h->h_err) {  // an earlier forward (already finished) saw a bad id: refuse to go on silently
      *h->h_err = 0;
      SAT_CUDA(cudaMemsetAsync(h->d_err, 0, sizeof(int), st));
      fail(SAT_EINVAL, "a symbol id outside [0, %d) reached the encoder in an earlier forward", hp.num_symbols);
    }
    const int64_t* mlen = h->mask_padding ? lengths : nullptr;  // masks padded frames (see sat_encoder::mask_padding)
    const size_t rows = (size_t)B * T;
    const int P0 = hp.encoder_prenet_units[0], P1 = hp.encoder_prenet_units[1];
    const int Cb = hp.conv_channels, KW = hp.max_filter_width, half = h->half;
    const int Ds = hp.self_attention_out_units;
    // ---- workspace ----
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off += (n + 63) / 64 * 64; return o; };
    const size_t o_p0 = take(rows * P0), o_p1 = take(rows * P1), o_bank = take(rows * (size_t)KW * Cb);
    const size_t o_q1 = take(rows * hp.projection1_out_channels), o_q2 = take(rows * hp.projection2_out_channels);
    const size_t o_ha = take(rows * half), o_hb = take(rows * half), o_xp = take(rows * 8 * (size_t)half);
    size_t o_sa0 = 0, o_sa1 = 0, o_mha = 0;
    if (sa) {
      o_sa0 = take(rows * Ds);
      o_sa1 = take(rows * Ds);
      o_mha = take(mha_scratch_floats(h->hops.empty() ? MhaDev() : h->hops[0], (int64_t)rows));
    }
    float* ws = h->ws.get<float>(off);
    // ---- prenet: per-symbol table for id inputs, else the two dense layers (embedding gather fused into layer 0) ----
    const bool use_tab = ids && h->pre_tab;
    if (use_tab && !h->tc_full) {
      const size_t n = rows * (P1 / 4);
      prenet_table_gather_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, st>>>(
          (const long long*)ids, h->pre_tab, hp.num_symbols, B, T, P1, 0, 0, (const long long*)mlen, ws + o_p1, nullptr, h->d_err);
      count_launch();
    }
    if (!use_tab) {
      GemmArgs g;
      g.X = ids ? h->emb : embedded; g.Bn = 1; g.T = (int)rows; g.Cin = hp.embedding_dim; g.ldx = hp.embedding_dim;
      g.gather_ids = ids; g.gather_offset = 0;
      if (ids) { g.gather_rows = hp.num_symbols; g.err_flag = h->d_err; }
      g.W = h->pre_w[0]; g.N = P0; g.shift = h->pre_b[0]; g.act = ACT_RELU; g.C = ws + o_p0; g.ldc = P0;
      gemm_launch(g, st);
      if (!h->tc_full) {
        GemmArgs g1;
        g1.X = ws + o_p0; g1.Bn = 1; g1.T = (int)rows; g1.Cin = P0; g1.ldx = P0;
        g1.W = h->pre_w[1]; g1.N = P1; g1.shift = h->pre_b[1]; g1.act = ACT_RELU; g1.C = ws + o_p1; g1.ldc = P1;
        gemm_launch(g1, st);
      }
    }
    const int ldb = KW * Cb;
    const float* hw_in = nullptr;
    if (h->tc_full) {
      // ---- prenet layer 1 .. BiLSTM input projection, all tcgen05 implicit GEMMs over haloed fp16 activations ----
      // fp32 copies are kept only where an fp32 residual is needed (prenet output, highway inputs).
      const int Tp = T + 2 * CONV_TC_HALO;
      const size_t R = (size_t)B * Tp;
      auto al8 = [](size_t n) { return (n + 7) / 8 * 8; };
      const int Q1 = h->proj1.cout;
      const size_t n_p0 = al8(R * P0), n_x = al8(R * P1), n_bank = al8(R * (size_t)ldb), n_q1 = al8(R * Q1), n_h = al8(R * half);
      __half* p0_16 = h->ws16.get<__half>(n_p0 + n_x + 2 * n_bank + n_q1 + 2 * n_h + 64);
      __half* x16 = p0_16 + n_p0;
      __half* bank16 = x16 + n_x;
      __half* pool16 = bank16 + n_bank;
      __half* q1_16 = pool16 + n_bank;
      __half* hw16[2] = {q1_16 + n_q1, q1_16 + n_q1 + n_h};
      // only the inputs of multi-tap convolutions need zero halos
      conv_tc_zero_halo(x16, B, T, P1, st);
      conv_tc_zero_halo(pool16, B, T, ldb, st);
      conv_tc_zero_halo(q1_16, B, T, Q1, st);
      if (use_tab) {
        const size_t n = rows * (P1 / 4);
        prenet_table_gather_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 16), 256, 0, st>>>(
            (const long long*)ids, h->pre_tab, hp.num_symbols, B, T, P1, Tp, CONV_TC_HALO, (const long long*)mlen, ws + o_p1, x16,
            h->d_err);
        count_launch();
      } else {
        conv_tc_pack(ws + o_p0, P0, B, T, P0, p0_16, P0, st);
        ConvTcLaunch L;  // prenet layer 1: relu(x W + b) -> fp16 (bank input) and fp32 (residual of proj2)
        L.xp = p0_16; L.ldxp = P0; L.B = B; L.T = T; L.cin = P0; L.cout = P1; L.taps = 1; L.pad_left = 0;
        L.wimg = h->pre1_img; L.shift = h->pre_b[1]; L.act = ACT_RELU;
        L.y16 = x16; L.ld16 = P1; L.y32 = ws + o_p1; L.ld32 = P1; L.lengths = mlen;
        conv_tc_launch(L, st);
      }
      if (h->bank_img) {
        ConvTcLaunch L;  // all widths of the bank in one launch
        L.xp = x16; L.ldxp = P1; L.B = B; L.T = T; L.cin = P1; L.cout = KW * Cb; L.bank = KW;
        L.wimg = h->bank_img; L.scale = h->bank_scale; L.shift = h->bank_shift; L.act = ACT_RELU;
        L.y16 = bank16; L.ld16 = ldb; L.lengths = mlen;
        conv_tc_launch(L, st);
      } else {
        for (int k = 1; k <= KW; ++k) {
          const Conv& c = h->bank[k - 1];
          ConvTcLaunch L;
          L.xp = x16; L.ldxp = P1; L.B = B; L.T = T; L.cin = P1; L.cout = Cb; L.taps = k; L.pad_left = (k - 1) / 2;
          L.wimg = c.wimg; L.scale = c.scale; L.shift = c.shift; L.act = ACT_RELU;
          L.y16 = bank16 + (size_t)(k - 1) * Cb; L.ld16 = ldb; L.lengths = mlen;
          conv_tc_launch(L, st);
        }
      }
      conv_tc_maxpool2(bank16, B, T, ldb, pool16, st);
      {
        ConvTcLaunch L;  // proj1
        L.xp = pool16; L.ldxp = ldb; L.B = B; L.T = T; L.cin = ldb; L.cout = Q1; L.taps = 3; L.pad_left = 1;
        L.wimg = h->proj1.wimg; L.scale = h->proj1.scale; L.shift = h->proj1.shift; L.act = ACT_RELU;
        L.y16 = q1_16; L.ld16 = Q1; L.lengths = mlen;
        conv_tc_launch(L, st);
      }
      {
        ConvTcLaunch L;  // proj2 (linear) + prenet output
        L.xp = q1_16; L.ldxp = Q1; L.B = B; L.T = T; L.cin = Q1; L.cout = h->proj2.cout; L.taps = 3; L.pad_left = 1;
        L.wimg = h->proj2.wimg; L.scale = h->proj2.scale; L.shift = h->proj2.shift; L.act = ACT_NONE;
        L.res = ws + o_p1; L.ldr = P1; L.y32 = ws + o_q2; L.ld32 = h->proj2.cout; L.y16 = hw16[0]; L.ld16 = half;
        conv_tc_launch(L, st);
      }
      const float* x32 = ws + o_q2;
      float* pp[2] = {ws + o_ha, ws + o_hb};
      int cur16 = 0;
      for (size_t i = 0; i < h->hw_w.size(); ++i) {
        ConvTcLaunch L;  // highway layer: gate epilogue over interleaved (H_u, T_u) columns
        L.xp = hw16[cur16]; L.ldxp = half; L.B = B; L.T = T; L.cin = half; L.cout = 2 * half; L.taps = 1; L.pad_left = 0;
        L.wimg = h->hw_img[i]; L.shift = h->hw_b[i]; L.highway = 1; L.res = x32; L.ldr = half;
        L.y32 = pp[i & 1]; L.ld32 = half; L.y16 = hw16[cur16 ^ 1]; L.ld16 = half;
        conv_tc_launch(L, st);
        x32 = pp[i & 1];
        cur16 ^= 1;
      }
      {
        ConvTcLaunch L;  // BiLSTM input projection (bias included), gate-interleaved [rows][2][U][4]
        L.xp = hw16[cur16]; L.ldxp = half; L.B = B; L.T = T; L.cin = half; L.cout = 8 * half; L.taps = 1; L.pad_left = 0;
        L.wimg = h->lstm_img; L.shift = h->lstm_bx; L.act = ACT_NONE; L.y32 = ws + o_xp; L.ld32 = 8 * half;
        conv_tc_launch(L, st);
      }
    } else {
      if (h->tc) {
        // ---- conv bank, max-pool and proj1 on the tensor cores over zero-haloed fp16 activations (conv_tc.cu) ----
        const int Tp = T + 2 * CONV_TC_HALO;
        const size_t R = (size_t)B * Tp;
        const size_t n_x = R * P1, n_bank = R * (size_t)ldb;
        __half* x16 = h->ws16.get<__half>(n_x + 2 * n_bank + 64);
        __half* bank16 = x16 + (n_x + 7) / 8 * 8;
        __half* pool16 = bank16 + n_bank;
        SAT_CUDA(cudaMemsetAsync(x16, 0, (n_x + 2 * n_bank + 64) * sizeof(__half), st));
        conv_tc_pack(ws + o_p1, P1, B, T, P1, x16, P1, st);
        if (h->bank_img) {
          ConvTcLaunch L;  // all widths of the bank in one launch
          L.xp = x16; L.ldxp = P1; L.B = B; L.T = T; L.cin = P1; L.cout = KW * Cb; L.bank = KW;
          L.wimg = h->bank_img; L.scale = h->bank_scale; L.shift = h->bank_shift; L.act = ACT_RELU;
          L.y16 = bank16; L.ld16 = ldb; L.lengths = mlen;
          conv_tc_launch(L, st);
        } else {
          for (int k = 1; k <= KW; ++k) {
            const Conv& c = h->bank[k - 1];
            ConvTcLaunch L;
            L.xp = x16; L.ldxp = P1; L.B = B; L.T = T; L.cin = P1; L.cout = Cb; L.taps = k; L.pad_left = (k - 1) / 2;
            L.wimg = c.wimg; L.scale = c.scale; L.shift = c.shift; L.act = ACT_RELU;
            L.y16 = bank16 + (size_t)(k - 1) * Cb; L.ld16 = ldb; L.lengths = mlen;
            conv_tc_launch(L, st);
          }
        }
        conv_tc_maxpool2(bank16, B, T, ldb, pool16, st);
        ConvTcLaunch L;
        L.xp = pool16; L.ldxp = ldb; L.B = B; L.T = T; L.cin = ldb; L.cout = h->proj1.cout; L.taps = 3; L.pad_left = 1;
        L.wimg = h->proj1.wimg; L.scale = h->proj1.scale; L.shift = h->proj1.shift; L.act = ACT_RELU;
        L.y32 = ws + o_q1; L.ld32 = h->proj1.cout; L.lengths = mlen;
        conv_tc_launch(L, st);
      } else {
        // ---- conv bank -> concat buffer [rows][KW*Cb] (fp32 SIMT form) ----
        for (int k = 1; k <= KW; ++k) {
          const Conv& c = h->bank[k - 1];
          conv_bn_act_launch(ws + o_p1, B, T, P1, c.w, c.scale, c.shift, k, P1, Cb, ACT_RELU, 0, nullptr, 0,
                             ws + o_bank + (size_t)(k - 1) * Cb, ldb, st, mlen);
        }
        // ---- maxpool (fused on load) + proj1 ----
        conv_bn_act_launch(ws + o_bank, B, T, ldb, h->proj1.w, h->proj1.scale, h->proj1.shift, 3, ldb, h->proj1.cout, ACT_RELU,
                           1, nullptr, 0, ws + o_q1, h->proj1.cout, st, mlen);
      }
      // ---- proj2 + residual ----
      conv_bn_act_launch(ws + o_q1, B, T, h->proj1.cout, h->proj2.w, h->proj2.scale, h->proj2.shift, 3, h->proj1.cout,
                         h->proj2.cout, ACT_NONE, 0, ws + o_p1, P1, ws + o_q2, h->proj2.cout, st, mlen);
      hw_in = ws + o_q2;
      float* ping = ws + o_ha;
      float* pong = ws + o_hb;
      if (h->adj_w) {
        GemmArgs g;
        g.X = hw_in; g.Bn = 1; g.T = (int)rows; g.Cin = h->proj2.cout; g.ldx = h->proj2.cout;
        g.W = h->adj_w; g.N = half; g.shift = h->adj_b; g.C = ping; g.ldc = half;
        gemm_launch(g, st);
        hw_in = ping;
        std::swap(ping, pong);
      }
      for (size_t i = 0; i < h->hw_w.size(); ++i) {
        GemmArgs g;
        g.X = hw_in; g.Bn = 1; g.T = (int)rows; g.Cin = half; g.ldx = half;
        g.W = h->hw_w[i]; g.N = 2 * half; g.shift = h->hw_b[i]; g.highway = 1; g.R = hw_in; g.ldr = half; g.C = ping; g.ldc = half;
        gemm_launch(g, st);
        hw_in = ping;
        std::swap(ping, pong);
      }
    }
    // ---- BiLSTM: time-batched input projection, then the recurrence ----
    {
      if (!h->tc_full) {
        GemmArgs g;
        g.X = hw_in; g.Bn = 1; g.T = (int)rows; g.Cin = half; g.ldx = half;
        g.W = h->lstm_wx; g.N = 8 * half; g.shift = h->lstm_bx; g.C = ws + o_xp; g.ldc = 8 * half;
        gemm_launch(g, st);
      }
      SAT_CUDA(cudaMemsetAsync(lstm_out, 0, rows * 2 * half * sizeof(float), st));
      // sequences per CTA: as few as still fit one wave of the SMs (2 directions x ceil(B / rows) CTAs)
      const float zc = hp.zoneout_factor_cell, zh = hp.zoneout_factor_output;
      if (2 * B <= h->sm_count) bilstm_launch<128, 1>(ws + o_xp, h->lstm_wh, (const long long*)lengths, B, T, zc, zh, lstm_out, st);
      else if (B <= h->sm_count) bilstm_launch<128, 2>(ws + o_xp, h->lstm_wh, (const long long*)lengths, B, T, zc, zh, lstm_out, st);
      else bilstm_launch<128, 4>(ws + o_xp, h->lstm_wh, (const long long*)lengths, B, T, zc, zh, lstm_out, st);
      count_launch();
      check_launch("bilstm_kernel");
    }
    // ---- self-attention branch (modules/module.py:429-437); the padding mask is OFF in the reference ----
    if (sa) {
      GemmArgs g;
      g.X = lstm_out; g.Bn = 1; g.T = (int)rows; g.Cin = 2 * half; g.ldx = 2 * half;
      g.W = h->sap_w; g.N = Ds; g.shift = h->sap_b; g.C = ws + o_sa0; g.ldc = Ds;
      gemm_launch(g, st);
      const float* x = ws + o_sa0;
      const int H = hp.self_attention_num_heads;
      for (size_t i = 0; i < h->hops.size(); ++i) {
        float* y = (i + 1 == h->hops.size()) ? sa_out : ws + ((i & 1) ? o_sa0 : o_sa1);
        float* pr = sa_probs ? sa_probs + i * (size_t)H * B * T * T : nullptr;
        sat_layer_forward(h->hops[i], x, B, T, 0, mlen, y, pr, ws + o_mha, st, 1);
        x = y;
      }
      if (h->hops.empty()) SAT_CUDA(cudaMemcpyAsync(sa_out, x, rows * Ds * sizeof(float), cudaMemcpyDeviceToDevice, st));
    }
    if (ids) SAT_CUDA(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
  });
}

extern "C" int sat_encoder_forward(sat_encoder* h, const int64_t* ids, const int64_t* lengths, int32_t B, int32_t T,
                                   float* lstm_out, float* sa_out, float* sa_probs, void* stream) {
  if (!ids) { set_last_error("ids is null"); return SAT_EINVAL; }
  return encoder_forward_impl(h, ids, nullptr, lengths, B, T, lstm_out, sa_out, sa_probs, stream);
}

extern "C" int sat_encoder_forward_embedded(sat_encoder* h, const float* embedded, const int64_t* lengths, int32_t B,
                                            int32_t T, float* lstm_out, float* sa_out, float* sa_probs, void* stream) {
  if (!embedded) { set_last_error("embedded is null"); return SAT_EINVAL; }
  return encoder_forward_impl(h, nullptr, embedded, lengths, B, T, lstm_out, sa_out, sa_probs, stream);
}
