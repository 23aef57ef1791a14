// decoder.cu -- host side of sat_decoder_*: weight repacking, workspace layout, launch.
#include "common.cuh"
#include "decoder_kernel.cuh"
#include <cstdlib>

namespace sat {
using namespace dec;

namespace {

struct PackedDense {
  float* Wp = nullptr;
  float* bias = nullptr;
  float* Wrm = nullptr;   // row-major [K][Np] copy (pack_dense only)
  void* Wrm16 = nullptr;  // row-major fp16 [K][Np8] copy
  int K = 0, N = 0, n_tiles = 0, Np = 0, Np8 = 0, reps = 1;
  size_t rep_stride = 0, rep_stride16 = 0;
  DenseDesc desc() const {
    return DenseDesc{Wp, bias, K, N, n_tiles, -1, 0, Wrm, Np, Wrm16, Np8, reps, rep_stride, rep_stride16, 0, 0};
  }
};

// generic column-tile packing: column c of tile i comes from source column colmap[i*8+c] (or -1 = zero)
PackedDense pack_cols(DevPool& pool, const float* W, const float* bias, int K, int Nsrc, const std::vector<int>& colmap) {
  PackedDense d;
  d.K = K;
  d.N = (int)colmap.size();
  d.n_tiles = ceil_div(d.N, TILE_N);
  std::vector<float> wp((size_t)d.n_tiles * K * TILE_N, 0.f), bp((size_t)d.n_tiles * TILE_N, 0.f);
  for (int n = 0; n < d.N; ++n) {
    const int src = colmap[n];
    if (src < 0) continue;
    const int tile = n / TILE_N, c = n % TILE_N;
    for (int k = 0; k < K; ++k) wp[((size_t)tile * K + k) * TILE_N + c] = W[(size_t)k * Nsrc + src];
    if (bias) bp[n] = bias[src];
  }
  d.Wp = pool.upload(wp);
  d.bias = pool.upload(bp);
  return d;
}
// row-major copies [K][Np] (fp32) and [K][Np8] (fp16) of the first N columns of W[K][Nsrc], zero-padded to Nout columns:
// the operands of the row-split phases (row_dense)
void add_row_major(DevPool& pool, PackedDense& d, const float* W, int K, int Nsrc, int N, int Nout) {
  int reps = 1;  // replicas (see DenseDesc::reps); measured on B200: 1, 4 and 8 give the same phase times -> 1
  if (const char* e = getenv("SAT_WEIGHT_REPLICAS")) reps = std::max(1, std::min(16, atoi(e)));
  d.reps = reps;
  d.Np = (Nout + 3) / 4 * 4;
  d.rep_stride = ((size_t)K * d.Np + 63) / 64 * 64 + 32;  // + 32 floats: successive replicas start on different lines
  std::vector<float> rm(d.rep_stride * reps, 0.f);
  for (int r = 0; r < reps; ++r)
    for (int k = 0; k < K; ++k)
      for (int n = 0; n < N; ++n) rm[r * d.rep_stride + (size_t)k * d.Np + n] = W[(size_t)k * Nsrc + n];
  d.Wrm = pool.upload(rm);
  d.Np8 = (Nout + 7) / 8 * 8;
  d.rep_stride16 = ((size_t)K * d.Np8 + 127) / 128 * 128 + 64;
  std::vector<__half> rh(d.rep_stride16 * reps + 8, __float2half(0.f));
  for (int r = 0; r < reps; ++r)
    for (int k = 0; k < K; ++k)
      for (int n = 0; n < N; ++n) rh[r * d.rep_stride16 + (size_t)k * d.Np8 + n] = __float2half_rn(W[(size_t)k * Nsrc + n]);
  d.Wrm16 = pool.upload((const float*)rh.data(), (rh.size() + 1) / 2);
}
PackedDense pack_dense(DevPool& pool, const float* W, const float* bias, int K, int N) {
  std::vector<int> cm(N);
  for (int n = 0; n < N; ++n) cm[n] = n;
  PackedDense d = pack_cols(pool, W, bias, K, N, cm);
  add_row_major(pool, d, W, K, N, N, N);
  return d;
}
// LSTM kernel [K][4*units] gate order i,j,f,o -> tiles of 2 units x 4 gates
PackedDense pack_lstm(DevPool& pool, const float* W, const float* bias, int K, int units) {
  const int tiles = ceil_div(units, 2);
  std::vector<int> cm((size_t)tiles * TILE_N, -1);
  for (int u = 0; u < units; ++u)
    for (int g = 0; g < 4; ++g) cm[(u / 2) * TILE_N + (u % 2) * 4 + g] = g * units + u;
  PackedDense d = pack_cols(pool, W, bias, K, 4 * units, cm);
  d.N = tiles * TILE_N;
  return d;
}

__global__ void f32_to_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = tc::to_f16(src[i]);
}

__global__ void init_state_kernel(float* alpha0, float* alpha1, float* u0, float* u1, int B, int T) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  if (alpha0) alpha0[(size_t)b * T] = 1.f;
  if (alpha1) alpha1[(size_t)b * T] = 1.f;
  if (u0) u0[b] = 0.5f;
  if (u1) u1[b] = 0.5f;
}

// spk_term[n][b] = softsign(speaker_embed[b] . Ws[:, n] + bs[n])   (multi_speaker_modules.py:29)
__global__ void spk_term_kernel(const float* spk, const float* Ws, const float* bs, int B, int E, int N, int ld, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * N) return;
  const int n = i / B, b = i % B;
  float v = bs[n];
  for (int e = 0; e < E; ++e) v = fmaf(spk[(size_t)b * E + e], Ws[(size_t)e * N + n], v);
  out[(size_t)n * ld + b] = v / (1.f + fabsf(v));
}

// teacher-forced input frames of all steps, step-major: x[t][b][f] = 0 (<GO>) for t = 0, else the last n_feed frames
// of target step min(t-1, S_target-1)   (modules/helpers.py:44-58)
__global__ void teacher_inputs_kernel(const float* __restrict__ target, int B, int S, int S_target, int Mr, int Mnf,
                                      float* __restrict__ x) {
  const size_t n = (size_t)S * B * Mnf;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int f = (int)(i % Mnf);
    const size_t row = i / Mnf;
    const int b = (int)(row % B), t = (int)(row / B);
    x[i] = t == 0 ? 0.f : __ldg(target + ((size_t)b * S_target + min(t - 1, S_target - 1)) * Mr + (Mr - Mnf) + f);
  }
}
// x[t][b][n] += spk_term[n][b]  (MultiSpeakerPreNet speaker term, feature-major source)
__global__ void add_spk_rows_kernel(float* __restrict__ x, int rows, int B, int N, int ldx, const float* __restrict__ spk, int ld) {
  const size_t n = (size_t)rows * N;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % N);
    const size_t row = i / N;
    x[row * ldx + c] += __ldg(spk + (size_t)c * ld + (int)(row % B));
  }
}
// time-batched projection [B*S][ldy] -> mel [B][S][Mr], stop [B][S]
__global__ void split_projection_kernel(const float* __restrict__ y, size_t rows, int ldy, int Mr, float* __restrict__ mel,
                                        float* __restrict__ stop) {
  const size_t n = rows * (size_t)(Mr + 1);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % (Mr + 1));
    const size_t row = i / (Mr + 1);
    const float v = __ldg(y + row * ldy + c);
    if (c == Mr) stop[row] = v;
    else mel[row * Mr + c] = v;
  }
}

// sample ids: argmax over M per r-slot (modules/helpers.py:49-52); first maximum wins like tf.argmax
__global__ void argmax_kernel(const float* mel, long long rows, int M, int* ids) {
  const long long row = (long long)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = lane; i < M; i += 32) {
    const float v = mel[row * M + i];
    if (v > best) { best = v; bi = i; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) ids[row] = bi;
}

}  // namespace
}  // namespace sat

using namespace sat;

struct sat_decoder {
  sat_hparams hp;
  DevPool pool;
  Scratch ws;
  Scratch ws_tail;  // time-batched buffers of the teacher-forced form
  Scratch ws_pack;  // packed fp16 activations of the tensor-core dense layers (dense_tc3_launch)
  // tensor-core forms (fp16 hi + lo operand pairs, fp32-grade) of the time-batched dense layers: attention keys,
  // multi-speaker dense / prenet layer 1 over all steps, the decoder self-attention projections, frame + stop projection
  DenseTc3 t_mem[2], t_msd, t_pre1, t_outp, t_qkv[MAX_HOPS], t_o[MAX_HOPS], t_tr[MAX_HOPS];
  bool dense_tc = true;
  PackedDense pre0, msd, pre1, lstm_a, lstm_1, lstm_2, outp, dq[2];
  PackedDense outp_f;  // outp + prenet layer 0 folded onto the fed-back frame (self feed): [d_out][Mr + 1 + P0]
  bool has_outp_f = false;
  PackedDense sa_qkv[MAX_HOPS], sa_o[MAX_HOPS], sa_tr[MAX_HOPS];
  float* mem_layer[2] = {nullptr, nullptr};  // [Dm][Up]
  float* dq_cl[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // [mech][0: cluster of 2, 1: cluster of 4]: query-layer
                                                                  // weights as contiguous column slices [cl][A][Up / cl]
  float *att_v[2], *att_ba[2], *conv_w[2], *conv_b[2], *Wl[2], *Wt[2];
  float bt[2] = {0.f, 0.f};
  float *ms_Ws = nullptr, *ms_bs = nullptr;
  int Up[2] = {0, 0};
  int n_mech = 1, hops = 0, sm_count = 0, grid = 0;
  int P0 = 0, P1 = 0, Dctx = 0, Mr = 0, Mnf = 0;
  unsigned* barrier = nullptr;
  bool prof_on = false;  // phase counters requested (sat_decoder_profile)
  bool retire_rows = true;  // per-row stop rule: retire stopped rows (sat_decoder_set_row_retirement)
  long long* prof = nullptr;
  size_t max_dyn_smem = 0;
  int last_tc = 0, last_cl = 1;
  int step_next = -1;  // sat_decoder_step: the step the state in the workspace is ready for
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // bracket the persistent launch of the last run
  int last_B = 0, last_T = 0, last_S = 0;
  ~sat_decoder() {
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
  }
};

extern "C" int sat_decoder_create(const sat_hparams* hpp, const sat_tensor* weights, int32_t n_weights, sat_decoder** out) {
  return guarded([&] {
    SAT_REQUIRE(hpp && weights && out, "null argument");
    const sat_hparams& hp = *hpp;
    Blob W(weights, n_weights);
    auto* h = new sat_decoder();
    std::unique_ptr<sat_decoder> guard(h);
    h->hp = hp;
    const int M = hp.num_mels, r = hp.outputs_per_step, nf = hp.n_feed_frame;
    const int A = hp.attention_out_units, Dd = hp.decoder_out_units;
    SAT_REQUIRE(nf >= 1 && nf <= r, "n_feed_frame must be in [1, outputs_per_step]");
    h->n_mech = hp.decoder_kind == SAT_DEC_EXTENDED ? 1 : 2;
    h->hops = hp.decoder_kind == SAT_DEC_DUAL_SOURCE_TRANSFORMER ? hp.decoder_self_attention_num_hop : 0;
    SAT_REQUIRE(h->hops <= MAX_HOPS, "at most %d decoder self-attention hops are supported", MAX_HOPS);
    h->Mr = M * r;
    h->Mnf = M * nf;
    h->P0 = hp.decoder_prenet_units[0];
    h->P1 = hp.decoder_prenet_units[1];
    h->Dctx = 0;
    for (int i = 0; i < h->n_mech; ++i) h->Dctx += hp.memory_dim[i];
    DevPool& pool = h->pool;
    const std::string D = "decoder/";
    // prenet (modules/module.py:1511-1517)
    if (hp.multi_speaker_prenet) {
      const int E = hp.speaker_embedding_dim;
      h->pre0 = pack_dense(pool, W.get(D + "prenet/ms/dense0/kernel", (int64_t)h->Mnf * h->P0),
                           W.get(D + "prenet/ms/dense0/bias", h->P0), h->Mnf, h->P0);
      h->msd = pack_dense(pool, W.get(D + "prenet/ms/dense/kernel", (int64_t)h->P0 * h->P0),
                          W.get(D + "prenet/ms/dense/bias", h->P0), h->P0, h->P0);
      h->t_msd = dense_tc3_upload(pool, W.get(D + "prenet/ms/dense/kernel", (int64_t)h->P0 * h->P0), h->P0,
                                  W.get(D + "prenet/ms/dense/bias", h->P0), h->P0, h->P0);
      h->ms_Ws = pool.upload(W.get(D + "prenet/ms/speaker_projection/kernel", (int64_t)E * h->P0), (size_t)E * h->P0);
      h->ms_bs = pool.upload(W.get(D + "prenet/ms/speaker_projection/bias", h->P0), h->P0);
    } else {
      h->pre0 = pack_dense(pool, W.get(D + "prenet/dense_0/kernel", (int64_t)h->Mnf * h->P0),
                           W.get(D + "prenet/dense_0/bias", h->P0), h->Mnf, h->P0);
    }
    h->pre1 = pack_dense(pool, W.get(D + "prenet/dense_1/kernel", (int64_t)h->P0 * h->P1),
                         W.get(D + "prenet/dense_1/bias", h->P1), h->P0, h->P1);
    h->t_pre1 = dense_tc3_upload(pool, W.get(D + "prenet/dense_1/kernel", (int64_t)h->P0 * h->P1), h->P1,
                                 W.get(D + "prenet/dense_1/bias", h->P1), h->P0, h->P1);
    if (const char* e = getenv("SAT_DENSE_TC3")) h->dense_tc = atoi(e) != 0;
    // cells
    const int Ka = h->P1 + h->Dctx + A, K1 = A + h->Dctx + Dd, K2 = 2 * Dd;
    h->lstm_a = pack_lstm(pool, W.get(D + "attention_lstm/kernel", (int64_t)Ka * 4 * A), W.get(D + "attention_lstm/bias", 4 * A), Ka, A);
    h->lstm_1 = pack_lstm(pool, W.get(D + "lstm_1/kernel", (int64_t)K1 * 4 * Dd), W.get(D + "lstm_1/bias", 4 * Dd), K1, Dd);
    h->lstm_2 = pack_lstm(pool, W.get(D + "lstm_2/kernel", (int64_t)K2 * 4 * Dd), W.get(D + "lstm_2/bias", 4 * Dd), K2, Dd);
    // attention mechanisms
    for (int i = 0; i < h->n_mech; ++i) {
      const std::string P = D + "attention" + std::to_string(i + 1) + "/";
      const int U = hp.attention_units[i], Dm = hp.memory_dim[i], Up = att_padded_units(U);
      const int kind = hp.attention_kind[i];
      SAT_REQUIRE(U <= 256, "attention num_units > 256 unsupported (got %d)", U);
      SAT_REQUIRE(Dm <= NT, "attention memory depth > %d unsupported", NT);
      h->Up[i] = Up;
      auto padU = [&](const float* src, int rows) {  // [rows][U] -> [rows][Up]
        std::vector<float> v((size_t)rows * Up, 0.f);
        for (int rr = 0; rr < rows; ++rr)
          for (int u = 0; u < U; ++u) v[(size_t)rr * Up + u] = src[(size_t)rr * U + u];
        return pool.upload(v);
      };
      h->mem_layer[i] = padU(W.get(P + "memory_layer/kernel", (int64_t)Dm * U), Dm);
      h->t_mem[i] = dense_tc3_upload(pool, W.get(P + "memory_layer/kernel", (int64_t)Dm * U), U, nullptr, Dm, U);
      {
        std::vector<int> cm(Up, -1);
        for (int u = 0; u < U; ++u) cm[u] = u;
        h->dq[i] = pack_cols(pool, W.get(P + "query_layer/kernel", (int64_t)A * U), nullptr, A, U, cm);
        add_row_major(pool, h->dq[i], W.get(P + "query_layer/kernel", (int64_t)A * U), A, U, U, Up);
        // column slices for the cluster-split rows of the small-batch form (decoder_kernel.cuh, DecParams::cl)
        const float* wq = W.get(P + "query_layer/kernel", (int64_t)A * U);
        for (int ci = 0; ci < 2; ++ci) {
          const int cl = 2 << ci, sw = Up / cl;
          if (Up % (32 * cl)) continue;
          std::vector<float> sl((size_t)A * Up, 0.f);
          for (int r = 0; r < cl; ++r)
            for (int k = 0; k < A; ++k)
              for (int n = 0; n < sw; ++n)
                if (r * sw + n < U) sl[((size_t)r * A + k) * sw + n] = wq[(size_t)k * U + r * sw + n];
          h->dq_cl[i][ci] = pool.upload(sl);
        }
      }
      h->att_v[i] = padU(W.get(P + "attention_v", U), 1);
      std::vector<float> zeros(Up, 0.f);
      h->att_ba[i] = pool.upload(zeros);
      h->conv_w[i] = h->conv_b[i] = h->Wl[i] = h->Wt[i] = nullptr;
      if (kind == SAT_ATT_FORWARD || kind == SAT_ATT_LOCATION_SENSITIVE) {
        const int Kc = hp.attention_kernel, F = hp.attention_filters;
        SAT_REQUIRE(Kc <= 2 * HALO && F <= 32, "attention_kernel <= %d and attention_filters <= 32 required", 2 * HALO);
        SAT_REQUIRE(F <= 8 || (128 / std::min(Up, 128)) * F <= 512, "attention_filters %d too large for %d attention units", F, U);
        h->att_ba[i] = padU(W.get(P + "attention_bias", U), 1);
        h->conv_w[i] = pool.upload(W.get(P + "location_conv/kernel", (int64_t)Kc * F), (size_t)Kc * F);
        h->conv_b[i] = pool.upload(W.get(P + "location_conv/bias", F), F);
        h->Wl[i] = padU(W.get(P + "location_layer/kernel", (int64_t)F * U), F);
        if (kind == SAT_ATT_FORWARD && hp.use_transition_agent) {
          h->Wt[i] = pool.upload(W.get(P + "transition/kernel", Dm + U), Dm + U);
          h->bt[i] = W.get(P + "transition/bias", 1)[0];
        }
      }
    }
    // decoder self-attention hops (modules/module.py:363-371 inside rnn_wrappers.py:111-124)
    int d_out = Dd;
    for (int i = 0; i < h->hops; ++i) {
      const int Ds = hp.decoder_self_attention_out_units;
      SAT_REQUIRE(Ds == Dd, "decoder self-attention units must equal decoder_out_units");
      SAT_REQUIRE(Ds % hp.decoder_self_attention_num_heads == 0 && (Ds / hp.decoder_self_attention_num_heads) % 4 == 0 &&
                      Ds / hp.decoder_self_attention_num_heads <= 512, "unsupported head size");
      const std::string P = D + "sa_" + std::to_string(i) + "/";
      std::vector<float> w((size_t)Dd * 3 * Ds), b(3 * Ds);
      const char* names[3] = {"key", "value", "query"};
      for (int j = 0; j < 3; ++j) {
        const float* k = W.get(P + "mha/" + names[j] + "/kernel", (int64_t)Dd * Ds);
        const float* bb = W.get(P + "mha/" + names[j] + "/bias", Ds);
        for (int kk = 0; kk < Dd; ++kk)
          for (int n = 0; n < Ds; ++n) w[(size_t)kk * 3 * Ds + j * Ds + n] = k[(size_t)kk * Ds + n];
        for (int n = 0; n < Ds; ++n) b[j * Ds + n] = bb[n];
      }
      h->sa_qkv[i] = pack_dense(pool, w.data(), b.data(), Dd, 3 * Ds);
      h->t_qkv[i] = dense_tc3_upload(pool, w.data(), 3 * Ds, b.data(), Dd, 3 * Ds);
      h->t_o[i] = dense_tc3_upload(pool, W.get(P + "mha/output/kernel", (int64_t)Ds * Ds), Ds, W.get(P + "mha/output/bias", Ds), Ds, Ds);
      h->t_tr[i] = dense_tc3_upload(pool, W.get(P + "transform/kernel", (int64_t)Ds * Ds), Ds, W.get(P + "transform/bias", Ds), Ds, Ds);
      h->sa_o[i] = pack_dense(pool, W.get(P + "mha/output/kernel", (int64_t)Ds * Ds), W.get(P + "mha/output/bias", Ds), Ds, Ds);
      h->sa_tr[i] = pack_dense(pool, W.get(P + "transform/kernel", (int64_t)Ds * Ds), W.get(P + "transform/bias", Ds), Ds, Ds);
      d_out = Ds;
    }
    // frame + stop projection fused into one [d_out][Mr + 1] layer
    {
      const float* ko = W.get(D + "out_projection/kernel", (int64_t)d_out * h->Mr);
      const float* bo = W.get(D + "out_projection/bias", h->Mr);
      const float* ks = W.get(D + "stop_token_projection/kernel", d_out);
      const float* bs = W.get(D + "stop_token_projection/bias", 1);
      const int N = h->Mr + 1;
      std::vector<float> w((size_t)d_out * N), b(N);
      for (int k = 0; k < d_out; ++k) {
        for (int n = 0; n < h->Mr; ++n) w[(size_t)k * N + n] = ko[(size_t)k * h->Mr + n];
        w[(size_t)k * N + h->Mr] = ks[k];
      }
      for (int n = 0; n < h->Mr; ++n) b[n] = bo[n];
      b[h->Mr] = bs[0];
      h->outp = pack_dense(pool, w.data(), b.data(), d_out, N);
      h->t_outp = dense_tc3_upload(pool, w.data(), N, b.data(), d_out, N);
      // Self feed: the next input frame is the last n_feed_frame frames of this projection, and prenet layer 0 is a
      // dense layer on it -- two GEMVs in sequence on the critical path of every step.  Their composition is one
      // linear map of the decoder output: column Mr + 1 + j = sum_f W_out[:, f0 + f] W_pre0[f][j] (bias likewise),
      // formed here in double precision; the relu is applied when the columns are read.  (modules/module.py:599-617,
      // helpers: feed slice [:, -M * n_feed_frame:]; PreNet layer 0 / MultiSpeakerPreNet dense0.)
      const char* p0name = hp.multi_speaker_prenet ? "prenet/ms/dense0" : "prenet/dense_0";
      const float* kp = W.get(D + p0name + "/kernel", (int64_t)h->Mnf * h->P0);
      const float* bp = W.get(D + p0name + "/bias", h->P0);
      const int f0 = h->Mr - h->Mnf, NF = N + h->P0;
      std::vector<float> wf((size_t)d_out * NF), bf(NF);
      std::vector<double> acc(h->P0);
      for (int k = 0; k < d_out; ++k) {
        for (int n = 0; n < N; ++n) wf[(size_t)k * NF + n] = w[(size_t)k * N + n];
        std::fill(acc.begin(), acc.end(), 0.0);
        for (int f = 0; f < h->Mnf; ++f) {
          const double a = ko[(size_t)k * h->Mr + f0 + f];
          for (int j = 0; j < h->P0; ++j) acc[j] += a * (double)kp[(size_t)f * h->P0 + j];
        }
        for (int j = 0; j < h->P0; ++j) wf[(size_t)k * NF + N + j] = (float)acc[j];
      }
      for (int n = 0; n < N; ++n) bf[n] = b[n];
      for (int j = 0; j < h->P0; ++j) {
        double a = bp[j];
        for (int f = 0; f < h->Mnf; ++f) a += (double)bo[f0 + f] * (double)kp[(size_t)f * h->P0 + j];
        bf[N + j] = (float)a;
      }
      if (NF <= 4 * NT && (size_t)d_out * NF <= (size_t)1 << 22) {
        h->outp_f = pack_dense(pool, wf.data(), bf.data(), d_out, NF);
        h->has_outp_f = true;
      }
    }
    // launch geometry
    int dev = 0;
    SAT_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    SAT_CUDA(cudaGetDeviceProperties(&prop, dev));
    SAT_REQUIRE(prop.cooperativeLaunch, "device lacks cooperative launch");
    h->sm_count = prop.multiProcessorCount;
    h->max_dyn_smem = prop.sharedMemPerBlockOptin - 4096;
    SAT_CUDA(cudaFuncSetAttribute(decoder_persistent_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->max_dyn_smem));
    SAT_CUDA(cudaFuncSetAttribute(decoder_persistent_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->max_dyn_smem));
    SAT_CUDA(cudaFuncSetAttribute(decoder_persistent_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->max_dyn_smem));
    SAT_CUDA(cudaFuncSetAttribute(decoder_persistent_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)h->max_dyn_smem));
    h->grid = h->sm_count;
    h->barrier = pool.alloc<unsigned>(1);
    h->prof = pool.alloc<long long>(N_PROF);
    SAT_CUDA(cudaEventCreate(&h->ev0));
    SAT_CUDA(cudaEventCreate(&h->ev1));
    guard.release();
    *out = h;
  });
}

extern "C" void sat_decoder_destroy(sat_decoder* h) { delete h; }

extern "C" double sat_decoder_step_bytes(const sat_decoder* h, int32_t B, int32_t T, int32_t avg_cache_len) {
  // bytes_step(B) = W * 4 + B * [T * sum_i (U_i + D_i) * 4 + avg_cache * 2 * d_sa * 4 + state + io]
  // W = decoder parameters actually read by a step (counted once per step, un-padded); T = (mean) memory length.
  if (!h) return 0.0;
  const sat_hparams& hp = h->hp;
  const int A = hp.attention_out_units, Dd = hp.decoder_out_units;
  double w = (double)h->Mnf * h->P0 + h->P0 + (double)h->P0 * h->P1 + h->P1;
  if (hp.multi_speaker_prenet) w += (double)h->P0 * h->P0 + h->P0;
  w += (double)(h->P1 + h->Dctx + A) * 4 * A + 4 * A;
  w += (double)(A + h->Dctx + Dd) * 4 * Dd + 4 * Dd + (double)2 * Dd * 4 * Dd + 4 * Dd;
  double per_utt = 0;
  for (int i = 0; i < h->n_mech; ++i) {
    const int U = hp.attention_units[i];
    w += (double)A * U + U;
    if (hp.attention_kind[i] == SAT_ATT_FORWARD || hp.attention_kind[i] == SAT_ATT_LOCATION_SENSITIVE)
      w += U + (double)hp.attention_kernel * hp.attention_filters + hp.attention_filters + (double)hp.attention_filters * U;
    per_utt += (double)T * (U + hp.memory_dim[i]) * 4.0;
  }
  int d_out = Dd;
  for (int i = 0; i < h->hops; ++i) {
    const int Ds = hp.decoder_self_attention_out_units;
    w += 5.0 * ((double)Ds * Ds + Ds);
    per_utt += (double)avg_cache_len * 2.0 * Ds * 4.0;
    d_out = Ds;
  }
  w += (double)d_out * (h->Mr + 1) + h->Mr + 1;
  const double state = 4.0 * (2.0 * (A + 2 * Dd) * 2 + h->Dctx + 3.0 * T * h->n_mech);
  const double io = 4.0 * (h->Mnf + h->Mr + 1);
  return w * 4.0 + B * (per_utt + state + io);
}

// steps [t_begin, t_begin + t_count) of the run `iop` describes; t_count < 0: the whole run
static int decoder_run_impl(sat_decoder* h, const sat_decoder_io* iop, void* stream, int t_begin, int t_count,
                            int32_t* finished = nullptr) {
  return guarded([&] {
    SAT_REQUIRE(h && iop, "null argument");
    const bool stepping = t_count >= 0;
    const sat_decoder_io& io = *iop;
    const sat_hparams& hp = h->hp;
    cudaStream_t st = (cudaStream_t)stream;
    const int B = io.B, T = io.T_enc;
    SAT_REQUIRE(B >= 1 && B <= MAX_B, "batch must be in [1, %d] per call (got %d); split larger batches", MAX_B, B);
    SAT_REQUIRE(T >= 1 && T <= MAX_T, "T_enc must be in [1, %d] (got %d)", MAX_T, T);
    SAT_REQUIRE(io.max_steps >= 1 && io.max_steps <= 8192, "max_steps must be in [1, 8192]");
    SAT_REQUIRE(io.memory_lengths && io.mel && io.stop && io.stop_steps && io.n_steps, "missing required buffer");
    SAT_REQUIRE(io.feed_mode != SAT_FEED_TEACHER || (io.target && io.target_steps >= 1), "teacher feed needs target");
    SAT_REQUIRE(!hp.multi_speaker_prenet || io.speaker_embed, "multi-speaker prenet needs speaker_embed");
    const int Bp = (B + 3) / 4 * 4, ld = Bp;
    const int R = ceil_div(Bp, ROWS_PER_GROUP);
    const int A = hp.attention_out_units, Dd = hp.decoder_out_units, Ds = hp.decoder_self_attention_out_units;
    const int S = io.max_steps;

    DecParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.Bp = Bp; p.ld = ld; p.R = R; p.T = T;
    p.M = hp.num_mels; p.r = hp.outputs_per_step; p.nf = hp.n_feed_frame; p.Mnf = h->Mnf; p.Mr = h->Mr;
    p.P0 = h->P0; p.P1 = h->P1; p.A = A; p.Dd = Dd; p.Dctx = h->Dctx; p.n_mech = h->n_mech;
    p.hops = h->hops; p.heads = hp.decoder_self_attention_num_heads; p.Ds = Ds; p.ms = hp.multi_speaker_prenet;
    p.feed_mode = io.feed_mode; p.stop_mode = io.stop_mode; p.max_steps = S; p.min_iters = io.min_iters;
    p.retire = io.stop_mode == SAT_STOP_PER_ROW && h->retire_rows && h->hops == 0 && !stepping;  // (the KV-cached hops walk all rows)
    p.finished = finished;
    p.t_begin = stepping ? t_begin : 0;
    p.t_end = stepping ? std::min(S, t_begin + t_count) : S;
    SAT_REQUIRE(!stepping || (t_begin >= 0 && t_begin < S && t_count >= 1), "step %d is outside [0, max_steps)", t_begin);
    SAT_REQUIRE(!stepping || t_begin == 0 || (h->step_next == t_begin && h->last_B == B && h->last_T == T && h->last_S == S),
                "sat_decoder_step: step %d does not continue the handle's run (next step %d)", t_begin, h->step_next);
    if (const char* e = getenv("SAT_RETIRE")) p.retire = p.retire && atoi(e) != 0;
    // the phase counters (sat_decoder_profile / SAT_PROF=1) run in an instantiation of their own, without row retirement
    const bool prof_run = h->prof_on || getenv("SAT_PROF");
    if (prof_run) p.retire = 0;
    const void* kernel_fn = prof_run   ? (const void*)decoder_persistent_kernel<false, true>
                            : p.retire ? (const void*)decoder_persistent_kernel<true>
                                       : (const void*)decoder_persistent_kernel<false>;
    p.S_target = io.target_steps; p.Smax = S;
    p.zc = hp.zoneout_factor_cell; p.zh = hp.zoneout_factor_output;
    p.pre0 = h->pre0.desc(); p.pre1 = h->pre1.desc(); p.outp = h->outp.desc();
    {
      bool fuse = h->has_outp_f && io.feed_mode == SAT_FEED_SELF;
      if (const char* e = getenv("SAT_FUSE_PRENET0")) fuse = fuse && atoi(e) != 0;
      if (fuse) {
        p.outp = h->outp_f.desc();
        p.fused_prenet0 = 1;
      }
    }
    if (p.ms) p.msd = h->msd.desc();
    p.lstm_a = h->lstm_a.desc(); p.lstm_1 = h->lstm_1.desc(); p.lstm_2 = h->lstm_2.desc();

    const int Ka = p.P1 + p.Dctx + A, K1 = A + p.Dctx + Dd, K2 = 2 * Dd;
    // ---- tensor-core mode for the gate GEMMs once the batch makes them a real contraction ----
    bool use_tc = B >= 64;
    if (const char* e = getenv("SAT_TC")) use_tc = atoi(e) != 0;
    if (stepping) use_tc = false;  // the tensor-core form keeps the cell state in shared memory for the length of a launch
    use_tc = use_tc && (A % 4 == 0) && (Dd % 4 == 0);
    const int RB = ceil_div(Bp, TC_M);
    size_t smem_bytes = 0;
    if (use_tc) {
      TcLstmDesc* td[3] = {&p.tc_a, &p.tc_1, &p.tc_2};
      const PackedDense* src[3] = {&h->lstm_a, &h->lstm_1, &h->lstm_2};
      const int Ks[3] = {Ka, K1, K2}, us[3] = {A, Dd, Dd};
      int off = 0;
      for (int i = 0; i < 3; ++i) {
        TcLstmDesc& d = *td[i];
        d.Wp = src[i]->Wp; d.bias = src[i]->bias; d.K = Ks[i]; d.units = us[i];
        d.n_tiles = us[i] / 4;
        d.kchunks = ceil_div(Ks[i], tc::CHUNK_K16);
        d.slots = ceil_div(d.n_tiles * RB, h->grid);
        d.img_off = off;
        off += d.slots * d.kchunks * 2048;
      }
      for (int i = 0; i < 3; ++i) {  // cell state of the CTA's own items, after the images
        td[i]->st_off = off;
        off += td[i]->slots * TC_STATE_BYTES;
      }
      // activation ring: `stages` bulk copies of G operand tiles (16 KB each) in flight.  The SM's TMA unit retires one
      // copy per ~0.28 us whatever its size (scripts/dev/tma_bench.cu), so the ring takes the largest copies that fit:
      // 2 x 64 KB, else 2 x 48 KB, 2 x 32 KB, 2 x 16 KB (SAT_TC_G / SAT_TC_STAGES override)
      const long long room = (long long)h->max_dyn_smem - 1024 - off - 1024;
      int G = 0, stages = 2;
      for (int g : {4, 3, 2, 1})
        if (G == 0 && (long long)stages * g * TC_TILE_BYTES <= room) G = g;
      if (const char* e = getenv("SAT_TC_G")) G = std::max(1, std::min(8, atoi(e)));
      if (const char* e = getenv("SAT_TC_STAGES")) stages = std::max(1, std::min((int)TC_MAX_STAGES, atoi(e)));
      const size_t scratch_bytes = std::max((size_t)SCRATCH_FLOATS * sizeof(float), (size_t)stages * G * TC_TILE_BYTES + 1024);
      p.tc_G = G; p.tc_stages = stages;
      // copy order: copies whose columns were all produced at least two phases before the LSTM go first and are requested
      // behind the grid barrier in front of the phase, at most one per ring stage (SAT_TC_PREFILL=0 switches that off)
      if (G > 0) {
        struct Range { int lo, hi; };
        const Range old_a[1] = {{p.P1, Ka}}, old_1[2] = {{0, A}, {A + p.Dctx, K1}}, old_2[1] = {{Dd, K2}};
        const Range* olds[3] = {old_a, old_1, old_2};
        const int nold[3] = {1, 2, 1};
        int pre_max = stages;
        if (const char* e = getenv("SAT_TC_PREFILL")) pre_max = std::min(stages, std::max(0, atoi(e)));
        for (int i = 0; i < 3; ++i) {
          TcLstmDesc& d = *td[i];
          d.ncopy = ceil_div(d.kchunks, G);
          if (d.ncopy > 16) { G = 0; break; }
          int n = 0, n_old = 0;
          for (int pass = 0; pass < 2; ++pass)
            for (int c = 0; c < d.ncopy; ++c) {
              const int lo = c * G * tc::CHUNK_K16, hi = std::min(lo + G * tc::CHUNK_K16, Ks[i]);
              bool old = false;
              for (int r = 0; r < nold[i]; ++r) old = old || (lo >= olds[i][r].lo && hi <= olds[i][r].hi);
              if (old == (pass == 0)) { d.order[n++] = (unsigned char)c; n_old += pass == 0; }
            }
          d.n_pre = std::min(n_old, pre_max);
        }
      }
      p.scratch_floats = (int)(scratch_bytes / sizeof(float));
      smem_bytes = scratch_bytes + 1024 + (size_t)off;
      if (G == 0 || smem_bytes > h->max_dyn_smem) use_tc = false;  // resident images do not fit: SIMT form
    }
    {
      const int need = frame_smem_floats(p.outp.K, p.outp.Np, p.Mnf, p.pre0.Np, p.pre1.Np);
      SAT_REQUIRE(need <= (int)SCRATCH_FLOATS, "frame phase needs %d floats of shared memory (limit %d): layer sizes too large",
                  need, (int)SCRATCH_FLOATS);
      SAT_REQUIRE(p.outp.Np <= 4 * NT && p.pre0.Np <= 4 * NT && p.pre1.Np <= 4 * NT,
                  "prenet / projection widths must be <= %d", 4 * NT);
    }
    if (const char* e = getenv("SAT_BENCH_BARRIERS")) p.bench_barriers = atoi(e);
    p.tc = use_tc ? 1 : 0;
    if (!use_tc) p.scratch_floats = SCRATCH_FLOATS;
    p.RB = RB;
    p.Ka_pad = ceil_div(Ka, tc::CHUNK_K16) * tc::CHUNK_K16;
    p.K1_pad = ceil_div(K1, tc::CHUNK_K16) * tc::CHUNK_K16;
    p.K2_pad = ceil_div(K2, tc::CHUNK_K16) * tc::CHUNK_K16;
    h->last_tc = p.tc;
    // ---- shared-memory plan (SIMT form): keep as many LSTM tiles resident as fit ----
    // rows per work item: the finest split of {128, 64, 32} whose items still fit one wave of the grid
    auto plan_rows = [&](DenseDesc& d, bool fixed128) {
      d.rpg = ROWS_PER_GROUP;
      if (!fixed128)
        for (int rpg : {64, 32})
          if (d.n_tiles * ceil_div(Bp, rpg) <= h->grid) d.rpg = rpg;
      d.R = ceil_div(Bp, d.rpg);
    };
    DenseDesc* lst[3] = {&p.lstm_a, &p.lstm_1, &p.lstm_2};
    for (DenseDesc* d : lst) plan_rows(*d, true);
    plan_rows(p.pre0, false); plan_rows(p.pre1, false); plan_rows(p.outp, false);
    if (p.ms) plan_rows(p.msd, false);
    auto plan_simt = [&](int grid) {  // resident LSTM tiles of the SIMT form for a launch of `grid` CTAs -> bytes
      size_t fl = SCRATCH_FLOATS;
      for (DenseDesc* d : lst) {
        d->smem_off = -1;
        d->slots = 0;
        const int slots = ceil_div(d->n_tiles * d->R, grid);
        const size_t need = (size_t)slots * d->K * TILE_N;
        if ((fl + need) * sizeof(float) <= h->max_dyn_smem) {
          d->smem_off = (int)(fl - SCRATCH_FLOATS);
          d->slots = slots;
          fl += need;
        }
      }
      return fl * sizeof(float);
    };
    if (!use_tc) smem_bytes = plan_simt(h->grid);
    // ---- attention row plan: per-position arrays sized by this batch's T; two concurrent rows per CTA when they fit ----
    int att_TS = (T + 15) / 16 * 16, att_FB = 8 * att_TS;
    for (int i = 0; i < h->n_mech; ++i) att_FB = std::max(att_FB, NW * hp.memory_dim[i]);
    if (att_smem_floats(att_TS, att_FB) > (int)ATT_SMEM_FLOATS) { att_TS = MAX_T; att_FB = 8 * MAX_T; }
    p.att_row_floats = (att_smem_floats(att_TS, att_FB) + 31) / 32 * 32;
    p.att_pair = (att_TS < MAX_T && 2 * p.att_row_floats <= p.scratch_floats && B * h->n_mech > h->grid) ? 1 : 0;
    if (const char* e = getenv("SAT_ATT_PAIR")) p.att_pair = p.att_pair && atoi(e) != 0;
    // ---- cluster-split attention rows: small batches leave most SMs idle in the row-split phase, so a cluster of 4
    // (or 2) CTAs shares each (row, mechanism) item (DecParams::cl).  fp32 SIMT form only. ----
    int grid_run = h->grid;
    p.cl = 1;
    if (!use_tc) {
      int want = 4;
      if (const char* e = getenv("SAT_CLUSTER")) want = atoi(e);
      for (int cl = want; cl >= 2 && p.cl == 1; cl >>= 1) {
        if (cl != 2 && cl != 4) continue;
        const int ci = cl == 4 ? 1 : 0;
        bool any = false;
        int split[2] = {0, 0};
        for (int i = 0; i < h->n_mech; ++i) {
          int kind = hp.attention_kind[i];
          if (io.forced_alignment) kind = SAT_ATT_TEACHER_FORCING;
          const int Dv = hp.memory_dim[i], win = Dv / cl, F = hp.attention_filters, Up = h->Up[i];
          const bool ctx_ok = Dv % cl == 0 && win % 4 == 0 && ((win / 4) & (win / 4 - 1)) == 0 && win / 4 <= 32 &&
                              io.memory[i] && (((uintptr_t)io.memory[i]) & 15) == 0;
          const bool loc = kind == SAT_ATT_FORWARD || kind == SAT_ATT_LOCATION_SENSITIVE;
          const bool mma_ok = loc && F <= 7 && (Up == 128 || Up == 256) && att_TS < MAX_T && (Up / 32) % cl == 0 &&
                              !h->Wt[i] && h->dq_cl[i][ci];
          split[i] = ctx_ok && (kind == SAT_ATT_TEACHER_FORCING || mma_ok);
          any = any || split[i];
        }
        int maxNp = 0;
        for (int i = 0; i < h->n_mech; ++i) maxNp = std::max(maxNp, h->dq[i].Np);
        const int att_off = (al4(A) + maxNp + 4 * NT + 31) / 32 * 32;
        if (!any || att_off + p.att_row_floats > (int)SCRATCH_FLOATS) continue;
        int grid_c = h->grid / cl * cl;
        for (int pass = 0; pass < 2 && grid_c >= B * h->n_mech * cl; ++pass) {
          const size_t bytes = plan_simt(grid_c);
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(grid_c); cfg.blockDim = dim3(NT); cfg.dynamicSmemBytes = bytes; cfg.stream = st;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension;
          at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          cfg.attrs = at; cfg.numAttrs = 1;
          int n_cl = 0;
          if (cudaOccupancyMaxActiveClusters(&n_cl, kernel_fn, &cfg) != cudaSuccess) {
            cudaGetLastError();
            break;
          }
          if (n_cl * cl >= grid_c) {
            p.cl = cl; grid_run = grid_c; smem_bytes = bytes; p.cl_att_off = att_off;
            for (int i = 0; i < h->n_mech; ++i) {
              p.cl_split[i] = split[i]; p.cl_wq[i] = h->dq_cl[i][ci]; p.cl_sw[i] = h->Up[i] / cl;
            }
            break;
          }
          grid_c = n_cl * cl;  // fewer clusters fit than CTAs / cl: shrink the grid and re-plan once
        }
      }
      if (p.cl == 1) smem_bytes = plan_simt(h->grid);
    }
    if (getenv("SAT_DEBUG_PLAN"))
      fprintf(stderr, "[sat decoder plan] B=%d T=%d tc=%d G=%d stages=%d att_pair=%d att_row_floats=%d scratch_floats=%d smem=%zu cl=%d "
                      "grid=%d\n", B, T, (int)use_tc, p.tc_G, p.tc_stages, p.att_pair, p.att_row_floats, p.scratch_floats, smem_bytes, p.cl,
              grid_run);
    h->last_cl = p.cl;

    // ---- workspace carve-up (floats) ----
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off += (n + 63) / 64 * 64; return o; };
    const size_t o_xin = take((size_t)p.Mnf * ld), o_p0 = take((size_t)p.P0 * ld), o_p0b = take((size_t)p.P0 * ld);
    // fp16 row-major [rows_tc][K*_pad] in tensor-core mode (two halves per float slot), else fp32 [K][ld]
    const size_t rows_tc = (size_t)RB * TC_M;
    const size_t n_att = use_tc ? rows_tc * p.Ka_pad / 2 : (size_t)Ka * ld, n_l1 = use_tc ? rows_tc * p.K1_pad / 2 : (size_t)K1 * ld;
    const size_t n_l2 = use_tc ? rows_tc * p.K2_pad / 2 : (size_t)K2 * ld;
    const size_t o_att0 = take(n_att), o_att1 = take(n_att);
    const size_t o_l10 = take(n_l1), o_l11 = take(n_l1);
    const size_t o_l20 = take(n_l2), o_l21 = take(n_l2);
    const size_t o_qfm = take((size_t)A * ld);
    const size_t o_ca = take((size_t)A * ld), o_c1 = take((size_t)Dd * ld), o_c2 = take((size_t)Dd * ld);
    const size_t o_dout = take((size_t)Dd * ld);
    const size_t o_saq = take((size_t)Ds * ld), o_sactx = take((size_t)Ds * ld), o_saa = take((size_t)Ds * ld);
    const size_t o_say0 = take((size_t)Ds * ld), o_say1 = take((size_t)Ds * ld);
    const size_t o_spk = take((size_t)p.P0 * ld);
    size_t o_pq[2], o_sal[2], o_sap[2], o_su[2], o_keys[2];
    for (int i = 0; i < 2; ++i) {
      o_pq[i] = take((size_t)256 * ld);
      o_sal[i] = take((size_t)B * T);
      o_sap[i] = take((size_t)B * T);
      o_su[i] = take((size_t)B);
    }
    const size_t state_floats = off;  // everything up to here is zeroed per run
    for (int i = 0; i < h->n_mech; ++i) o_keys[i] = take((size_t)B * T * h->Up[i]);
    size_t o_v16[2] = {0, 0};
    if (use_tc)
      for (int i = 0; i < h->n_mech; ++i) o_v16[i] = take(((size_t)B * T * hp.memory_dim[i] + 1) / 2);
    size_t o_kc[MAX_HOPS], o_vc[MAX_HOPS];
    for (int i = 0; i < h->hops; ++i) {
      o_kc[i] = take((size_t)B * Ds * S);
      o_vc[i] = take((size_t)B * Ds * S);
    }
    float* ws = h->ws.get<float>(off);
    const bool fresh = p.t_begin == 0;  // a continued run keeps the state the previous launch left in the workspace
    if (fresh) SAT_CUDA(cudaMemsetAsync(ws, 0, state_floats * sizeof(float), st));
    SAT_CUDA(cudaMemsetAsync(h->barrier, 0, sizeof(unsigned), st));
    if (fresh) SAT_CUDA(cudaMemsetAsync(io.stop_steps, 0xff, sizeof(int32_t) * B, st));

    p.xin = ws + o_xin; p.p0 = ws + o_p0; p.p0b = ws + o_p0b;
    p.att_in[0] = ws + o_att0; p.att_in[1] = ws + o_att1;
    p.l1_in[0] = ws + o_l10; p.l1_in[1] = ws + o_l11;
    p.l2_in[0] = ws + o_l20; p.l2_in[1] = ws + o_l21;
    p.qfm = ws + o_qfm;
    p.c_a = ws + o_ca; p.c_1 = ws + o_c1; p.c_2 = ws + o_c2; p.dout = ws + o_dout;
    p.sa_q = ws + o_saq; p.sa_ctx = ws + o_sactx; p.sa_a = ws + o_saa; p.sa_y[0] = ws + o_say0; p.sa_y[1] = ws + o_say1;
    p.target = io.target; p.lengths = (const long long*)io.memory_lengths;
    p.mel = io.mel; p.stop = io.stop; p.dec_out = io.decoder_outputs; p.stop_steps = io.stop_steps; p.n_steps = io.n_steps;
    p.barrier = h->barrier; p.prof = prof_run ? h->prof : nullptr;

    // tensor-core form of the time-batched dense layers (dense_tc3_launch): worth it from a few hundred rows on
    const size_t rows_pack = std::max((size_t)B * T, (size_t)B * S);
    __half* pack = nullptr;
    if (h->dense_tc) {
      int Kmax = std::max(std::max(p.P0, Dd), Ds);
      for (int i = 0; i < h->n_mech; ++i) Kmax = std::max(Kmax, (int)hp.memory_dim[i]);
      pack = h->ws_pack.get<__half>(dense_tc3_pack_halves(rows_pack, (Kmax + 63) / 64 * 64));
    }
    auto tc3 = [&](const DenseTc3& d, size_t rows) { return pack && d.img && rows >= 256; };
    if (p.ms) {
      spk_term_kernel<<<ceil_div(B * p.P0, 256), 256, 0, st>>>(io.speaker_embed, h->ms_Ws, h->ms_bs, B, hp.speaker_embedding_dim,
                                                              p.P0, ld, ws + o_spk);
      count_launch();
      p.spk_term = ws + o_spk;
    }
    for (int i = 0; i < h->n_mech; ++i) {
      int kind = hp.attention_kind[i];
      if (io.forced_alignment) kind = SAT_ATT_TEACHER_FORCING;
      SAT_REQUIRE(io.memory[i], "memory[%d] is null", i);
      AttMech& m = p.mech[i];
      m.kind = kind; m.U = hp.attention_units[i]; m.Up = h->Up[i]; m.Dv = hp.memory_dim[i];
      m.F = hp.attention_filters; m.Kc = hp.attention_kernel; m.pad_left = (hp.attention_kernel - 1) / 2;
      m.cumulative = hp.cumulative_weights; m.agent = (h->Wt[i] != nullptr);
      m.values = io.memory[i];
      if (use_tc && m.Dv % 8 == 0) {
        // fp16 copy of the attention values: halves the bytes of the context reduction (see attention_context_h8)
        const size_t n = (size_t)B * T * m.Dv;
        __half* v16 = (__half*)(ws + o_v16[i]);
        f32_to_f16_kernel<<<(unsigned)std::min<size_t>(ceil_div64((long long)n, 256), 148 * 8), 256, 0, st>>>(io.memory[i], v16, n);
        count_launch();
        m.values16 = v16;
      }
      m.v = h->att_v[i]; m.ba = h->att_ba[i]; m.conv_w = h->conv_w[i]; m.conv_b = h->conv_b[i]; m.Wl = h->Wl[i];
      m.Wt = h->Wt[i]; m.bt = h->bt[i];
      m.pq = ws + o_pq[i]; m.pq_su = 1; m.pq_sb = m.Up;  // row-major [b][Up]: one 16-byte load per lane
      m.st_align = ws + o_sal[i]; m.st_alpha = ws + o_sap[i]; m.st_u = ws + o_su[i];
      m.out_align = io.alignments[i]; m.out_S = S;
      m.TS = att_TS; m.FB = att_FB;
      p.dq[i] = h->dq[i].desc();
      plan_rows(p.dq[i], false);
      if (kind == SAT_ATT_TEACHER_FORCING) {
        SAT_REQUIRE(io.teacher_alignments[i] && io.target_steps >= 1, "teacher-forcing attention needs teacher_alignments[%d]", i);
        m.teacher = io.teacher_alignments[i];
        m.teacher_S = io.target_steps;
      } else {
        // keys = (memory * mask) @ memory_layer  (TF BahdanauAttention; computed once per batch)
        if (tc3(h->t_mem[i], (size_t)B * T) && h->t_mem[i].Npad == m.Up) {
          dense_tc3_launch(h->t_mem[i], io.memory[i], m.Dv, (size_t)B * T, ACT_NONE, nullptr, 0, ws + o_keys[i], m.Up, pack,
                           st, io.memory_lengths, T);
        } else {
          GemmArgs g;
          g.X = io.memory[i]; g.Bn = B; g.T = T; g.Cin = m.Dv; g.ldx = m.Dv;
          g.W = h->mem_layer[i]; g.N = m.Up; g.C = ws + o_keys[i]; g.ldc = m.Up;
          g.row_lengths = io.memory_lengths;
          gemm_launch(g, st);
        }
        m.keys = ws + o_keys[i];
      }
    }
    if (fresh) {
      init_state_kernel<<<ceil_div(B, 128), 128, 0, st>>>(ws + o_sap[0], ws + o_sap[1], ws + o_su[0], ws + o_su[1], B, T);
      count_launch();
    }
    for (int i = 0; i < h->hops; ++i) {
      p.hop[i].qkv = h->sa_qkv[i].desc(); p.hop[i].o = h->sa_o[i].desc(); p.hop[i].tr = h->sa_tr[i].desc();
      plan_rows(p.hop[i].qkv, false); plan_rows(p.hop[i].o, false); plan_rows(p.hop[i].tr, false);
      p.hop[i].kcache = ws + o_kc[i]; p.hop[i].vcache = ws + o_vc[i];
      p.hop[i].probs = io.self_attention_probs[i];
    }

    // ---- teacher-forced time-batched form: prenet of all steps before the loop, hops + projections after it ----
    bool tail = io.feed_mode == SAT_FEED_TEACHER && io.stop_mode == SAT_STOP_FIXED && !stepping;
    if (const char* e = getenv("SAT_BATCHED_TAIL")) tail = tail && atoi(e) != 0;
    auto dense_rows = [&](const PackedDense& d, const DenseTc3& t, const float* x, int ldx, size_t rows, int act, float* y,
                          int ldy) {
      if (tc3(t, rows) && ldy >= t.Npad) {
        dense_tc3_launch(t, x, ldx, rows, act, nullptr, 0, y, ldy, pack, st);
        return;
      }
      SAT_REQUIRE(ldy == d.Np, "dense_rows: output stride");
      GemmArgs g;
      g.X = x; g.Bn = 1; g.T = (int)rows; g.Cin = d.K; g.ldx = ldx;
      g.W = d.Wrm; g.N = d.Np; g.shift = d.bias; g.act = act; g.C = y; g.ldc = d.Np;
      gemm_launch(g, st);
    };
    const size_t rows_all = (size_t)B * S;
    float *t_dec = nullptr, *t_y[2] = {nullptr, nullptr}, *t_mha = nullptr, *t_proj = nullptr;
    int ld_proj = h->outp.Np;
    if (tail) {
      for (int i = 0; i < h->hops; ++i)
        tail = tail && h->sa_qkv[i].Np == h->sa_qkv[i].N && h->sa_o[i].Np == h->sa_o[i].N && h->sa_tr[i].Np == h->sa_tr[i].N;
    }
    if (tail) {
      size_t toff = 0;
      auto ttake = [&](size_t n) { size_t o = toff; toff += (n + 63) / 64 * 64; return o; };
      const size_t o_x = ttake(rows_all * p.Mnf), o_a = ttake(rows_all * h->pre0.Np), o_b = ttake(rows_all * h->pre0.Np);
      const size_t o_p1 = ttake(rows_all * h->pre1.Np), o_dec = ttake(rows_all * Dd);
      const size_t o_y0 = ttake(rows_all * Ds), o_y1 = ttake(rows_all * Ds);
      MhaDev m0;
      m0.D = Ds;
      const size_t o_mha = ttake(h->hops ? mha_scratch_floats(m0, (int64_t)rows_all) : 0);
      ld_proj = tc3(h->t_outp, rows_all) ? h->t_outp.Npad : h->outp.Np;
      const size_t o_pr = ttake(rows_all * ld_proj);
      float* tw = h->ws_tail.get<float>(toff);
      teacher_inputs_kernel<<<(unsigned)std::min<size_t>((rows_all * p.Mnf + 255) / 256, 148 * 16), 256, 0, st>>>(
          io.target, B, S, io.target_steps, p.Mr, p.Mnf, tw + o_x);
      count_launch();
      dense_rows(h->pre0, DenseTc3(), tw + o_x, p.Mnf, rows_all, ACT_RELU, tw + o_a, h->pre0.Np);
      const float* pin = tw + o_a;
      if (p.ms) {
        add_spk_rows_kernel<<<(unsigned)std::min<size_t>((rows_all * p.P0 + 255) / 256, 148 * 16), 256, 0, st>>>(
            tw + o_a, (int)rows_all, B, p.P0, h->pre0.Np, p.spk_term, ld);
        count_launch();
        dense_rows(h->msd, h->t_msd, tw + o_a, h->pre0.Np, rows_all, ACT_RELU, tw + o_b, h->pre0.Np);
        pin = tw + o_b;
      }
      dense_rows(h->pre1, h->t_pre1, pin, h->pre0.Np, rows_all, ACT_RELU, tw + o_p1, h->pre1.Np);
      p.batched_tail = 1;
      p.p1_all = tw + o_p1;
      p.p1_ld = h->pre1.Np;
      t_dec = io.decoder_outputs ? io.decoder_outputs : tw + o_dec;
      p.dec_out = t_dec;
      t_y[0] = tw + o_y0; t_y[1] = tw + o_y1; t_mha = tw + o_mha; t_proj = tw + o_pr;
    }

    void* args[] = {&p};
    SAT_CUDA(cudaEventRecord(h->ev0, st));
    {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(grid_run); cfg.blockDim = dim3(NT); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = st;
      cudaLaunchAttribute at[2];
      at[0].id = cudaLaunchAttributeCooperative;
      at[0].val.cooperative = 1;
      at[1].id = cudaLaunchAttributeClusterDimension;
      at[1].val.clusterDim.x = p.cl; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = p.cl > 1 ? 2 : 1;
      // the specialised instantiation of the plain kernel (free-running baseline decoder, B >= 64; decoder_kernel.cuh)
      const void* launch_fn = kernel_fn;
      if (!prof_run && !p.retire && p.tc && p.cl == 1 && p.hops == 0 && p.n_mech == 1 && !p.batched_tail && p.fused_prenet0 && !p.ms &&
          !getenv("SAT_NO_SPEC"))
        launch_fn = (const void*)decoder_persistent_kernel<false, false, true>;
      SAT_CUDA(cudaLaunchKernelExC(&cfg, launch_fn, args));
    }
    SAT_CUDA(cudaEventRecord(h->ev1, st));
    h->last_B = B; h->last_T = T; h->last_S = S;
    h->step_next = stepping ? p.t_end : -1;
    count_launch();
    if (tail) {
      // decoder self-attention hops as causal time-batched attention, then the frame + stop projection
      const float* y = t_dec;
      for (int i = 0; i < h->hops; ++i) {
        MhaDev m;
        m.Wqkv = h->sa_qkv[i].Wrm; m.bqkv = h->sa_qkv[i].bias; m.Wo = h->sa_o[i].Wrm; m.bo = h->sa_o[i].bias;
        m.Wt = h->sa_tr[i].Wrm; m.bt = h->sa_tr[i].bias;
        m.Din = Dd; m.D = Ds; m.H = hp.decoder_self_attention_num_heads;
        if (pack && rows_all >= 256) {
          m.tq = h->t_qkv[i]; m.to = h->t_o[i]; m.tt = h->t_tr[i];
          m.xpack = pack;
        }
        sat_layer_forward(m, y, B, S, 1, nullptr, t_y[i & 1], io.self_attention_probs[i], t_mha, st);
        y = t_y[i & 1];
      }
      dense_rows(h->outp, h->t_outp, y, h->outp.K, rows_all, ACT_NONE, t_proj, ld_proj);
      split_projection_kernel<<<(unsigned)std::min<size_t>((rows_all * (p.Mr + 1) + 255) / 256, 148 * 16), 256, 0, st>>>(
          t_proj, rows_all, ld_proj, p.Mr, io.mel, io.stop);
      count_launch();
    }
    if (io.sample_ids) {
      const long long rows = (long long)B * S * p.r;
      argmax_kernel<<<(unsigned)ceil_div64(rows, 8), 256, 0, st>>>(io.mel, rows, p.M, io.sample_ids);
      count_launch();
    }
    check_launch("decoder_persistent_kernel");
  });
}

extern "C" int sat_decoder_run(sat_decoder* h, const sat_decoder_io* iop, void* stream) {
  return decoder_run_impl(h, iop, stream, 0, -1);
}

extern "C" int sat_decoder_step(sat_decoder* h, const sat_decoder_io* iop, int32_t step, int32_t* finished, void* stream) {
  return decoder_run_impl(h, iop, stream, step, 1, finished);
}

// cycles per phase accumulated by CTA 0 over all runs since the last reset (debug / profiling aid)
extern "C" int sat_decoder_profile(sat_decoder* h, int64_t* out16 /* N_PROF = 48 slots */, int32_t reset) {
  return guarded([&] {
    SAT_REQUIRE(h && out16, "null argument");
    h->prof_on = true;  // the counters run from the next sat_decoder_run on
    SAT_CUDA(cudaDeviceSynchronize());
    SAT_CUDA(cudaMemcpy(out16, h->prof, sizeof(long long) * N_PROF, cudaMemcpyDeviceToHost));
    if (reset) SAT_CUDA(cudaMemset(h->prof, 0, sizeof(long long) * N_PROF));
  });
}

// device time of the persistent kernel of the last sat_decoder_run (CUDA events on the launching stream); blocks
// until that kernel has finished
extern "C" int sat_decoder_used_tensor_cores(sat_decoder* h, int32_t* used) {
  return guarded([&] {
    SAT_REQUIRE(h && used, "null argument");
    *used = h->last_tc;
  });
}

extern "C" int sat_decoder_set_row_retirement(sat_decoder* h, int32_t on) {
  return guarded([&] {
    SAT_REQUIRE(h, "null argument");
    h->retire_rows = on != 0;
  });
}

extern "C" int sat_decoder_cluster_size(sat_decoder* h, int32_t* size) {
  return guarded([&] {
    SAT_REQUIRE(h && size, "null argument");
    *size = h->last_cl;
  });
}

extern "C" int sat_decoder_kernel_ms(sat_decoder* h, float* ms) {
  return guarded([&] {
    SAT_REQUIRE(h && ms, "null argument");
    SAT_CUDA(cudaEventSynchronize(h->ev1));
    SAT_CUDA(cudaEventElapsedTime(ms, h->ev0, h->ev1));
  });
}
