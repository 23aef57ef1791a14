// postnet.cu -- sat_postnet_*: tacotron2 PostNetV2 (imported models/models.py:4; hparams hparams.py:146-150;
// consumer predict_mel.py:50,60).  Its definition is absent from /root/reference; the form here is
// n x [Conv1d(k, C, no bias) -> BN -> tanh (last layer: none)] -> Dense(num_mels) -> + input.
// The convolutions run as tcgen05 implicit GEMMs (conv_tc.cu: fp16 operands, fp32 accumulation) when the channel
// count is a multiple of 128 -- the mel input is zero-padded to 128 channels -- and as fp32 SIMT GEMMs otherwise
// (or with SAT_POSTNET_TC=0); the final Dense + residual stays fp32.
#include "common.cuh"
#include <cstdlib>

namespace sat {
void conv_bn_act_launch(const float* x, int B, int T, int ldx, const float* w, const float* scale, const float* shift,
                        int k, int cin, int cout, int act, int maxpool2, const float* R, int ldr, float* y, int ldy,
                        cudaStream_t st, const int64_t* row_lengths = nullptr);
}
using namespace sat;

struct sat_postnet {
  sat_hparams hp;
  DevPool pool;
  Scratch ws;
  std::vector<float*> w, scale, shift;
  std::vector<__half*> wimg;  // tensor-core weight images (first layer: input channels padded to min_cin)
  float *pw = nullptr, *pb = nullptr;
  bool tc = false;
  int cin0 = 0;  // padded input channels of the first layer on the tensor-core path
  Scratch ws16;
};

extern "C" int sat_postnet_create(const sat_hparams* hpp, const sat_tensor* weights, int32_t n_weights, sat_postnet** out) {
  return guarded([&] {
    SAT_REQUIRE(hpp && weights && out, "null argument");
    const sat_hparams& hp = *hpp;
    SAT_REQUIRE(hp.num_postnet_layers >= 1, "post-net needs at least one layer");
    Blob W(weights, n_weights);
    auto* h = new sat_postnet();
    std::unique_ptr<sat_postnet> guard(h);
    h->hp = hp;
    const int M = hp.num_mels, C = hp.postnet_channels, k = hp.postnet_kernel_size;
    h->cin0 = (M + 63) / 64 * 64;
    h->tc = conv_tc_supported(h->cin0, C) && conv_tc_supported(C, C) && M % 4 == 0 && k <= 2 * CONV_TC_HALO;
    if (const char* e = getenv("SAT_POSTNET_TC")) h->tc = h->tc && atoi(e) != 0;
    for (int i = 0; i < hp.num_postnet_layers; ++i) {
      const std::string P = "postnet/conv_" + std::to_string(i);
      const int cin = i == 0 ? M : C;
      h->w.push_back(h->pool.upload(W.get(P + "/kernel", (int64_t)k * cin * C), (size_t)k * cin * C));
      if (h->tc) {
        const float* w = W.get(P + "/kernel", (int64_t)k * cin * C);
        const int cp = i == 0 ? h->cin0 : C;
        std::vector<float> wp((size_t)k * cp * C, 0.f);  // zero rows for the padded input channels
        for (int t = 0; t < k; ++t)
          for (int c = 0; c < cin; ++c) memcpy(&wp[((size_t)t * cp + c) * C], &w[((size_t)t * cin + c) * C], sizeof(float) * C);
        h->wimg.push_back(conv_tc_weight_image(h->pool, wp.data(), k, cp, C));
      }
      const float* gamma = W.get(P + "/bn/gamma", C);
      const float* beta = W.get(P + "/bn/beta", C);
      const float* mean = W.get(P + "/bn/mean", C);
      const float* var = W.get(P + "/bn/var", C);
      std::vector<float> sc(C), sh(C);
      for (int c = 0; c < C; ++c) {
        sc[c] = gamma[c] / sqrtf(var[c] + 1e-3f);
        sh[c] = beta[c] - mean[c] * sc[c];
      }
      h->scale.push_back(h->pool.upload(sc));
      h->shift.push_back(h->pool.upload(sh));
    }
    h->pw = h->pool.upload(W.get("postnet/projection/kernel", (int64_t)C * M), (size_t)C * M);
    h->pb = h->pool.upload(W.get("postnet/projection/bias", M), M);
    guard.release();
    *out = h;
  });
}

extern "C" void sat_postnet_destroy(sat_postnet* h) { delete h; }

extern "C" int sat_postnet_uses_tensor_cores(sat_postnet* h, int32_t* used) {
  return guarded([&] {
    SAT_REQUIRE(h && used, "null argument");
    *used = h->tc ? 1 : 0;
  });
}

extern "C" int sat_postnet_forward(sat_postnet* h, const float* mel, int32_t B, int32_t T, float* mel_postnet, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(h && mel && mel_postnet && B >= 1 && T >= 1, "bad argument");
    const sat_hparams& hp = h->hp;
    cudaStream_t st = (cudaStream_t)stream;
    const int M = hp.num_mels, C = hp.postnet_channels, k = hp.postnet_kernel_size, n = hp.num_postnet_layers;
    const size_t rows = (size_t)B * T;
    float* ws = h->ws.get<float>(2 * rows * C);
    float* buf[2] = {ws, ws + rows * C};
    const float* x = mel;
    int ldx = M, cin = M;
    if (h->tc) {
      const size_t R = (size_t)B * (T + 2 * CONV_TC_HALO);
      const size_t n_in = R * h->cin0, n_act = R * C;
      __half* in16 = h->ws16.get<__half>(n_in + 2 * n_act + 64);
      __half* act16[2] = {in16 + (n_in + 7) / 8 * 8, in16 + (n_in + 7) / 8 * 8 + n_act};
      SAT_CUDA(cudaMemsetAsync(in16, 0, (n_in + 2 * n_act + 64) * sizeof(__half), st));
      conv_tc_pack(mel, M, B, T, M, in16, h->cin0, st);
      const __half* x16 = in16;
      int ld16 = h->cin0;
      for (int i = 0; i < n; ++i) {
        ConvTcLaunch L;
        L.xp = x16; L.ldxp = ld16; L.B = B; L.T = T; L.cin = ld16; L.cout = C; L.taps = k; L.pad_left = (k - 1) / 2;
        L.wimg = h->wimg[i]; L.scale = h->scale[i]; L.shift = h->shift[i];
        if (i < n - 1) {
          L.act = ACT_TANH; L.y16 = act16[i & 1]; L.ld16 = C;
        } else {
          L.act = ACT_NONE; L.y32 = buf[0]; L.ld32 = C;  // the projection below reads fp32
        }
        conv_tc_launch(L, st);
        x16 = act16[i & 1];
        ld16 = C;
      }
      x = buf[0];
      ldx = cin = C;
    }
    for (int i = 0; i < n && !h->tc; ++i) {
      conv_bn_act_launch(x, B, T, ldx, h->w[i], h->scale[i], h->shift[i], k, cin, C, i < n - 1 ? ACT_TANH : ACT_NONE, 0,
                         nullptr, 0, buf[i & 1], C, st);
      x = buf[i & 1];
      ldx = cin = C;
    }
    GemmArgs g;
    g.X = x; g.Bn = 1; g.T = (int)rows; g.Cin = C; g.ldx = C;
    g.W = h->pw; g.N = M; g.shift = h->pb; g.R = mel; g.ldr = M; g.C = mel_postnet; g.ldc = M;
    gemm_launch(g, st);
  });
}
