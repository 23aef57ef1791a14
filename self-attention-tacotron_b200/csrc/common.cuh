// common.cuh -- host-side plumbing shared by every translation unit of libsat.so.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cstdarg>
#include <string>
#include <vector>
#include <map>
#include <stdexcept>
#include <memory>
#include <algorithm>
#include "../../include/sat.h"

namespace sat {

// ---- error plumbing: C++ exceptions inside, int codes at the ABI ----------
struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
void set_last_error(const std::string& m);
[[noreturn]] void fail(int code, const char* fmt, ...);

#define SAT_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess)                                                              \
      ::sat::fail(SAT_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

#define SAT_REQUIRE(cond, ...)                        \
  do {                                                \
    if (!(cond)) ::sat::fail(SAT_EINVAL, __VA_ARGS__); \
  } while (0)

// Wraps an ABI body: catches everything, records the message, returns a code.
template <class F>
int guarded(F&& f) {
  try {
    f();
    return SAT_OK;
  } catch (const Error& e) {
    set_last_error(e.what());
    return e.code;
  } catch (const std::exception& e) {
    set_last_error(e.what());
    return SAT_EINVAL;
  }
}

// every kernel launch of the library goes through this counter (bench: gpu_launches)
void count_launch(int n = 1);
void check_launch(const char* what);

// ---- weight blob lookup -----------------------------------------------------
struct Blob {
  std::map<std::string, std::pair<const float*, int64_t>> t;
  Blob(const sat_tensor* w, int n) {
    for (int i = 0; i < n; ++i) t[w[i].name] = {w[i].data, w[i].numel};
  }
  bool has(const std::string& name) const { return t.count(name) != 0; }
  const float* get(const std::string& name, int64_t numel) const {
    auto it = t.find(name);
    if (it == t.end()) fail(SAT_EMISSING, "weight '%s' missing from the blob", name.c_str());
    if (it->second.second != numel)
      fail(SAT_EINVAL, "weight '%s' has %lld elements, expected %lld", name.c_str(),
           (long long)it->second.second, (long long)numel);
    return it->second.first;
  }
};

// ---- device memory owned by a handle ---------------------------------------
struct DevPool {
  std::vector<void*> ptrs;
  size_t bytes = 0;
  ~DevPool() { release(); }
  void release() {
    for (void* p : ptrs) cudaFree(p);
    ptrs.clear();
    bytes = 0;
  }
  template <class T>
  T* alloc(size_t n, bool zero = true) {
    void* p = nullptr;
    size_t b = (n ? n : 1) * sizeof(T);
    SAT_CUDA(cudaMalloc(&p, b));
    if (zero) SAT_CUDA(cudaMemset(p, 0, b));
    ptrs.push_back(p);
    bytes += b;
    return (T*)p;
  }
  float* upload(const float* host, size_t n) {
    float* d = alloc<float>(n, false);
    SAT_CUDA(cudaMemcpy(d, host, n * sizeof(float), cudaMemcpyHostToDevice));
    return d;
  }
  float* upload(const std::vector<float>& v) { return upload(v.data(), v.size()); }
};

// grow-only scratch buffer (re-used across calls of a handle)
struct Scratch {
  void* p = nullptr;
  size_t cap = 0;
  ~Scratch() {
    if (p) cudaFree(p);
  }
  template <class T>
  T* get(size_t n) {
    size_t b = (n ? n : 1) * sizeof(T);
    if (b > cap) {
      if (p) SAT_CUDA(cudaFree(p));
      p = nullptr;
      SAT_CUDA(cudaMalloc(&p, b));
      cap = b;
    }
    return (T*)p;
  }
};

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

enum Act { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2, ACT_SIGMOID = 3, ACT_SOFTSIGN = 4 };

// ---- row-major GEMM / implicit conv1d (gemm.cu) -----------------------------
// C[(b,t), n] = act( (sum_tap sum_c X[b, t + tap - pad_left, c] * W[tap][c][n]) * scale[n] + shift[n] ) (+ R[(b,t), n])
// X is [Bn, T, ldx] with Cin <= ldx used columns starting at x_col0; zero outside [0,T).
// Output row stride ldc, column offset c_col0.  scale may be null (=1), shift may be null (=0).
struct GemmArgs {
  const float* X = nullptr;
  int Bn = 1, T = 1, Cin = 0, ldx = 0;
  const float* W = nullptr;  // device [taps][Cin][N]
  int taps = 1, pad_left = 0, N = 0;
  const float* scale = nullptr;
  const float* shift = nullptr;
  int act = ACT_NONE;
  const float* R = nullptr;  // residual, row stride ldr
  int ldr = 0;
  float* C = nullptr;
  int ldc = 0;
  const int64_t* row_lengths = nullptr;  // optional [Bn]: rows with t >= length read as zero input (values mask)
  const int64_t* gather_ids = nullptr;   // optional [Bn*T] int64: X row index = ids - gather_offset (embedding lookup); T-shift disabled
  int64_t gather_offset = 0;
  int64_t gather_rows = 0;               // rows of the gathered table: an id outside [offset, offset + rows) reads as a zero
  int* err_flag = nullptr;               // row and sets *err_flag = 1 (device int, sticky) instead of reading out of bounds
  int maxpool2 = 0;                      // 1: input element = max(X[t], X[t+1]) (MaxPooling1D k2 s1 SAME) fused on load
  int highway = 0;                       // 1: HighwayNet epilogue over interleaved (H_u, T_u) column pairs; R = layer input; writes N/2 columns
};
void gemm_launch(const GemmArgs& a, cudaStream_t st);

// ---- tensor-core implicit-GEMM conv1d (conv_tc.cu) -----------------------------
constexpr int CONV_TC_HALO = 8;  // zero rows on each side of every utterance in the haloed fp16 layout (taps <= 16)
struct ConvTcLaunch {
  const __half* xp = nullptr;  // haloed fp16 input [B][T + 2*HALO][ldxp]
  int ldxp = 0, B = 0, T = 0, cin = 0, cout = 0, taps = 1, pad_left = 0;
  const __half* wimg = nullptr;  // conv_tc_weight_image
  const float* scale = nullptr;
  const float* shift = nullptr;
  int act = ACT_NONE;
  __half* y16 = nullptr;  // haloed fp16 output (pointer already offset to the first output channel), row stride ld16
  int ld16 = 0;
  float* y32 = nullptr;   // dense fp32 output [B*T][ld32]
  int ld32 = 0;
  const float* res = nullptr;  // optional fp32 residual [B*T][ldr] (joins before the stores; needs y32)
  int ldr = 0;
  int highway = 0;  // HighwayNet epilogue over interleaved (H_u, T_u) columns: x = res, bias = shift, writes cout/2 columns
  // conv-bank form: ONE launch computes the convolutions of widths 1 .. bank (128 channels each, 'SAME' padding) of the
  // same input.  wimg = conv_tc_bank_image: [bank * 128][bank * cin], tap slot s of every width = input offset
  // s - (bank - 1) / 2; column tile w computes width w + 1 over its own tap slots only.  cout = bank * 128;
  // scale / shift / y16 cover all widths (channel offset 128 * w); taps / pad_left are ignored.
  int bank = 0;
  // optional [B] int64 utterance lengths: output rows t >= length are written as ZEROS (both outputs), so that a
  // following multi-tap convolution sees what it would see for the un-padded utterance ('SAME' zero padding)
  const int64_t* lengths = nullptr;
};
__half* conv_tc_bank_image(DevPool& pool, const std::vector<const float*>& w, int cin);  // w[k-1] = [k][cin][128]
bool conv_tc_supported(int cin, int cout);
__half* conv_tc_weight_image(DevPool& pool, const float* w, int taps, int cin, int n);
void conv_tc_pack(const float* x, int ldx, int B, int T, int C, __half* xp, int ldp, cudaStream_t st);
void conv_tc_maxpool2(const __half* x, int B, int T, int C, __half* y, cudaStream_t st);
void conv_tc_zero_halo(__half* x, int B, int T, int C, cudaStream_t st);  // C % 8 == 0
void conv_tc_launch(const ConvTcLaunch& L, cudaStream_t st);

// A time-batched dense layer y[rows, N] = act(x[rows, K] W[K, N] + bias) (+ residual) on the tensor cores with fp32-grade
// accuracy from fp16 operands: x = xh + xl and W = Wh + Wl as fp16 pairs, y = xh Wh + xl Wh + xh Wl as ONE conv_tc_kernel
// launch (taps = 1) over the packed activations [xh | xl | xh] (3K columns) and the image [Wh | Wh | Wl] -- the dropped
// xl Wl term is ~2^-22 relative.  K % 64 == 0; N is padded to a multiple of 128 (Y needs ldy >= Npad columns).
struct DenseTc3 {
  const __half* img = nullptr;  // [Npad][3K] K-major
  const float* bias = nullptr;  // [Npad] (zeros where absent)
  int K = 0, N = 0, Npad = 0;
};
DenseTc3 dense_tc3_upload(DevPool& pool, const float* w, int ldw, const float* bias, int K, int N);  // w: host [K][ldw]
size_t dense_tc3_pack_halves(size_t rows, int K);  // halves of the packed-activation scratch for `rows` rows
// lengths / T (optional): rows (b, t) with t >= lengths[b] are packed as zeros (masked memory of the attention keys)
void dense_tc3_launch(const DenseTc3& d, const float* x, int ldx, size_t rows, int act, const float* res, int ldr, float* y,
                      int ldy, __half* pack, cudaStream_t st, const int64_t* lengths = nullptr, int T = 0);

// ---- multi-head self-attention (mha.cu) --------------------------------------
// Device copy of one MultiHeadAttention (modules/self_attention.py:89-128) and, optionally, the
// tanh transform layer of the enclosing SelfAttentionTransformer (modules/module.py:358).
struct MhaDev {
  DenseTc3 tq, to, tt;    // optional tensor-core forms of the three dense layers (dense_tc3_upload) ...
  __half* xpack = nullptr;  // ... and their packed-activation scratch (dense_tc3_pack_halves(rows, max(Din, D)))
  float* Wqkv = nullptr;  // [Din][3D], column blocks key | value | query
  float* bqkv = nullptr;  // [3D]
  float* Wo = nullptr;    // [D][D]
  float* bo = nullptr;
  float* Wt = nullptr;    // [D][D] transform (null when absent)
  float* bt = nullptr;
  int Din = 0, D = 0, H = 1;
};
// mha_prefix: ".../mha" holding {key,value,query,output}/{kernel,bias}; transform_prefix may be empty
MhaDev mha_upload(DevPool& pool, const Blob& W, const std::string& mha_prefix, const std::string& transform_prefix,
                  int Din, int D, int H);
size_t mha_scratch_floats(const MhaDev& m, int64_t rows);
// out [B,T,D] = output Dense(attention(x)); probs [B,H,T,T] or null; pad_lengths != null enables the padding mask
void mha_forward(const MhaDev& m, const float* x, int B, int T, int causal, const int64_t* pad_lengths,
                 float* out, float* probs, float* scratch, cudaStream_t st, int head_major = 0);
// SelfAttentionTransformer: out = x + tanh(Dense(mha(x)))   (needs Din == D and a transform layer)
void sat_layer_forward(const MhaDev& m, const float* x, int B, int T, int causal, const int64_t* pad_lengths,
                       float* out, float* probs, float* scratch, cudaStream_t st, int head_major = 0);

}  // namespace sat
