// abi.cu -- error plumbing, launch accounting and the stand-alone module entry points.
#include "common.cuh"
#include "decoder_attn.cuh"

namespace sat {

static thread_local std::string g_last_error;
static thread_local int64_t g_launches = 0;

void set_last_error(const std::string& m) { g_last_error = m; }

void fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw Error(code, buf);
}

void count_launch(int n) { g_launches += n; }

void check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) fail(SAT_ECUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
}

namespace {
using namespace dec;

__global__ void __launch_bounds__(NT) attention_step_kernel(AttMech m, int B, int T, const long long* lengths) {
  extern __shared__ __align__(16) float sm[];
  for (int b = blockIdx.x; b < B; b += gridDim.x) attention_item<NT>(m, b, 0, T, (int)lengths[b], sm);
}

__global__ void embedding_kernel(const float* table, long long rows, int dim, long long offset, const long long* ids,
                                 long long n, float* out, int* bad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * dim) return;
  const long long r = ids[i / dim] - offset;
  if (r < 0 || r >= rows) { *bad = 1; out[i] = 0.f; return; }
  out[i] = table[r * dim + i % dim];
}

__global__ void mask_rows_kernel(const float* x, const long long* lengths, int B, int T, int D, float* y) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * T * D) return;
  const long long row = i / D;
  const int b = (int)(row / T), t = (int)(row % T);
  y[i] = t < lengths[b] ? x[i] : 0.f;
}
}  // namespace
}  // namespace sat

using namespace sat;

extern "C" int sat_abi_version(void) { return SAT_ABI_VERSION; }
extern "C" const char* sat_last_error(void) { return g_last_error.c_str(); }
extern "C" int64_t sat_launch_count(void) { return g_launches; }
extern "C" void sat_launch_count_reset(void) { g_launches = 0; }

// Host-side helper of the sharded path (sharding.make_buckets): cuts of one rank's share, sorted longest first, into
// batches of at most batch_size utterances minimising sum_batches longest * (c0 + c1 * size) -- an O(n * batch_size)
// dynamic programme from the back (ties: the smaller batch), written here because it sits in front of the first launch
// of every predict() call.
extern "C" int sat_bucket_cuts(const int64_t* sorted_lengths, int64_t n, int32_t batch_size, double c0, double c1,
                               int32_t* cut) {
  return guarded([&] {
    SAT_REQUIRE(n >= 0 && batch_size > 0 && (n == 0 || (sorted_lengths && cut)), "bad argument");
    std::vector<double> best((size_t)n + 1, 0.0);
    for (int64_t i = n - 1; i >= 0; --i) {
      const double L = (double)sorted_lengths[i];
      const int64_t kmax = std::min<int64_t>(batch_size, n - i);
      double bc = 0.0;
      int32_t bk = 0;
      for (int64_t k = 1; k <= kmax; ++k) {
        const double c = L * (c0 + c1 * (double)k) + best[i + k];
        if (bk == 0 || c < bc) bc = c, bk = (int32_t)k;
      }
      best[i] = bc;
      cut[i] = bk;
    }
  });
}

extern "C" int sat_dense_forward(const float* kernel_host, const float* bias_host, int32_t K, int32_t N, int32_t act,
                                 const float* x, int64_t rows, float* y, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(kernel_host && x && y && K > 0 && N > 0, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    DevPool pool;
    GemmArgs g;
    g.X = x; g.Bn = 1; g.T = (int)rows; g.Cin = K; g.ldx = K;
    g.W = pool.upload(kernel_host, (size_t)K * N); g.N = N;
    g.shift = bias_host ? pool.upload(bias_host, N) : nullptr;
    g.act = act; g.C = y; g.ldc = N;
    // rows are independent: run as one "utterance" of `rows` frames with a single tap
    gemm_launch(g, st);
    SAT_CUDA(cudaStreamSynchronize(st));  // the temporary weight copies die with `pool`
  });
}

extern "C" int sat_attention_prepare(const sat_tensor* weights, int32_t n_weights, const char* prefix,
                                     const float* memory, const int64_t* lengths, int32_t B, int32_t T, int32_t Dm,
                                     int32_t U, float* keys, float* values, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(weights && prefix && memory && lengths && keys, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    Blob W(weights, n_weights);
    DevPool pool;
    GemmArgs g;
    g.X = memory; g.Bn = B; g.T = T; g.Cin = Dm; g.ldx = Dm;
    g.W = pool.upload(W.get(std::string(prefix) + "/memory_layer/kernel", (int64_t)Dm * U), (size_t)Dm * U);
    g.N = U; g.C = keys; g.ldc = U; g.row_lengths = lengths;
    gemm_launch(g, st);
    if (values) {
      const long long n = (long long)B * T * Dm;
      mask_rows_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, st>>>(memory, (const long long*)lengths, B, T, Dm, values);
      count_launch();
    }
    check_launch("attention_prepare");
    SAT_CUDA(cudaStreamSynchronize(st));
  });
}

extern "C" int sat_attention_step(const sat_tensor* weights, int32_t n_weights, const char* prefix, int32_t kind,
                                  int32_t B, int32_t T, int32_t U, int32_t Dv, int32_t A,
                                  int32_t attention_kernel, int32_t attention_filters,
                                  int32_t cumulative_weights, int32_t use_transition_agent,
                                  const float* keys, const float* values, const int64_t* lengths, const float* query,
                                  float* state_alignments, float* state_alpha, float* state_u,
                                  float* alignments, float* context, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(weights && prefix && keys && values && lengths && query && alignments && context, "null argument");
    SAT_REQUIRE(kind == SAT_ATT_ADDITIVE || kind == SAT_ATT_LOCATION_SENSITIVE || kind == SAT_ATT_FORWARD,
                "sat_attention_step handles additive / location_sensitive / forward");
    SAT_REQUIRE(U >= 1 && U <= 256 && Dv <= dec::NT && T <= dec::MAX_T, "unsupported shape (U<=256, Dv<=512, T<=1024)");
    const int Up = dec::att_padded_units(U);
    cudaStream_t st = (cudaStream_t)stream;
    Blob W(weights, n_weights);
    const std::string P = std::string(prefix) + "/";
    DevPool pool;
    // processed query = query @ query_layer (row-major [B][U])
    float* pq = pool.alloc<float>((size_t)B * U);
    GemmArgs g;
    g.X = query; g.Bn = 1; g.T = B; g.Cin = A; g.ldx = A;
    g.W = pool.upload(W.get(P + "query_layer/kernel", (int64_t)A * U), (size_t)A * U);
    g.N = U; g.C = pq; g.ldc = U;
    gemm_launch(g, st);
    dec::AttMech m;
    memset(&m, 0, sizeof(m));
    m.kind = kind; m.U = U; m.Up = Up; m.Dv = Dv; m.F = attention_filters; m.Kc = attention_kernel;
    m.pad_left = (attention_kernel - 1) / 2; m.cumulative = cumulative_weights;
    // keys and per-unit vectors padded to the kernel's unit axis (zeros in the padding)
    float* keys_p = pool.alloc<float>((size_t)B * T * Up);
    SAT_CUDA(cudaMemcpy2DAsync(keys_p, (size_t)Up * 4, keys, (size_t)U * 4, (size_t)U * 4, (size_t)B * T,
                               cudaMemcpyDeviceToDevice, st));
    auto padU = [&](const float* src, int rows) {
      std::vector<float> v((size_t)rows * Up, 0.f);
      for (int r = 0; r < rows; ++r)
        for (int u = 0; u < U; ++u) v[(size_t)r * Up + u] = src[(size_t)r * U + u];
      return pool.upload(v);
    };
    m.keys = keys_p; m.values = values;
    m.v = padU(W.get(P + "attention_v", U), 1);
    std::vector<float> zeros(Up, 0.f);
    m.ba = pool.upload(zeros);
    if (kind != SAT_ATT_ADDITIVE) {
      SAT_REQUIRE(state_alignments, "location-sensitive kinds need state_alignments");
      SAT_REQUIRE(attention_kernel <= 2 * dec::HALO && attention_filters <= 32, "attention_kernel/filters too large");
      SAT_REQUIRE(attention_filters <= 8 || (128 / std::min(Up, 128)) * attention_filters <= 512, "too many filters");
      m.ba = padU(W.get(P + "attention_bias", U), 1);
      m.conv_w = pool.upload(W.get(P + "location_conv/kernel", (int64_t)attention_kernel * attention_filters),
                             (size_t)attention_kernel * attention_filters);
      m.conv_b = pool.upload(W.get(P + "location_conv/bias", attention_filters), attention_filters);
      m.Wl = padU(W.get(P + "location_layer/kernel", (int64_t)attention_filters * U), attention_filters);
    }
    if (kind == SAT_ATT_FORWARD) {
      SAT_REQUIRE(state_alpha && state_u, "forward attention needs alpha and u state");
      if (use_transition_agent) {
        m.agent = 1;
        m.Wt = pool.upload(W.get(P + "transition/kernel", Dv + U), Dv + U);
        m.bt = W.get(P + "transition/bias", 1)[0];
      }
    }
    float* dummy_state = pool.alloc<float>((size_t)B * T);
    m.pq = pq; m.pq_su = 1; m.pq_sb = U;
    m.st_align = state_alignments ? state_alignments : dummy_state;
    m.st_alpha = state_alpha; m.st_u = state_u;
    m.out_align = alignments; m.out_S = 1;
    m.ctx_dst[0] = context; m.ctx_dst[1] = nullptr; m.ctx_sd[0] = 1; m.ctx_sb[0] = Dv;
    m.TS = dec::MAX_T; m.FB = 8 * dec::MAX_T;
    const size_t smem = dec::ATT_SMEM_FLOATS * sizeof(float);
    SAT_CUDA(cudaFuncSetAttribute(attention_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attention_step_kernel<<<B < 148 ? B : 148, dec::NT, smem, st>>>(m, B, T, (const long long*)lengths);
    count_launch();
    check_launch("attention_step_kernel");
    SAT_CUDA(cudaStreamSynchronize(st));
  });
}

extern "C" int sat_embedding_lookup(const float* table_host, int64_t table_rows, int32_t dim, int64_t index_offset,
                                    const int64_t* ids, int64_t n, float* out, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(table_host && ids && out && table_rows > 0 && dim > 0 && n >= 0, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    DevPool pool;
    float* table = pool.upload(table_host, (size_t)table_rows * dim);
    int* bad = pool.alloc<int>(1);
    if (n > 0) {
      embedding_kernel<<<(unsigned)ceil_div64(n * dim, 256), 256, 0, st>>>(table, table_rows, dim, index_offset,
                                                                          (const long long*)ids, n, out, bad);
      count_launch();
      check_launch("embedding_kernel");
    }
    int hbad = 0;
    SAT_CUDA(cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    SAT_CUDA(cudaStreamSynchronize(st));
    SAT_REQUIRE(!hbad, "embedding id out of range [%lld, %lld)", (long long)index_offset, (long long)(index_offset + table_rows));
  });
}
