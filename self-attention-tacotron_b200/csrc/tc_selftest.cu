// tc_selftest.cu -- host helper for TMA tensor maps and a one-CTA self-test of the tcgen05 building blocks
// (TMA SWIZZLE_128B loads -> tcgen05.mma kind::tf32 -> TMEM -> tcgen05.ld).  D[128,16] = A[128,K] . Wt[16,K]^T.
#include "common.cuh"
#include "tc.cuh"

namespace sat {

// cuTensorMapEncodeTiled is resolved through the runtime (cudaGetDriverEntryPoint) instead of linking libcuda, so
// that libsat.so still loads -- and its symbol table can be checked -- on a machine without a GPU driver.
namespace {
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  if (!fn) fail(SAT_ECUDA, "cuTensorMapEncodeTiled is not available from the installed CUDA driver");
  return fn;
}

void encode_2d(CUtensorMap* map, CUtensorMapDataType dt, size_t elem, uint32_t box_k, const void* base, uint64_t rows,
               uint64_t cols, uint64_t row_stride_elems, uint32_t box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * elem};
  cuuint32_t box[2] = {box_k, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = encode_tiled()(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail(SAT_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
}
}  // namespace

// un-swizzled fp32 box {box_cols, box_rows}: a staging load (the consumer converts / re-lays-out in shared memory)
void make_tensor_map_2d_plain(CUtensorMap* map, const float* base, uint64_t rows, uint64_t cols, uint64_t row_stride_floats,
                              uint32_t box_cols, uint32_t box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_floats * 4};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = encode_tiled()(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail(SAT_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
}

void make_tensor_map_2d(CUtensorMap* map, const float* base, uint64_t rows, uint64_t cols, uint64_t row_stride_floats,
                        uint32_t box_rows) {
  encode_2d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, tc::CHUNK_K, base, rows, cols, row_stride_floats, box_rows);
}

void make_tensor_map_2d_f16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_elems,
                            uint32_t box_rows) {
  encode_2d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, tc::CHUNK_K16, base, rows, cols, row_stride_elems, box_rows);
}

namespace {
constexpr int ST_STAGES = 2;

// M = 128: D row i = TMEM lane i.  M = 64 (SAT_SELFTEST_M=64; A rows 64..127 are not read): the raw dump of all 128
// lanes shows where the 64 accumulator rows live (rows 16 q .. 16 q + 15 in lanes 32 q .. 32 q + 15).
__global__ void __launch_bounds__(128) tc_selftest_kernel(const CUtensorMap* __restrict__ mapA, const float* __restrict__ Wt,
                                                          int K, float* __restrict__ D, int M) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int nchunks = K / tc::CHUNK_K;
  uint8_t* a_ring = smem;                             // ST_STAGES x 16 KB
  uint8_t* b_img = a_ring + ST_STAGES * 16384;        // nchunks x 2 KB
  uint64_t* bars = (uint64_t*)(b_img + nchunks * 2048);
  uint64_t* full = bars;                              // [ST_STAGES]
  uint64_t* empty = bars + ST_STAGES;                 // [ST_STAGES]
  uint64_t* done = bars + 2 * ST_STAGES;
  uint32_t* tmem_slot = (uint32_t*)(bars + 2 * ST_STAGES + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < ST_STAGES; ++s) {
      tc::mbar_init(full + s, 1);
      tc::mbar_init(empty + s, 1);
    }
    tc::mbar_init(done, 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(tmem_slot, 32);
  // resident B image: Wt [16][K] row-major -> per 32-k chunk, K-major SW128
  for (int i = tid; i < 16 * K; i += blockDim.x) {
    const int n = i / K, k = i % K;
    *(float*)(b_img + (k / tc::CHUNK_K) * 2048 + tc::sw128_offset(n, k % tc::CHUNK_K)) = Wt[i];
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (tid == 0) {
    for (int c = 0; c < nchunks; ++c) {
      const int s = c % ST_STAGES;
      const uint32_t use = c / ST_STAGES;
      if (use > 0) tc::mbar_wait(empty + s, (use - 1) & 1);
      tc::mbar_arrive_expect_tx(full + s, M * 128);
      tc::tma_load_2d(a_ring + s * 16384, mapA, full + s, c * tc::CHUNK_K, 0);
    }
  } else if (tid == 32) {
    const uint32_t idesc = tc::idesc_tf32(M, 16);
    for (int c = 0; c < nchunks; ++c) {
      const int s = c % ST_STAGES;
      tc::mbar_wait(full + s, (c / ST_STAGES) & 1);
      tc::tc_fence_after();
      const uint32_t a0 = tc::smem_u32(a_ring + s * 16384), b0 = tc::smem_u32(b_img + c * 2048);
#pragma unroll
      for (int k8 = 0; k8 < tc::CHUNK_K / tc::UMMA_K; ++k8)
        tc::mma_tf32(tmem, tc::smem_desc_sw128(a0 + k8 * 32), tc::smem_desc_sw128(b0 + k8 * 32), idesc, (c | k8) != 0);
      tc::mma_commit(empty + s);
    }
    tc::mma_commit(done);
  }
  __syncwarp();
  tc::mbar_wait(done, 0);
  tc::tc_fence_after();
  float v[16];
  tc::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16), v);
#pragma unroll
  for (int i = 0; i < 16; ++i) D[(size_t)(warp * 32 + lane) * 16 + i] = v[i];
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 32);
}
}  // namespace
}  // namespace sat

using namespace sat;

// A [128][K] and Wt [16][K] device row-major fp32 (K multiple of 32, <= 1024); D [128][16] device
extern "C" int sat_tc_selftest(const float* A, const float* Wt, int32_t K, float* D, void* stream) {
  return guarded([&] {
    SAT_REQUIRE(A && Wt && D && K >= 32 && K % 32 == 0 && K <= 1024, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CUtensorMap hmap;
    int M = 128;
    if (const char* e = getenv("SAT_SELFTEST_M")) M = atoi(e) == 64 ? 64 : 128;
    make_tensor_map_2d(&hmap, A, 128, (uint64_t)K, (uint64_t)K, (uint32_t)M);
    DevPool pool;
    CUtensorMap* dmap = (CUtensorMap*)pool.alloc<uint8_t>(sizeof(CUtensorMap) + 64);
    dmap = (CUtensorMap*)(((uintptr_t)dmap + 63) & ~(uintptr_t)63);
    SAT_CUDA(cudaMemcpyAsync(dmap, &hmap, sizeof(hmap), cudaMemcpyHostToDevice, st));
    const size_t smem = 1024 + ST_STAGES * 16384 + (size_t)(K / 32) * 2048 + 256;
    SAT_CUDA(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tc_selftest_kernel<<<1, 128, smem, st>>>(dmap, Wt, K, D, M);
    count_launch();
    check_launch("tc_selftest_kernel");
    SAT_CUDA(cudaStreamSynchronize(st));
  });
}
