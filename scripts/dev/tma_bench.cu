// Developer micro-benchmark (not part of the library): how fast does ONE SM's TMA unit deliver [rows x 64] fp16 operand
// boxes out of L2 when (a) every CTA reads the same rows, (b) every CTA reads its own rows, with the row stride of the
// decoder's activation buffers (1536 B) or contiguous tiles?  148 CTAs, one producer thread each, a ring of `stages`
// boxes, the consumer only waits for the bytes.   Build: nvcc -arch=sm_100a -O3 -o tma_bench tma_bench.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}" ::"r"(smem_u32(b)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma2d(void* dst, const CUtensorMap* m, uint64_t* bar, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
               "l"(m), "r"(smem_u32(bar)), "r"(x), "r"(y)
               : "memory");
}

__device__ __forceinline__ void bulk1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// 1-D bulk copies of `bytes` each, `stages` in flight, every CTA its own (or the same) source block
__global__ void __launch_bounds__(128) bench1d(const uint8_t* src, uint32_t bytes, int n_ops, int stages, int own, size_t src_bytes,
                                               long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* ring = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[16];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) mbar_init(full + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t* base = src + (own ? ((size_t)blockIdx.x * bytes * 4) % (src_bytes - (size_t)bytes * 4) : 0);
    int issued = 0, waited = 0;
    const long long t0 = clock64();
    while (waited < n_ops) {
      while (issued < n_ops && issued - waited < stages) {
        const int s = issued % stages;
        mbar_expect(full + s, bytes);
        bulk1d(ring + (size_t)s * bytes, base + (size_t)(issued % 4) * bytes, bytes, full + s);
        ++issued;
      }
      mbar_wait(full + waited % stages, (waited / stages) & 1);
      ++waited;
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

// rewrite the source region (coalesced or 2-byte scattered stores) so that the copies that follow read freshly written lines
__global__ void dirty(uint16_t* x, size_t n_halves, int scattered, uint16_t v) {
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (size_t)gridDim.x * blockDim.x;
  if (scattered) {
    // thread t writes element (row = t % 128, col = t / 128) of consecutive [128][64] tiles: 2-byte stores, 128 B apart within a warp
    for (size_t i = tid; i < n_halves; i += nt) {
      const size_t tile = i >> 13, e = i & 8191;
      x[(tile << 13) + (e & 127) * 64 + (e >> 7)] = v;
    }
  } else {
    for (size_t i = tid; i < n_halves; i += nt) x[i] = v;
  }
}

// mode bit 0: every CTA its own rows; kchunks boxes per "phase", `phases` phases
__global__ void __launch_bounds__(128) bench(const __grid_constant__ CUtensorMap map, int box_rows, int kchunks, int phases, int stages,
                                             int own_rows, int rows_total, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* ring = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[16];
  const int sbytes = box_rows * 128;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) mbar_init(full + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int row0 = own_rows ? (int)((blockIdx.x * box_rows) % (rows_total - box_rows + 1)) : 0;
    const long long t0 = clock64();
    int issued = 0, waited = 0;
    const int total = kchunks * phases;
    while (waited < total) {
      while (issued < total && issued - waited < stages) {
        const int s = issued % stages;
        mbar_expect(full + s, sbytes);
        tma2d(ring + s * sbytes, &map, full + s, (issued % kchunks) * 64, row0);
        ++issued;
      }
      mbar_wait(full + waited % stages, (waited / stages) & 1);
      ++waited;
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

int main(int argc, char** argv) {
  const int box_rows = argc > 1 ? atoi(argv[1]) : 128, stages = argc > 2 ? atoi(argv[2]) : 6, own = argc > 3 ? atoi(argv[3]) : 0;
  const int K = argc > 4 ? atoi(argv[4]) : 768, grid = argc > 5 ? atoi(argv[5]) : 148;
  const int rows_total = 148 * 128, kchunks = K / 64, phases = 200;
  __half* x;
  CK(cudaMalloc(&x, (size_t)rows_total * K * 2));
  CK(cudaMemset(x, 0, (size_t)rows_total * K * 2));
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  auto enc = (CUresult(*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                          const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill))fn;
  CUtensorMap map;
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows_total}, strides[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows}, es[2] = {1, 1};
  CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, x, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  long long* cyc;
  CK(cudaMalloc(&cyc, 148 * 8));
  const size_t smem = (size_t)stages * box_rows * 128 + 1024;
  CK(cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    bench<<<grid, 128, smem>>>(map, box_rows, kchunks, phases, stages, own, rows_total, cyc);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)grid * kchunks * phases * box_rows * 128;
    printf("box_rows %d stages %d own_rows %d K %d grid %d: %.3f ms, %.1f GB/s per SM, %.2f TB/s total, %.3f us per box\n", box_rows, stages,
           own, K, grid, ms, bytes / grid / (ms * 1e-3) / 1e9, bytes / (ms * 1e-3) / 1e12, ms * 1e3 / (kchunks * phases));
  }
  if (argc > 6) {  // 1-D bulk copies: argv[6] = bytes per copy, argv[2] = stages
    const uint32_t bytes = (uint32_t)atoi(argv[6]);
    const int n_ops = 2000;
    const size_t smem1 = (size_t)stages * bytes + 1024;
    CK(cudaFuncSetAttribute(bench1d, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      bench1d<<<grid, 128, smem1>>>((const uint8_t*)x, bytes, n_ops, stages, own, (size_t)rows_total * K * 2, cyc);
      cudaEventRecord(e1);
      CK(cudaDeviceSynchronize());
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      printf("1-D bulk %u bytes, stages %d, own %d, grid %d: %.3f us per copy, %.1f GB/s per SM, %.2f TB/s total\n", bytes, stages, own, grid,
             ms * 1e3 / n_ops, (double)bytes * n_ops / (ms * 1e-3) / 1e9, (double)bytes * n_ops * grid / (ms * 1e-3) / 1e12);
    }
  }
  if (argc > 7) {  // argv[7] = 1: coalesced rewrite, 2: scattered 2-byte rewrite of the source before every short pass of 4 copies
    const uint32_t bytes = (uint32_t)atoi(argv[6]);
    const int mode = atoi(argv[7]);
    const size_t smem1 = (size_t)stages * bytes + 1024;
    const size_t region = own ? (size_t)grid * bytes * 4 + (size_t)bytes * 4 : (size_t)bytes * 4;
    std::vector<long long> hc(148);
    double sum = 0;
    int cnt = 0;
    for (int rep = 0; rep < 20; ++rep) {
      if (mode) dirty<<<148 * 4, 512>>>((uint16_t*)x, region / 2, mode == 2, (uint16_t)rep);
      bench1d<<<grid, 128, smem1>>>((const uint8_t*)x, bytes, 4, stages, own, (size_t)rows_total * K * 2, cyc);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(hc.data(), cyc, grid * 8, cudaMemcpyDeviceToHost));
      if (rep >= 4) for (int i = 0; i < grid; ++i) { sum += hc[i]; ++cnt; }
    }
    printf("pass of 4 copies of %u B after rewrite mode %d (own %d, grid %d, stages %d): %.0f cycles per pass -> %.1f B/cycle per SM\n", bytes, mode, own,
           grid, stages, sum / cnt, 4.0 * bytes / (sum / cnt));
  }
  return 0;
}
