// tc.cuh -- sm_100a tensor-core building blocks (inline PTX): mbarrier, TMA (cp.async.bulk.tensor), TMEM
// allocation, tcgen05.mma kind::tf32 with shared-memory descriptors, tcgen05.ld.
//
// Operand convention used everywhere in this library: both operands K-major in shared memory with the 128-byte
// swizzle (one row = 32 tf32 = 128 B; 8-row groups of 1024 B), exactly what a TMA box {32 floats, rows} with
// CU_TENSOR_MAP_SWIZZLE_128B writes.  D[M=128, N] accumulates in TMEM (row i = lane i, column j = column j).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace sat {
namespace tc {

constexpr int CHUNK_K = 32;                 // tf32 elements per 128-byte swizzle row
constexpr int UMMA_K = 8;                   // K of one tcgen05.mma kind::tf32
constexpr uint32_t SPIN_LIMIT = 1u << 28;   // bounded waits: a protocol bug traps instead of hanging the GPU

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t n = 0;
  while (!mbar_try_wait(bar, parity))
    if (++n > SPIN_LIMIT) __trap();
}

// ---- TMA: 2-D tiled load global -> shared, completion on an mbarrier ------------------------------------------
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int32_t x, int32_t y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}
// 1-D bulk copy global -> shared (src and dst 16-byte aligned, bytes a multiple of 16), completing on the mbarrier.
// Measured on B200 (scripts/dev/tma_bench.cu): the SM's TMA unit retires ONE copy or tensor box per ~0.28 us whatever
// its size up to 32 KB (64 / 128 / 256-row boxes alike, 1 or 148 CTAs, 2 or 12 in flight) and moves larger copies at
// 130 - 220 GB/s -- so a stream is paid for in instructions: few, large copies.
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// ---- TMEM ---------------------------------------------------------------------------------------------------
// one full warp; ncols = power of two >= 32; the base address lands in *slot (shared memory)
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- descriptors -----------------------------------------------------------------------------------------------
// K-major operand, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO); LBO unused.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(1024u >> 4) << 32;  // stride byte offset
  d |= 1ull << 46;                    // descriptor version (sm_100)
  d |= 2ull << 61;                    // SWIZZLE_128B
  return d;
}
// kind::tf32, fp32 accumulate, A and B K-major, dense
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// kind::f16 with fp16 operands, fp32 accumulate, A and B K-major, dense.  fp16 carries the same 11-bit significand
// as tf32 in half the bytes, so the fp16 form halves the L2 -> shared-memory stream of a K-major operand.
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
constexpr int CHUNK_K16 = 64;  // fp16 elements per 128-byte swizzle row
constexpr int UMMA_K16 = 16;   // K of one tcgen05.mma kind::f16

// D[tmem] (+)= A[smem] * B[smem]^T ; issued by ONE thread
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The same, called by EVERY lane of a warp with warp-uniform arguments; one elected lane issues.  A loop run by a single
// thread (if (tid == 32)) keeps descriptors and addresses in per-thread registers, and every tcgen05.mma then pays for
// R2UR transfers behind elect loops: ~130 cycles per instruction on B200 instead of ~20 -- with N = 16 tiles that issue
// time, not the data, bounded the gate GEMMs.
// alo / blo: low words of the SW128 operand descriptors ((shared address & 0x3FFFF) >> 4; consecutive K-steps of 32
// bytes are +2); the high word is the same for every K-major SW128 operand
constexpr uint32_t DESC_SW128_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr) { return (saddr & 0x3FFFFu) >> 4; }
__device__ __forceinline__ void mma_f16_warp(uint32_t tmem_d, uint32_t alo, uint32_t blo, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 da, {%1, %5};\n\t"
      "mov.b64 db, {%2, %5};\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(alo), "r"(blo), "r"(idesc), "r"(accumulate), "r"(DESC_SW128_HI)
      : "memory");
}
__device__ __forceinline__ void mma_commit_warp(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on `bar` once every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// 16 consecutive fp32 columns of this thread's TMEM lane (warp w may touch lanes 32*(w%4) .. +31 only)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// 4 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
  uint32_t r[4];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}

// fp32 -> fp16, round to nearest, saturating at the largest finite half instead of overflowing to inf
__device__ __forceinline__ __half to_f16(float x) { return __float2half_rn(fminf(fmaxf(x, -65504.f), 65504.f)); }

// round-to-nearest fp32 -> tf32 (the tensor core itself TRUNCATES fp32 operands, which biases every product
// towards zero; operands stored pre-rounded make the error unbiased and half as large)
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}

// byte offset of element (row n, k) inside one 32-k chunk image of a K-major SW128 operand (chunk base 1024-aligned)
__device__ __forceinline__ uint32_t sw128_offset(int n, int kk) {
  return (uint32_t)(n * 128 + ((((kk >> 2) ^ (n & 7)) << 4) | ((kk & 3) << 2)));
}

// the same for an fp16 operand: 64-k chunk image, 2-byte elements
__device__ __forceinline__ uint32_t sw128_offset16(int n, int kk) {
  return (uint32_t)(n * 128 + ((((kk >> 3) ^ (n & 7)) << 4) | ((kk & 7) << 1)));
}

}  // namespace tc

// host: 2-D tensor map over a row-major fp32 matrix [rows][cols] (cols contiguous), box = {32 floats, box_rows},
// 128-byte swizzle (tc_host.cu)
void make_tensor_map_2d(CUtensorMap* map, const float* base, uint64_t rows, uint64_t cols, uint64_t row_stride_floats,
                        uint32_t box_rows);
// un-swizzled fp32 box {box_cols, box_rows} (box_cols * 4 a multiple of 16 bytes): staging loads
void make_tensor_map_2d_plain(CUtensorMap* map, const float* base, uint64_t rows, uint64_t cols, uint64_t row_stride_floats,
                              uint32_t box_cols, uint32_t box_rows);
// the same over a row-major fp16 matrix, box = {64 halves, box_rows}
void make_tensor_map_2d_f16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_elems,
                            uint32_t box_rows);

}  // namespace sat
