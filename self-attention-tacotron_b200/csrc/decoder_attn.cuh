// decoder_attn.cuh -- one attention-mechanism call for one utterance row, executed by a
// whole CTA (row-split phase of the persistent decoder kernel; also the body of the
// stand-alone sat_attention_step entry point).
//
// Restates, per row b (citations into /root/reference):
//   forward            modules/forward_attention.py:88-122 (+ _location_sensitive_score :13-26)
//   location_sensitive tacotron2 LocationSensitiveAttention (modules/attentions.py:35-42)
//   additive           TF BahdanauAttention (modules/attentions.py:53-57)
//   teacher forcing    modules/teacher_forcing_attention.py:29-35
//   context            _calculate_context modules/forward_attention.py:29-41
#pragma once
#include "decoder_dev.cuh"
#include "tc.cuh"
#include "../../include/sat.h"

namespace sat {
namespace dec {

constexpr int HALO = 32;         // zero halo each side of the previous alignments (attention_kernel <= 64)
constexpr int MAX_T = 1024;      // padded memory length bound of the row phase
constexpr int MAX_DV = 512;       // attention memory depth bound (context staging)
constexpr int ATT_SMEM_FLOATS = 13 * MAX_T + 2 * HALO + NT + MAX_DV + 64;

// unit axis padding of keys / per-unit vectors: 4 * 2^k lanes-worth for U <= 128, else 256 (two warps per row)
inline int att_padded_units(int U) {
  if (U > 128) return 256;
  int c = 1;
  while (4 * c < U) c <<= 1;
  return 4 * c;
}

struct AttMech {
  int kind;             // SAT_ATT_*
  int U, Up, Dv, F, Kc, pad_left, cumulative, agent;
  const float* keys;    // [B][T][Up]   (Up = U rounded up to 4, zero padded)
  const float* values;  // [B][T][Dv]
  const float* v;       // [Up]
  const float* ba;      // [Up] attention_bias (zeros for additive)
  const float* conv_w;  // [Kc][F]
  const float* conv_b;  // [F]
  const float* Wl;      // [F][Up] location_layer
  const float* Wt;      // [Dv + U] transition agent kernel
  float bt;
  const float* pq;      // processed query, element (u, b) at pq[u * pq_su + b * pq_sb]
  int pq_su, pq_sb;
  float* st_align;      // [B][T] previous alignments (state)
  float* st_alpha;      // [B][T]
  float* st_u;          // [B]
  const float* teacher; // [B][S][T] or null
  int teacher_S;
  float* out_align;     // [B][out_S][T] or null
  int out_S;
  long long* prof;      // optional stage cycle counters (8 slots), CTA 0 only
  const __half* values16;  // optional fp16 copy of `values` (tensor-core mode)
  void* ctx_dst[2];     // context element (d, b) written at ctx_dst[i][d * ctx_sd[i] + b * ctx_sb[i]]; may be null
  int ctx_sd[2], ctx_sb[2];
  int ctx_f16;          // 1: the context destinations are fp16 operand-tile buffers of the tensor-core GEMMs (element
                        // tc_act_index(ctx_sd[i] + d, b, ctx_sb[i]): ctx_sd = feature offset, ctx_sb = chunks per row), 0: fp32
  // shared-memory plan of the row phase: TS = capacity of the per-position arrays (>= T, multiple of 16; MAX_T for the
  // stand-alone entry point), FB = floats of the location-feature / context-partial buffer (>= 8 TS and >= 16 Dv)
  int TS, FB;
};

// floats of shared memory one attention row needs for a given plan
__host__ __device__ inline int att_smem_floats(int TS, int FB) {
  return 5 * TS + 2 * HALO + FB + NT + MAX_DV + 64 + (TS < MAX_T ? 8 * TS : 0);
}

// A row is processed either by the whole CTA (GT = NT, CTA barrier) or, when the CTA owns two rows of the phase, by
// one half of it (GT = NT / 2, named barrier 1 + half): the two rows then run concurrently, each stage of one hiding
// the barrier / L2 latencies of the other, instead of back to back.
template <int GT>
__device__ __forceinline__ void group_sync(int gid) {
  if (GT == NT) __syncthreads();
  else asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "n"(GT) : "memory");
}

// ---- thread-block cluster helpers (cluster-split attention rows of the small-batch form) ----------------------------
// A row's attention is split over the CS CTAs of a cluster: each CTA owns Up / CS attention units (its slice of the
// query layer and of the energies) and Dv / CS columns of the context; the per-unit-group partial energies are
// all-gathered through distributed shared memory (every CTA stores its partials into the same offset of every CTA's
// shared memory), one cluster barrier, then each CTA runs the (cheap) softmax / alpha recursion redundantly.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
// all threads of all CTAs of the cluster; release / acquire: shared-memory stores (local and remote) before it are
// visible after it
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// store v at the address of `local` in the shared memory of CTA `rank` of the cluster
__device__ __forceinline__ void st_cluster_f32(float* local, uint32_t rank, float v) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"((uint32_t)__cvta_generic_to_shared(local)), "r"(rank));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(v) : "memory");
}


// Energies of one row.  v . tanh(x) = sum(v) - 2 sum_u v_u / (exp(2 x_u) + 1); the constant sum(v) is the same
// for every j of the row and cancels in the softmax, so the kernel evaluates
//     e'_j = sum_u (-2 v_u) * rcp(ex2(C x_u) + 1),  C = 2 log2(e)
// with C folded into the processed query / bias / location weights (registers) and applied to the key with the
// FFMA that adds it.  Abs error of the implied tanh ~1e-7 (ex2.approx, rcp.approx); saturates cleanly.
//
// Lane mapping: one lane owns ONE float4 of the unit axis (keeps the per-lane constants at 8 + 4 F registers).
// Up = 4 * lpr with lpr a power of two <= 32 lanes per memory row (32 / lpr rows per warp pass), or Up = 256:
// two warps per memory row (units [0,128) and [128,256)), partial sums combined through e2.
// F_T: 0 = no location term (additive); 5 / 8 = location filters held in registers (8 = zero-padded 6..8);
// -1 = generic (any F <= 32, location features recomputed per row).
template <int F_T>
__device__ void attention_energies(const AttMech& m, int b, int T, int len, const float* aprev,
                                   const float* fbuf, float* e, float* e2) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr float C2 = 2.8853900817779268f;  // 2 * log2(e)
  constexpr int FR = F_T > 0 ? F_T : 1;
  const int halves = m.Up > 128 ? 2 : 1;
  const int lpr = halves == 2 ? 32 : (m.Up >> 2);  // lanes per memory row
  const int G = 32 / lpr;                          // memory rows per warp pass
  const int half = halves == 2 ? (warp & 1) : 0;
  const int wrow = halves == 2 ? (warp >> 1) : warp;
  const int nwr = halves == 2 ? NW / 2 : NW;       // warps along the row axis
  const int u0 = half * 128 + 4 * (lane % lpr);
  const int g = lane / lpr;
  float pqb[4], vv[4], wl[FR][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int u = u0 + i;
    pqb[i] = u < m.U ? C2 * (__ldcg(m.pq + (size_t)u * m.pq_su + (size_t)b * m.pq_sb) + __ldg(m.ba + u)) : 0.f;
    vv[i] = -2.f * __ldg(m.v + u);  // zero in the padding
    if (F_T > 0) {
#pragma unroll
      for (int c = 0; c < FR; ++c) wl[c][i] = c < m.F ? C2 * __ldg(m.Wl + c * m.Up + u) : 0.f;
    }
  }
  long long c0 = 0, c1 = 0, c2 = 0;
  const bool prof_on = m.prof && blockIdx.x == 0 && threadIdx.x == 0;
  if (prof_on) {
    float dep = pqb[0] + vv[0] + wl[0][0] + wl[FR - 1][3];
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c0) : "f"(dep));
  }
  const float* kbase = m.keys + (size_t)b * T * m.Up + u0;
  const int stride = nwr * G;
  constexpr int PF = 4;  // key rows in flight per lane
  float4 kq[PF];
  const int jbase = wrow * G + g;
#pragma unroll
  for (int s = 0; s < PF; ++s) {
    const int j = jbase + s * stride;
    kq[s] = j < len ? __ldg((const float4*)(kbase + (size_t)j * m.Up)) : make_float4(0, 0, 0, 0);
  }
  if (prof_on) {
    float dep = kq[0].x + kq[PF - 1].w;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c1) : "f"(dep));
  }
  float* eo = half ? e2 : e;
  for (int j0 = jbase; j0 - g < T; j0 += PF * stride) {
#pragma unroll
    for (int s = 0; s < PF; ++s) {
      const int j = j0 + s * stride;
      if (j - g >= T) break;  // warp-uniform
      const float4 kc = kq[s];
      {
        const int jn = j + PF * stride;
        kq[s] = jn < len ? __ldg((const float4*)(kbase + (size_t)jn * m.Up)) : make_float4(0, 0, 0, 0);
      }
      float sum = 0.f;
      float x[4] = {fmaf(kc.x, C2, pqb[0]), fmaf(kc.y, C2, pqb[1]), fmaf(kc.z, C2, pqb[2]), fmaf(kc.w, C2, pqb[3])};
      if (F_T > 0) {
        if (j < len) {
          const float4 f0 = *(const float4*)(fbuf + j * 8);
          const float4 f1 = FR > 4 ? *(const float4*)(fbuf + j * 8 + 4) : make_float4(0, 0, 0, 0);
          const float fj[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
#pragma unroll
          for (int c = 0; c < FR; ++c)
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = fmaf(fj[c], wl[c][i], x[i]);
        }
      } else if (F_T < 0) {
        // generic (9..32 filters): the warp first computes the location features of the G rows of this pass
        // into its private stash (fbuf is free in this mode), then every lane reads its row's features.
        float* stash = const_cast<float*>(fbuf) + warp * 512;  // G * F <= 512 enforced at create time
        __syncwarp();
        for (int idx = lane; idx < G * m.F; idx += 32) {
          const int gg = idx / m.F, c = idx % m.F;
          const int jj = j - g + gg;
          float fc = 0.f;
          if (jj < T) {
            fc = __ldg(m.conv_b + c);
            for (int i = 0; i < m.Kc; ++i) fc = fmaf(__ldg(m.conv_w + i * m.F + c), aprev[HALO + jj + i - m.pad_left], fc);
          }
          stash[idx] = fc * C2;
        }
        __syncwarp();
        for (int c2 = 0; c2 < m.F; ++c2) {
          const float f = stash[g * m.F + c2];
          const float4 w = __ldg((const float4*)(m.Wl + c2 * m.Up + u0));
          x[0] = fmaf(f, w.x, x[0]); x[1] = fmaf(f, w.y, x[1]); x[2] = fmaf(f, w.z, x[2]); x[3] = fmaf(f, w.w, x[3]);
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) sum = fmaf(vv[i], rcp_approx(ex2_approx(x[i]) + 1.f), sum);
      for (int o = 1; o < lpr; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      if ((lane % lpr) == 0 && j < T) eo[j] = (j < len) ? sum : (half ? 0.f : -INFINITY);
    }
  }
  if (prof_on) {
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c2));
    m.prof[6] += c1 - c0;  // first key rows landed
    m.prof[7] += c2 - c1;  // main loop of warp 0
  }
}

// Fast path of the energies for Up = 128 or 256 (a full warp spans 128 units; Up = 256: two warps per row).
// Four memory rows per pass and lane, software-pipelined key loads (next four rows in flight while the current
// four are evaluated), and ONE transpose-reduction per four rows (6 shuffles instead of 20).
// Only rows j < len are visited; the masked tail is filled by the caller.
template <int F_T>
__device__ void attention_energies_w32(const AttMech& m, int b, int T, int len, const float* fbuf, float* e, float* e2) {
  static_assert(F_T >= 0 && F_T <= 8, "register-resident location filters");
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr float C2 = 2.8853900817779268f;  // 2 * log2(e)
  constexpr int FR = F_T > 0 ? F_T : 1;
  const bool two = m.Up > 128;
  const int half = two ? (warp & 1) : 0;
  const int wrow = two ? (warp >> 1) : warp;
  const int nwr = two ? NW / 2 : NW;
  const int u0 = half * 128 + 4 * lane;
  float pqb[4], vv[4], wl[FR][4];
  {
    // per-unit constants: one 16-byte load each (all arrays are padded to Up with zeros)
    float4 q4;
    if (m.pq_su == 1) {
      q4 = __ldcg((const float4*)(m.pq + (size_t)b * m.pq_sb + u0));
    } else {
      q4.x = __ldcg(m.pq + (size_t)(u0 + 0) * m.pq_su + (size_t)b * m.pq_sb);
      q4.y = __ldcg(m.pq + (size_t)(u0 + 1) * m.pq_su + (size_t)b * m.pq_sb);
      q4.z = __ldcg(m.pq + (size_t)(u0 + 2) * m.pq_su + (size_t)b * m.pq_sb);
      q4.w = __ldcg(m.pq + (size_t)(u0 + 3) * m.pq_su + (size_t)b * m.pq_sb);
    }
    const float4 b4 = __ldg((const float4*)(m.ba + u0)), v4 = __ldg((const float4*)(m.v + u0));
    const float q[4] = {q4.x, q4.y, q4.z, q4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      pqb[i] = u0 + i < m.U ? C2 * (q[i] + bb[i]) : 0.f;
      vv[i] = -2.f * v[i];
    }
    if (F_T > 0) {
#pragma unroll
      for (int c = 0; c < FR; ++c) {
        float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < m.F) w4 = __ldg((const float4*)(m.Wl + c * m.Up + u0));
        wl[c][0] = C2 * w4.x; wl[c][1] = C2 * w4.y; wl[c][2] = C2 * w4.z; wl[c][3] = C2 * w4.w;
      }
    }
  }
  const float* kbase = m.keys + (size_t)b * T * m.Up + u0;
  const int last = len - 1;
  float* eo = half ? e2 : e;
  // rows of this warp: j = wrow + nwr * r, r = 0, 1, ...; out-of-range rows are clamped to `last` and discarded
  auto row_ptr = [&](int r) { return kbase + (size_t)min(wrow + nwr * r, last) * m.Up; };
  float4 kn[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) kn[s] = __ldg((const float4*)row_ptr(s));
  for (int r0 = 0; wrow + nwr * r0 < len; r0 += 4) {
    float4 kc[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      kc[s] = kn[s];
      kn[s] = __ldg((const float4*)row_ptr(r0 + 4 + s));
    }
    float sum[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const int j = min(wrow + nwr * (r0 + s), last);
      float x[4] = {fmaf(kc[s].x, C2, pqb[0]), fmaf(kc[s].y, C2, pqb[1]), fmaf(kc[s].z, C2, pqb[2]), fmaf(kc[s].w, C2, pqb[3])};
      if (F_T > 0) {
        const float4 f0 = *(const float4*)(fbuf + j * 8);
        const float4 f1 = FR > 4 ? *(const float4*)(fbuf + j * 8 + 4) : make_float4(0, 0, 0, 0);
        const float fj[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
#pragma unroll
        for (int c = 0; c < FR; ++c)
#pragma unroll
          for (int i = 0; i < 4; ++i) x[i] = fmaf(fj[c], wl[c][i], x[i]);
      }
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) acc = fmaf(vv[i], rcp_approx(ex2_approx(x[i]) + 1.f), acc);
      sum[s] = acc;
    }
    // transpose-reduce 4 rows x 32 lanes: afterwards lanes [8q, 8q+8) hold the total of row q
    const bool b4 = lane & 16, b3 = lane & 8;
    float a0 = b4 ? sum[2] : sum[0], a1 = b4 ? sum[3] : sum[1];
    a0 += __shfl_xor_sync(0xffffffffu, b4 ? sum[0] : sum[2], 16);
    a1 += __shfl_xor_sync(0xffffffffu, b4 ? sum[1] : sum[3], 16);
    float c = b3 ? a1 : a0;
    c += __shfl_xor_sync(0xffffffffu, b3 ? a0 : a1, 8);
    c += __shfl_xor_sync(0xffffffffu, c, 4);
    c += __shfl_xor_sync(0xffffffffu, c, 2);
    c += __shfl_xor_sync(0xffffffffu, c, 1);
    const int jq = wrow + nwr * (r0 + (lane >> 3));
    if ((lane & 7) == 0 && jq < len) eo[jq] = c;
  }
}

// ---- energies with the location term on the tensor cores (mma.sync m16n8k8, 3xTF32) ---------------------------
// x[j][u] = C2 * (key[j][u] + pq[u] + b[u] + sum_c f[j][c] Wl[c][u]) is a [T x 8] . [8 x Up] product added to the keys:
// A = [f[j][0..F) | 1 | 0..] (the constant-1 slot carries pq + b), B = C2 * [Wl ; pq + b ; 0], C = C2 * keys.  Both
// operands are split hi + lo (3 MMAs per tile) so the result keeps fp32 accuracy.  That leaves ex2 / rcp / a handful
// of FMAs per element for the SIMT pipes, and four reciprocals share one MUFU.RCP:
//   v0/y0 + v1/y1 + v2/y2 + v3/y3 = ((v0 y1 + v1 y0) y2 y3 + (v2 y3 + v3 y2) y0 y1) / (y0 y1 y2 y3),  y = 2^x + 1 <= 2^30 + 1.
// Warp w owns units [32 (w % NUG), +32) and row blocks of 16 positions; partial row sums over the warp's 32 units go
// to epart[w % NUG][j] and are summed by the softmax pass.
__device__ __forceinline__ void mma_tf32_16x8x8(float* c, const uint32_t* a, const uint32_t* b) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ float quad_recip_dot(float y0, float y1, float y2, float y3, float v0, float v1, float v2,
                                                float v3) {
  const float ab = y0 * y1, cd = y2 * y3;
  const float nab = fmaf(v0, y1, v1 * y0), ncd = fmaf(v2, y3, v3 * y2);
  return fmaf(nab, cd, ncd * ab) * rcp_approx(ab * cd);
}

template <int NUG, int GT>
__device__ void attention_energies_mma(const AttMech& m, int b, int T, int len, const float* fbuf, float* epart, int Tpad,
                                       int gtid, int ug_off = 0, int csize = 1) {
  constexpr float C2 = 2.8853900817779268f;  // 2 * log2(e)
  constexpr int NRG = (GT / 32) / NUG;
  const int lane = gtid & 31, warp = gtid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int ug = warp % NUG + ug_off, rg = warp / NUG;  // ug_off: first unit group of this CTA (cluster-split rows)
  const int ubase = 32 * ug;
  // first key block in flight before anything else (rows past the end are clamped to the last valid row)
  const float* kb = m.keys + (size_t)b * T * m.Up + ubase + 2 * t;
  const int last = len - 1;
  auto kptr = [&](int jb, int h) { return (const float2*)(kb + (size_t)min(16 * jb + g + 8 * h, last) * m.Up); };
  float2 kn[4][2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const float2* kp = kptr(rg, h);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) kn[nt][h] = __ldg(kp + 4 * nt);
  }
  // per-warp constants: all loads are issued unconditionally and up front (one L2 round trip), selected afterwards
  uint32_t bh[4][2], bl[4][2];
  float vv[4][2];
  {
    float wraw[4][2], qraw[4], braw[4];
    float2 vraw[4];
    const int k0 = min(t, max(m.F - 1, 0)), k1 = min(t + 4, max(m.F - 1, 0));
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int ub = ubase + 8 * nt + g;  // B fragment: feature slot k = t (+4), unit n = g
      wraw[nt][0] = __ldg(m.Wl + k0 * m.Up + ub);
      wraw[nt][1] = __ldg(m.Wl + k1 * m.Up + ub);
      qraw[nt] = __ldcg(m.pq + (size_t)ub * m.pq_su + (size_t)b * m.pq_sb);
      braw[nt] = __ldg(m.ba + ub);
      vraw[nt] = __ldg((const float2*)(m.v + ubase + 8 * nt + 2 * t));  // C fragment columns 2t, 2t+1
    }
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int ub = ubase + 8 * nt + g;
      const float pqb = ub < m.U ? C2 * (qraw[nt] + braw[nt]) : 0.f;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int k = t + 4 * h;
        const float w = k < m.F ? C2 * wraw[nt][h] : (k == m.F ? pqb : 0.f);
        split_tf32(w, bh[nt][h], bl[nt][h]);
      }
      vv[nt][0] = -2.f * vraw[nt].x;
      vv[nt][1] = -2.f * vraw[nt].y;
    }
  }
  for (int jb = rg; 16 * jb < len; jb += NRG) {
    float c[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      c[nt][0] = C2 * kn[nt][0].x; c[nt][1] = C2 * kn[nt][0].y;
      c[nt][2] = C2 * kn[nt][1].x; c[nt][3] = C2 * kn[nt][1].y;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float2* kp = kptr(jb + NRG, h);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) kn[nt][h] = __ldg(kp + 4 * nt);
    }
    // A fragment: rows g / g+8 of the block, feature slots t / t+4
    const float* fa = fbuf + (16 * jb + g) * 8 + t;
    uint32_t ah[4], al[4];
    split_tf32(fa[0], ah[0], al[0]);
    split_tf32(fa[64], ah[1], al[1]);
    split_tf32(fa[4], ah[2], al[2]);
    split_tf32(fa[68], ah[3], al[3]);
    float y[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      mma_tf32_16x8x8(c[nt], al, bh[nt]);
      mma_tf32_16x8x8(c[nt], ah, bl[nt]);
      mma_tf32_16x8x8(c[nt], ah, bh[nt]);
#pragma unroll
      for (int i = 0; i < 4; ++i) y[nt][i] = ex2_approx(fminf(c[nt][i], 30.f)) + 1.f;
    }
    // row g: elements [nt][0..1], row g+8: elements [nt][2..3]
    float r0 = quad_recip_dot(y[0][0], y[0][1], y[1][0], y[1][1], vv[0][0], vv[0][1], vv[1][0], vv[1][1]) +
               quad_recip_dot(y[2][0], y[2][1], y[3][0], y[3][1], vv[2][0], vv[2][1], vv[3][0], vv[3][1]);
    float r1 = quad_recip_dot(y[0][2], y[0][3], y[1][2], y[1][3], vv[0][0], vv[0][1], vv[1][0], vv[1][1]) +
               quad_recip_dot(y[2][2], y[2][3], y[3][2], y[3][3], vv[2][0], vv[2][1], vv[3][0], vv[3][1]);
    // reduce over the 4 lanes of a quad: odd t ends up with row g+8, even t with row g
    const bool odd = t & 1;
    float s = odd ? r1 : r0;
    s += __shfl_xor_sync(0xffffffffu, odd ? r0 : r1, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    const int j = 16 * jb + g + (odd ? 8 : 0);
    if (t < 2 && j < len) {
      float* dst = epart + ug * Tpad + j;
      if (csize == 1) *dst = s;
      else
        for (int cr = 0; cr < csize; ++cr) st_cluster_f32(dst, (uint32_t)cr, s);  // all-gather over the cluster
    }
  }
}

// context = alignments @ values for one row: warp per memory row, lanes over float4 chunks of the depth.
// CH = float4 chunks per lane per memory row (Dv = 4 * Cq; Cq <= 32: CH = 1 and 32 / Cq rows per warp pass;
// Cq = 64 / 128: CH = 2 / 4).  part = NW * Dv floats of shared memory.
template <int CH>
__device__ void attention_context_vec(const float* __restrict__ vb, const float* align, int len, int Dv,
                                      float* part, float* ctx_out /* smem [Dv] */, int rstride = 0 /* 0: Dv */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int Cq = Dv >> 2;
  const int lpr = CH > 1 ? 32 : Cq;      // lanes per memory row
  const int G = 32 / lpr;                // memory rows per warp pass
  const int c0 = lane % lpr, g = lane / lpr;
  const size_t rs = rstride ? rstride : Dv;  // floats between memory rows (a column window of a wider memory)
  float4 acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  constexpr int UN = CH > 1 ? 4 : 8;     // rows in flight per lane
  const int stride = NW * G;
  for (int j0 = warp * G + g; j0 < len; j0 += UN * stride) {
    float4 v[UN][CH];
    float a[UN];
#pragma unroll
    for (int s = 0; s < UN; ++s) {
      const int j = j0 + s * stride;
      const bool ok = j < len;
      a[s] = ok ? align[j] : 0.f;
#pragma unroll
      for (int c = 0; c < CH; ++c)
        v[s][c] = ok ? __ldg((const float4*)(vb + (size_t)j * rs) + c0 + 32 * c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int s = 0; s < UN; ++s)
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        acc[c].x = fmaf(a[s], v[s][c].x, acc[c].x); acc[c].y = fmaf(a[s], v[s][c].y, acc[c].y);
        acc[c].z = fmaf(a[s], v[s][c].z, acc[c].z); acc[c].w = fmaf(a[s], v[s][c].w, acc[c].w);
      }
  }
  for (int o = lpr; o < 32; o <<= 1) {
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      acc[c].x += __shfl_xor_sync(0xffffffffu, acc[c].x, o); acc[c].y += __shfl_xor_sync(0xffffffffu, acc[c].y, o);
      acc[c].z += __shfl_xor_sync(0xffffffffu, acc[c].z, o); acc[c].w += __shfl_xor_sync(0xffffffffu, acc[c].w, o);
    }
  }
  if (g == 0) {
#pragma unroll
    for (int c = 0; c < CH; ++c) *((float4*)(part + (size_t)warp * Dv) + c0 + 32 * c) = acc[c];
  }
  __syncthreads();
  for (int d = threadIdx.x; d < Dv; d += NT) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) s += part[w * Dv + d];
    ctx_out[d] = s;
  }
  __syncthreads();
}

// The same over an fp16 copy of the values (tensor-core mode: the context is rounded to fp16 on its way into the
// LSTM inputs anyway, and this phase is bound by the L2 -> SM bytes): a lane owns 8 halves, Dv / 8 lanes per row.
template <int GT>
__device__ inline void attention_context_h8(const __half* __restrict__ vb, const float* align, int len, int Dv, float* part,
                                            float* ctx_out /* smem [Dv] */, int gtid, int gid) {
  constexpr int NW = GT / 32;  // warps of the group (shadows the CTA-wide constant)
  const int lane = gtid & 31, warp = gtid >> 5;
  const int lpr = Dv >> 3, G = 32 / lpr;
  const int c0 = lane % lpr, g = lane / lpr;
  float acc[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) acc[c] = 0.f;
  constexpr int UN = 8;  // rows in flight per lane (16 measured no faster)
  const int stride = NW * G;
  for (int j0 = warp * G + g; j0 < len; j0 += UN * stride) {
    uint4 v[UN];
    float a[UN];
#pragma unroll
    for (int s = 0; s < UN; ++s) {
      const int j = j0 + s * stride;
      const bool ok = j < len;
      a[s] = ok ? align[j] : 0.f;
      v[s] = ok ? __ldg((const uint4*)(vb + (size_t)j * Dv) + c0) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int s = 0; s < UN; ++s) {
      const float2 f0 = __half22float2(*(const __half2*)&v[s].x), f1 = __half22float2(*(const __half2*)&v[s].y);
      const float2 f2 = __half22float2(*(const __half2*)&v[s].z), f3 = __half22float2(*(const __half2*)&v[s].w);
      acc[0] = fmaf(a[s], f0.x, acc[0]); acc[1] = fmaf(a[s], f0.y, acc[1]);
      acc[2] = fmaf(a[s], f1.x, acc[2]); acc[3] = fmaf(a[s], f1.y, acc[3]);
      acc[4] = fmaf(a[s], f2.x, acc[4]); acc[5] = fmaf(a[s], f2.y, acc[5]);
      acc[6] = fmaf(a[s], f3.x, acc[6]); acc[7] = fmaf(a[s], f3.y, acc[7]);
    }
  }
  for (int o = lpr; o < 32; o <<= 1) {
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
  }
  if (g == 0) {
    float4* dst = (float4*)(part + (size_t)warp * Dv + 8 * c0);
    dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
  group_sync<GT>(gid);
  for (int d = gtid; d < Dv; d += GT) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) s += part[w * Dv + d];
    ctx_out[d] = s;
  }
  group_sync<GT>(gid);
}

// One (row, mechanism) item, processed by a group of GT threads (GT = NT: the whole CTA, gid 0; GT = NT / 2: half
// gid of the CTA, only for the tensor-core energy form + fp16 context, no transition agent -- see att_row_pairable).
// `sm` = att_smem_floats(m.TS, m.FB) floats of shared memory owned by the group.  t = decoder step.
__device__ __forceinline__ bool att_row_pairable(const AttMech& m) {
  const int C8 = m.Dv >> 3;
  return (m.kind == SAT_ATT_FORWARD || m.kind == SAT_ATT_LOCATION_SENSITIVE) && m.F <= 7 && (m.Up == 128 || m.Up == 256) &&
         m.TS < MAX_T && !m.agent && m.values16 && (m.Dv & 7) == 0 && (C8 & (C8 - 1)) == 0 && C8 <= 32;
}

// crank / csize: rank of this CTA in, and size of, the cluster that shares the row (csize = 1: the row is not split;
// csize > 1 needs GT == NT, the tensor-core energy form (use_mma below) and no transition agent -- the host checks).
// previous alignments (zero-haloed) and forward variable of row b -> the row's shared memory (no barrier): lets a
// caller put these loads in flight together with other loads of its own (attention_item with preloaded = true)
template <int GT>
__device__ __forceinline__ void attention_load_state(const AttMech& m, int b, int T, float* sm, int gid = 0) {
  if (m.kind == SAT_ATT_TEACHER_FORCING) return;
  const int TS = m.TS;
  float* aprev = sm + TS;
  float* alpha = aprev + TS + 2 * HALO;
  const int tid = threadIdx.x - gid * GT;
  const bool loc = (m.kind == SAT_ATT_FORWARD || m.kind == SAT_ATT_LOCATION_SENSITIVE);
  for (int j = tid; j < T + 2 * HALO; j += GT) {
    const int jj = j - HALO;
    aprev[j] = (loc && jj >= 0 && jj < T) ? __ldcg(m.st_align + (size_t)b * T + jj) : 0.f;
  }
  if (m.kind == SAT_ATT_FORWARD)
    for (int j = tid; j < T; j += GT) alpha[j] = __ldcg(m.st_alpha + (size_t)b * T + j);
}

template <int GT>
__device__ __forceinline__ void attention_item_inl(const AttMech& m, int b, int t, int T, int len, float* sm, int gid = 0,
                                                   int crank = 0, int csize = 1, bool preloaded = false) {
  constexpr int GW = GT / 32;
  const int TS = m.TS;
  float* e = sm;                         // [TS] energies -> softmax probs a
  float* aprev = e + TS;                 // [TS + 2*HALO]
  float* alpha = aprev + TS + 2 * HALO;  // [TS]
  float* fbuf = alpha + TS;              // [FB] location features [TS * 8]; later the context partials [GW * Dv]
  float* nal = fbuf + m.FB;              // [TS] normalised forward variable
  float* e2 = nal + TS;                  // [TS] partial energies of units [128,256)
  float* part = e2 + TS;                 // [NT] context partials of the scalar path
  float* ctxs = part + NT;               // [MAX_DV] context of this row
  float* s32 = ctxs + MAX_DV;            // [64] block-reduce scratch
  float* epart_own = s32 + 64;           // [8 * TS] per-unit-group partial energies (TS < MAX_T only)
  const int tid = threadIdx.x - gid * GT;
  const bool loc = (m.kind == SAT_ATT_FORWARD || m.kind == SAT_ATT_LOCATION_SENSITIVE);
  const bool fwd = (m.kind == SAT_ATT_FORWARD);
  float* align = e;  // the alignments handed to the context / history
  const int lane = tid & 31, warp = tid >> 5;
  float u_prev = 0.f;
  long long tp = 0;
  int ps = 0;
  const bool profiling = m.prof && blockIdx.x == 0 && threadIdx.x == 0;
  if (profiling) tp = clock64();
  auto stamp = [&]() {
    if (profiling) {
      const long long now = clock64();
      m.prof[ps] += now - tp;
      tp = now;
    }
    ++ps;
  };

  if (m.kind == SAT_ATT_TEACHER_FORCING) {
    // index = prev_index + 1 = t ; alignments = teacher_alignments[:, index]
    const int ti = min(t, m.teacher_S - 1);
    for (int j = tid; j < T; j += GT) e[j] = __ldg(m.teacher + ((size_t)b * m.teacher_S + ti) * T + j);
    group_sync<GT>(gid);
  } else {
    if (fwd) u_prev = __ldcg(m.st_u + b);
    if (!preloaded) {
      attention_load_state<GT>(m, b, T, sm, gid);
      group_sync<GT>(gid);
    }
    stamp();
    if (loc && m.F <= 8) {
      // location features f[j][c] = b_c + sum_k w[k][c] a[j - pad + k]: a thread owns filter c = tid & 7 and four
      // consecutive positions, sliding a 4-wide register window over the taps (1 weight + 1 alignment LDS per tap)
      float* wsm = part;  // [Kc * 8] taps, zero-filled for c >= F
      for (int i = tid; i < m.Kc * 8; i += GT) wsm[i] = (i & 7) < m.F ? __ldg(m.conv_w + (i >> 3) * m.F + (i & 7)) : 0.f;
      group_sync<GT>(gid);
      const int c = tid & 7;
      const float cb = c < m.F ? __ldg(m.conv_b + c) : (c == m.F ? 1.f : 0.f);  // slot F = 1: carries pq + b in the MMA form
      for (int j0 = 4 * (tid >> 3); j0 < T; j0 += 4 * (GT / 8)) {
        const float* a = aprev + HALO + j0 - m.pad_left;
        float f0 = cb, f1 = cb, f2 = cb, f3 = cb;
        float x0 = a[0], x1 = a[1], x2 = a[2];
        for (int k = 0; k < m.Kc; ++k) {
          const float x3 = a[k + 3], w = wsm[k * 8 + c];
          f0 = fmaf(w, x0, f0);
          f1 = fmaf(w, x1, f1);
          f2 = fmaf(w, x2, f2);
          f3 = fmaf(w, x3, f3);
          x0 = x1;
          x1 = x2;
          x2 = x3;
        }
        fbuf[j0 * 8 + c] = f0;
        if (j0 + 1 < T) fbuf[(j0 + 1) * 8 + c] = f1;
        if (j0 + 2 < T) fbuf[(j0 + 2) * 8 + c] = f2;
        if (j0 + 3 < T) fbuf[(j0 + 3) * 8 + c] = f3;
      }
      group_sync<GT>(gid);
    }
    stamp();
    const int Tpad = (T + 15) & ~15;
    const int nug = m.Up >> 5;
    // partial energies per 32-unit group: own region, or (TS == MAX_T) the dead nal .. ctxs region of 3 * MAX_T floats
    const bool use_mma = loc && m.F <= 7 && (m.Up == 128 || m.Up == 256) && (TS < MAX_T || nug * Tpad <= 3 * MAX_T);
    float* epart = TS < MAX_T ? epart_own : nal;
    if (use_mma && csize > 1) {
      if constexpr (GT == NT) {
        // cluster-split row: this CTA evaluates nug / csize unit groups and stores its partials into every CTA
        const int nl = nug / csize;
        if (nl == 4) attention_energies_mma<4, GT>(m, b, T, len, fbuf, epart, Tpad, tid, 4 * crank, csize);
        else if (nl == 2) attention_energies_mma<2, GT>(m, b, T, len, fbuf, epart, Tpad, tid, 2 * crank, csize);
        else attention_energies_mma<1, GT>(m, b, T, len, fbuf, epart, Tpad, tid, crank, csize);
      }
    } else if (use_mma) {
      if (m.Up == 256) attention_energies_mma<8, GT>(m, b, T, len, fbuf, epart, Tpad, tid);
      else attention_energies_mma<4, GT>(m, b, T, len, fbuf, epart, Tpad, tid);
    } else if (GT == NT) {
      if (m.Up >= 128 && (!loc || m.F <= 8)) {
        if (!loc) attention_energies_w32<0>(m, b, T, len, fbuf, e, e2);
        else if (m.F <= 4) attention_energies_w32<4>(m, b, T, len, fbuf, e, e2);
        else if (m.F == 5) attention_energies_w32<5>(m, b, T, len, fbuf, e, e2);
        else attention_energies_w32<8>(m, b, T, len, fbuf, e, e2);
      } else if (!loc) attention_energies<0>(m, b, T, len, aprev, fbuf, e, e2);
      else if (m.F <= 4) attention_energies<4>(m, b, T, len, aprev, fbuf, e, e2);
      else if (m.F == 5) attention_energies<5>(m, b, T, len, aprev, fbuf, e, e2);
      else if (m.F <= 8) attention_energies<8>(m, b, T, len, aprev, fbuf, e, e2);
      else attention_energies<-1>(m, b, T, len, aprev, fbuf, e, e2);
    }
    if (csize > 1) cluster_sync_all();  // the partial energies of the other CTAs have landed
    else group_sync<GT>(gid);
    stamp();
    // masked softmax (TF: softmax(where(j < len, score, -inf))), the forward-variable recursion and the state update
    // with every element in registers (T <= NJ * GT) and one group barrier per reduction
    constexpr int NJ = MAX_T / NT * (NT / GT);
    float x[NJ], pr[NJ];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      const int j = tid + i * GT;
      x[i] = -INFINITY;
      if (j < len) {
        if (use_mma) {
          float a = 0.f;
          for (int q = 0; q < nug; ++q) a += epart[q * Tpad + j];
          x[i] = a;
        } else {
          x[i] = m.Up > 128 ? e[j] + e2[j] : e[j];
        }
      }
      mx = fmaxf(mx, x[i]);
    }
    mx = warp_max(mx);
    if (lane == 0) s32[warp] = mx;
    group_sync<GT>(gid);
    mx = warp_max(lane < GW ? s32[lane] : -INFINITY);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      pr[i] = (tid + i * GT < len) ? expf(x[i] - mx) : 0.f;
      sum += pr[i];
    }
    sum = warp_sum(sum);
    if (lane == 0) s32[NW + warp] = sum;
    group_sync<GT>(gid);
    sum = warp_sum(lane < GW ? s32[NW + lane] : 0.f);
    ps = 6;
    stamp();
    const float inv_sum = 1.f / sum;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      const int j = tid + i * GT;
      pr[i] = pr[i] * inv_sum;
      if (j < T) {
        e[j] = pr[i];
        if (crank == 0) m.st_align[(size_t)b * T + j] = m.cumulative ? pr[i] + aprev[HALO + j] : pr[i];
      }
    }
    if (fwd) {
      // alpha' = ((1-u) alpha + u shift(alpha) + 1e-7) * a ; alpha = alpha' / sum(alpha')
      float val[NJ], s2 = 0.f;
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        const int j = tid + i * GT;
        val[i] = 0.f;
        if (j < T) {
          const float sh = j > 0 ? alpha[j - 1] : 0.f;
          val[i] = (((1.f - u_prev) * alpha[j] + u_prev * sh) + 1e-7f) * pr[i];
        }
        s2 += val[i];
      }
      s2 = warp_sum(s2);
      if (lane == 0) s32[2 * NW + warp] = s2;
      group_sync<GT>(gid);
      s2 = warp_sum(lane < GW ? s32[2 * NW + lane] : 0.f);
      const float inv_s2 = 1.f / s2;
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        const int j = tid + i * GT;
        if (j < T) {
          const float a = val[i] * inv_s2;
          nal[j] = a;
          if (crank == 0) m.st_alpha[(size_t)b * T + j] = a;
        }
      }
      align = nal;
    }
    group_sync<GT>(gid);
    ps = 7;
    stamp();
  }
  if (m.out_align && crank == 0)
    for (int j = tid; j < T; j += GT) m.out_align[((size_t)b * m.out_S + t) * T + j] = align[j];
  ps = 3;
  stamp();

  // context = alignments @ values
  const int Dv = m.Dv, Cq = Dv >> 2;
  const float* vb = m.values + (size_t)b * T * Dv;
  const bool vec = (Dv & 3) == 0 && (Cq & (Cq - 1)) == 0 && Cq <= 128 && ((((size_t)vb) & 15) == 0);
  const int C8 = Dv >> 3;
  if (csize > 1) {
    if constexpr (GT == NT) {
      // cluster-split row: this CTA reduces its window of Dv / csize context columns over all memory rows
      const int win = Dv / csize, w0 = crank * win;
      attention_context_vec<1>(vb + w0, align, len, win, fbuf, ctxs, Dv);
      stamp();
      for (int d = tid; d < win; d += GT) {
        const float ctx = ctxs[d];
#pragma unroll
        for (int i = 0; i < 2; ++i)
          if (m.ctx_dst[i]) {
            if (m.ctx_f16) ((__half*)m.ctx_dst[i])[tc_act_index(m.ctx_sd[i] + w0 + d, b, m.ctx_sb[i])] = tc::to_f16(ctx);
            else ((float*)m.ctx_dst[i])[(size_t)(w0 + d) * m.ctx_sd[i] + (size_t)b * m.ctx_sb[i]] = ctx;
          }
      }
      group_sync<GT>(gid);
      stamp();
    }
    return;
  }
  if (m.values16 && (Dv & 7) == 0 && (C8 & (C8 - 1)) == 0 && C8 <= 32) {
    attention_context_h8<GT>(m.values16 + (size_t)b * T * Dv, align, len, Dv, fbuf, ctxs, tid, gid);
  } else if (GT == NT) {
    if (vec) {
      float* cpart = fbuf;  // location features are dead by now
      if (Cq <= 32) attention_context_vec<1>(vb, align, len, Dv, cpart, ctxs);
      else if (Cq == 64) attention_context_vec<2>(vb, align, len, Dv, cpart, ctxs);
      else attention_context_vec<4>(vb, align, len, Dv, cpart, ctxs);
    } else {
      const int G = max(1, NT / Dv);
      const int d = tid % Dv, grp = tid / Dv;
      float c = 0.f;
      if (grp < G && tid < G * Dv) {
#pragma unroll 4
        for (int j = grp; j < len; j += G) c = fmaf(align[j], __ldg(vb + (size_t)j * Dv + d), c);
      }
      part[tid] = c;
      __syncthreads();
      if (tid < Dv) {
        float ctx = 0.f;
        for (int g = 0; g < G; ++g) ctx += part[g * Dv + tid];
        ctxs[tid] = ctx;
      }
      __syncthreads();
    }
  }
  stamp();
  for (int d = tid; d < Dv; d += GT) {
    const float ctx = ctxs[d];
#pragma unroll
    for (int i = 0; i < 2; ++i)
      if (m.ctx_dst[i]) {
        if (m.ctx_f16) ((__half*)m.ctx_dst[i])[tc_act_index(m.ctx_sd[i] + d, b, m.ctx_sb[i])] = tc::to_f16(ctx);
        else ((float*)m.ctx_dst[i])[(size_t)d * m.ctx_sd[i] + (size_t)b * m.ctx_sb[i]] = ctx;
      }
  }
  if (GT == NT && fwd && m.agent) {
    // u = sigmoid([context, processed_query] . Wt + bt)   forward_attention.py:111-114
    float z = 0.f;
    for (int d = tid; d < Dv; d += NT) z = fmaf(ctxs[d], __ldg(m.Wt + d), z);
    for (int u = tid; u < m.U; u += NT)
      z = fmaf(__ldcg(m.pq + (size_t)u * m.pq_su + (size_t)b * m.pq_sb), __ldg(m.Wt + Dv + u), z);
    z = block_sum(z, s32);
    if (tid == 0) m.st_u[b] = sigmoidf_(z + m.bt);
  }
  group_sync<GT>(gid);
  stamp();
}

// The same, left to the compiler's inlining heuristics (the instantiation of the kernel without row retirement sits on
// the 128-register limit and spills less this way; the one with it would otherwise keep the AttMech copy in local memory).
template <int GT>
__device__ inline void attention_item(const AttMech& m, int b, int t, int T, int len, float* sm, int gid = 0, int crank = 0,
                                      int csize = 1, bool preloaded = false) {
  attention_item_inl<GT>(m, b, t, T, len, sm, gid, crank, csize, preloaded);
}

}  // namespace dec
}  // namespace sat
