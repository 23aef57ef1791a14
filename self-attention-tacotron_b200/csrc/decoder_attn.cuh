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
#include "../../include/sat.h"

namespace sat {
namespace dec {

constexpr int HALO = 32;         // zero halo each side of the previous alignments (attention_kernel <= 64)
constexpr int MAX_T = 1024;      // padded memory length bound of the row phase
constexpr int ATT_SMEM_FLOATS = 11 * MAX_T + 2 * HALO + NT + 64;

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
  float* ctx_dst[2];    // context element (d, b) written at ctx_dst[i][d * ctx_sd + b * ctx_sb]; may be null
  int ctx_sd, ctx_sb;
};

__device__ __forceinline__ float fast_tanh(float x) {
  // 1 - 2 / (exp(2x) + 1); abs error ~1e-6 (ex2.approx + fast divide), saturates cleanly
  const float e = __expf(2.f * x);
  return 1.f - __fdividef(2.f, e + 1.f);
}

// F_T: 0 = no location term (additive); 5 / 8 = location filters held in registers
// (8 = zero-padded for 6..8); -1 = generic (any F, location features recomputed per row).
template <int F_T>
__device__ void attention_energies(const AttMech& m, int b, int T, int len, const float* aprev,
                                   const float* fbuf, float* e) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NCH = 2;
  constexpr int FR = F_T > 0 ? F_T : 1;
  float pqb[NCH][4], vv[NCH][4], wl[FR][NCH][4];
#pragma unroll
  for (int ch = 0; ch < NCH; ++ch) {
    const int u0 = ch * 128 + 4 * lane;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int u = u0 + i;
      const bool ok = u < m.Up;
      pqb[ch][i] = (ok && u < m.U) ? __ldcg(m.pq + (size_t)u * m.pq_su + (size_t)b * m.pq_sb) + __ldg(m.ba + u) : 0.f;
      vv[ch][i] = ok ? __ldg(m.v + u) : 0.f;
      if (F_T > 0) {
#pragma unroll
        for (int c = 0; c < FR; ++c) wl[c][ch][i] = (ok && c < m.F) ? __ldg(m.Wl + c * m.Up + u) : 0.f;
      }
    }
  }
  const float* kbase = m.keys + (size_t)b * T * m.Up;
  float4 kn[NCH];
  auto load_keys = [&](int j) {
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) {
      const int u0 = ch * 128 + 4 * lane;
      kn[ch] = (j < len && u0 < m.Up) ? __ldg((const float4*)(kbase + (size_t)j * m.Up + u0)) : make_float4(0, 0, 0, 0);
    }
  };
  load_keys(warp);
  for (int j = warp; j < T; j += NW) {
    float4 kc[NCH];
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) kc[ch] = kn[ch];
    if (j + NW < T) load_keys(j + NW);
    float s = 0.f;
    if (j < len) {
      float fj[FR];
      if (F_T > 0) {
#pragma unroll
        for (int c = 0; c < FR; ++c) fj[c] = fbuf[j * 8 + c];
      }
      float fc = 0.f;
      if (F_T < 0 && lane < m.F) {
        // generic: recompute the location features of row j, lane c holds f[j][c] (F <= 32)
        fc = __ldg(m.conv_b + lane);
        for (int i = 0; i < m.Kc; ++i) fc = fmaf(__ldg(m.conv_w + i * m.F + lane), aprev[HALO + j + i - m.pad_left], fc);
      }
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        const int u0 = ch * 128 + 4 * lane;
        const bool ok = u0 < m.Up;  // lane-divergent: only guards loads/accumulation, never a shuffle
        float x[4] = {kc[ch].x + pqb[ch][0], kc[ch].y + pqb[ch][1], kc[ch].z + pqb[ch][2], kc[ch].w + pqb[ch][3]};
        if (F_T > 0) {
#pragma unroll
          for (int c = 0; c < FR; ++c)
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = fmaf(fj[c], wl[c][ch][i], x[i]);
        } else if (F_T < 0) {
          for (int c = 0; c < m.F; ++c) {
            const float f = __shfl_sync(0xffffffffu, fc, c);
            if (ok) {
#pragma unroll
              for (int i = 0; i < 4; ++i) x[i] = fmaf(f, __ldg(m.Wl + c * m.Up + u0 + i), x[i]);
            }
          }
        }
        if (ok) {
#pragma unroll
          for (int i = 0; i < 4; ++i) s = fmaf(vv[ch][i], fast_tanh(x[i]), s);
        }
      }
    }
    s = warp_sum(s);
    if (lane == 0) e[j] = (j < len) ? s : -INFINITY;
  }
}

// One (row, mechanism) item.  `sm` = ATT_SMEM_FLOATS floats of shared memory.  t = decoder step.
__device__ inline void attention_item(const AttMech& m, int b, int t, int T, int len, float* sm) {
  float* e = sm;                         // [MAX_T] energies -> softmax probs a
  float* aprev = e + MAX_T;              // [MAX_T + 2*HALO]
  float* alpha = aprev + MAX_T + 2 * HALO;  // [MAX_T]
  float* fbuf = alpha + MAX_T;           // [MAX_T * 8] location features; later scratch
  float* part = fbuf + 8 * MAX_T;        // [NT] context partials
  float* s32 = part + NT;                // [32] block-reduce scratch
  const int tid = threadIdx.x;
  const bool loc = (m.kind == SAT_ATT_FORWARD || m.kind == SAT_ATT_LOCATION_SENSITIVE);
  const bool fwd = (m.kind == SAT_ATT_FORWARD);
  float* align = e;  // the alignments handed to the context / history

  if (m.kind == SAT_ATT_TEACHER_FORCING) {
    // index = prev_index + 1 = t ; alignments = teacher_alignments[:, index]
    const int ti = min(t, m.teacher_S - 1);
    for (int j = tid; j < T; j += NT) e[j] = __ldg(m.teacher + ((size_t)b * m.teacher_S + ti) * T + j);
    __syncthreads();
  } else {
    for (int j = tid; j < T + 2 * HALO; j += NT) {
      const int jj = j - HALO;
      aprev[j] = (loc && jj >= 0 && jj < T) ? __ldcg(m.st_align + (size_t)b * T + jj) : 0.f;
    }
    if (fwd)
      for (int j = tid; j < T; j += NT) alpha[j] = __ldcg(m.st_alpha + (size_t)b * T + j);
    __syncthreads();
    if (loc && m.F <= 8) {
      for (int i = tid; i < T * 8; i += NT) {
        const int j = i >> 3, c = i & 7;
        float f = 0.f;
        if (c < m.F) {
          f = __ldg(m.conv_b + c);
          for (int k = 0; k < m.Kc; ++k) f = fmaf(__ldg(m.conv_w + k * m.F + c), aprev[HALO + j + k - m.pad_left], f);
        }
        fbuf[i] = f;
      }
      __syncthreads();
    }
    if (!loc) attention_energies<0>(m, b, T, len, aprev, fbuf, e);
    else if (m.F == 5) attention_energies<5>(m, b, T, len, aprev, fbuf, e);
    else if (m.F <= 8) attention_energies<8>(m, b, T, len, aprev, fbuf, e);
    else attention_energies<-1>(m, b, T, len, aprev, fbuf, e);
    __syncthreads();
    // masked softmax (TF: softmax(where(j < len, score, -inf)))
    float mx = -INFINITY;
    for (int j = tid; j < T; j += NT) mx = fmaxf(mx, e[j]);
    mx = block_max(mx, s32);
    float sum = 0.f;
    for (int j = tid; j < T; j += NT) {
      const float ex = (j < len) ? expf(e[j] - mx) : 0.f;
      e[j] = ex;
      sum += ex;
    }
    sum = block_sum(sum, s32);
    for (int j = tid; j < T; j += NT) e[j] = e[j] / sum;
    __syncthreads();
    if (fwd) {
      // alpha' = ((1-u) alpha + u shift(alpha) + 1e-7) * a ; alpha = alpha' / sum(alpha')
      const float u = __ldcg(m.st_u + b);
      float* nal = fbuf;
      float s2 = 0.f;
      for (int j = tid; j < T; j += NT) {
        const float sh = j > 0 ? alpha[j - 1] : 0.f;
        const float val = (((1.f - u) * alpha[j] + u * sh) + 1e-7f) * e[j];
        nal[j] = val;
        s2 += val;
      }
      s2 = block_sum(s2, s32);
      for (int j = tid; j < T; j += NT) nal[j] = nal[j] / s2;
      __syncthreads();
      align = nal;
    }
    // next state
    for (int j = tid; j < T; j += NT) {
      m.st_align[(size_t)b * T + j] = m.cumulative ? e[j] + aprev[HALO + j] : e[j];
      if (fwd) m.st_alpha[(size_t)b * T + j] = align[j];
    }
  }
  if (m.out_align)
    for (int j = tid; j < T; j += NT) m.out_align[((size_t)b * m.out_S + t) * T + j] = align[j];

  // context = alignments @ values  (threads: d fastest, groups split j)
  const int Dv = m.Dv;
  const int G = max(1, NT / Dv);
  const int d = tid % Dv, grp = tid / Dv;
  float c = 0.f;
  if (grp < G && tid < G * Dv) {
    const float* vb = m.values + (size_t)b * T * Dv + d;
#pragma unroll 4
    for (int j = grp; j < len; j += G) c = fmaf(align[j], __ldg(vb + (size_t)j * Dv), c);
  }
  part[tid] = c;
  __syncthreads();
  float ctx = 0.f;
  if (tid < Dv) {
    for (int g = 0; g < G; ++g) ctx += part[g * Dv + tid];
#pragma unroll
    for (int i = 0; i < 2; ++i)
      if (m.ctx_dst[i]) m.ctx_dst[i][(size_t)tid * m.ctx_sd + (size_t)b * m.ctx_sb] = ctx;
  }
  if (fwd && m.agent) {
    // u = sigmoid([context, processed_query] . Wt + bt)   forward_attention.py:111-114
    float z = (tid < Dv) ? ctx * __ldg(m.Wt + tid) : 0.f;
    for (int u = tid; u < m.U; u += NT)
      z = fmaf(__ldcg(m.pq + (size_t)u * m.pq_su + (size_t)b * m.pq_sb), __ldg(m.Wt + Dv + u), z);
    z = block_sum(z, s32);
    if (tid == 0) m.st_u[b] = sigmoidf_(z + m.bt);
  }
  __syncthreads();
}

}  // namespace dec
}  // namespace sat
