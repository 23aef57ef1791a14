// decoder_kernel.cuh -- the persistent autoregressive decoder kernel.
//
// One cooperative launch runs ALL decoder steps.  Per step:
//   C  attention ZoneoutLSTM                                     weight-stationary LSTM tiles
//   D+E query layers of the attention mechanisms, then energies, softmax, alpha recursion, context
//      -- all local to an utterance row                          row split, no barrier inside
//   F  decoder ZoneoutLSTM 1                                     weight-stationary LSTM tiles
//   G  decoder ZoneoutLSTM 2                                     weight-stationary LSTM tiles
//   H-K decoder self-attention hops with a KV cache (transformer decoder only)
//   L  frame + stop-token projection, fed-back frame, prenet of the NEXT step (MultiSpeakerPreNet included),
//      stop rule                                                 row split, no barrier inside
// with one grid-wide barrier between consecutive phases.  The prenet of the <GO> frame runs before the loop.
#pragma once
#include "decoder_attn.cuh"
#include "decoder_tc.cuh"

namespace sat {
namespace dec {

constexpr int N_PROF = 48;  // [0,16) phases of a step, [16,24) attention-item stages, [24,27) tc, [32,40) frame-item stages (CTA 0)
constexpr int SCRATCH_FLOATS = (RED_FLOATS > ATT_SMEM_FLOATS ? RED_FLOATS : ATT_SMEM_FLOATS);

struct SaHop {
  DenseDesc qkv, o, tr;
  float* kcache;  // [B][H][Smax][dh]
  float* vcache;
  float* probs;   // [B][H][Smax][Smax] attention probabilities (row t written at step t) or null
};

struct DecParams {
  int B, Bp, ld, R, T;
  int M, r, nf, Mnf, Mr, P0, P1, A, Dd, Dctx, n_mech, hops, heads, Ds, ms;
  int feed_mode, stop_mode, max_steps, min_iters, S_target, Smax;
  float zc, zh;
  DenseDesc pre0, msd, pre1, lstm_a, lstm_1, lstm_2, outp;
  DenseDesc dq[2];
  AttMech mech[2];
  SaHop hop[MAX_HOPS];
  float *xin, *p0, *p0b, *att_in[2], *l1_in[2], *l2_in[2], *c_a, *c_1, *c_2, *dout;
  float *sa_q, *sa_ctx, *sa_a, *sa_y[2];
  const float* spk_term;  // [P0][ld] softsign(Ws spk + bs), feature-major
  const float* target;    // [B][S_target][Mr]
  const long long* lengths;
  float *mel, *stop, *dec_out;
  int *stop_steps, *n_steps;
  unsigned* barrier;
  long long* prof;  // [N_PROF] accumulated SM cycles per phase (CTA 0), may be null
  // ---- tensor-core mode (B >= 64): LSTM inputs att_in / l1_in / l2_in are fp16 in operand-tile order
  // [RB row blocks][K*_pad / 64 chunks][128 rows][64 k] (tc_act_index) ----
  int tc, RB, Ka_pad, K1_pad, K2_pad;
  int tc_G, tc_stages;      // 64-k chunks per bulk copy and copies in flight of the gate GEMMs' activation ring
  // ---- teacher-forced time-batched form (feed_mode teacher + fixed step count): everything that does not feed the
  // recurrence leaves the loop.  The prenet of all (known) input frames is precomputed into p1_all[S][B][p1_ld]; the
  // loop runs C, D+E, F, G only and records the LSTM2 outputs in dec_out; the decoder self-attention hops (as ONE
  // causal batched attention per hop) and the frame / stop projections run afterwards as time-batched GEMMs --
  // the structure of the reference's own training branch (modules/module.py:743-765).
  int batched_tail;
  const float* p1_all;
  int p1_ld;
  int bench_barriers;       // diagnostic: time this many empty grid barriers before the loop (prof[30])
  int scratch_floats;       // phase scratch at the head of dynamic smem (>= SCRATCH_FLOATS; holds the TMA ring in tc mode)
  TcLstmDesc tc_a, tc_1, tc_2;
  float* qfm;               // feature-major copy of the attention-LSTM output (query) for the SIMT query layers
  int fused_prenet0;        // 1: `outp` carries P0 extra columns = prenet layer 0 applied to the fed-back frame, folded
                            //    into the projection (self feed only): column Mr + 1 + j = sum_f W_out[:, f0 + f] W_pre0[f][j]
  int att_pair;             // 1: a CTA that owns two rows of the attention phase runs them concurrently, one per half
  int att_row_floats;       // shared-memory floats of one attention row (second row of a pair starts there)
  // ---- cluster-split attention rows (small batches: B * n_mech * cl <= grid; launched with cluster dimension cl) ----
  // The cl CTAs of a cluster share one (row, mechanism) item: each computes Up / cl columns of the query layer, the
  // energies of those units (partials all-gathered through distributed shared memory) and Dv / cl context columns.
  int retire;               // per-row stop rule: stopped rows leave the row-split phases (s_live in the kernel)
  int* finished;            // optional (sat_decoder_step): 1 if the stop rule ended the run inside this launch, else 0
  int t_begin, t_end;       // decoder steps [t_begin, t_end) of the run (sat_decoder_step: one step per launch, the state
                            // of the fp32 form lives in the workspace between launches); normally [0, max_steps)
  int cl;                   // cluster size of the launch (1: rows are not split)
  int cl_split[2];          // mechanism i is split over the cluster (else rank 0 takes the whole item)
  const float* cl_wq[2];    // query-layer weights as cl contiguous column slices [cl][A][cl_sw] (fp32)
  int cl_sw[2];             // columns per slice = Up / cl
  int cl_att_off;           // float offset of the attention row's shared memory (behind the query-layer arrays, which a
                            // slower CTA of the cluster may still be using when remote partial energies arrive)
};

struct DenseEpi {
  const float* bias; int N, ld, act;
  Act2D dst; float* dst2; const float* addend;  // dst2 / addend feature-major [n][ld]
  __device__ __forceinline__ void operator()(int b, int tile, int ul, const float* g) const {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int n = tile * TILE_N + ul * 4 + c;
      if (n >= N) continue;
      float v = g[c] + __ldg(bias + n);
      if (act == 1) v = fmaxf(v, 0.f);
      else if (act == 2) v = tanhf(v);
      if (addend) v += __ldcg(addend + (size_t)n * ld + b);
      dst.put(n, b, v);
      if (dst2) dst2[(size_t)n * ld + b] = v;
    }
  }
};

struct LstmEpi {
  const float* bias; int units, ld, B;
  float* c; const float* hprev; float* out; float* hz;
  float* rm_out; int rm_stride; int rm_off;  // optional row-major copy of the output: rm_out[b*rm_stride + rm_off + u]
  float zc, zh;
  __device__ __forceinline__ void operator()(int b, int tile, int ul, const float* g) const {
    const int u = tile * 2 + ul;
    if (u >= units) return;
    const float* bb = bias + tile * TILE_N + ul * 4;
    const float gi = g[0] + __ldg(bb + 0), gj = g[1] + __ldg(bb + 1), gf = g[2] + __ldg(bb + 2), go = g[3] + __ldg(bb + 3);
    const size_t idx = (size_t)u * ld + b;
    const float cp = c[idx], hp = __ldcg(hprev + idx);
    const float cn = sigmoid_fast(gf + 1.0f) * cp + sigmoid_fast(gi) * tanh_fast(gj);
    const float hn = sigmoid_fast(go) * tanh_fast(cn);
    c[idx] = cn - zc * (cn - cp);
    hz[idx] = hn - zh * (hn - hp);
    out[idx] = hn;
    if (rm_out && b < B) rm_out[(size_t)b * rm_stride + rm_off + u] = hn;
  }
};

struct QkvEpi {
  const float* bias; int Ds, dh, H, Smax, t, ld, B;
  float *kc, *vc, *q;
  __device__ __forceinline__ void operator()(int b, int tile, int ul, const float* g) const {
    const int n0 = tile * TILE_N + ul * 4;
    if (n0 >= 3 * Ds) return;
    float v[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) v[c] = g[c] + __ldg(bias + n0 + c);
    const int which = n0 / Ds, d = n0 % Ds;
    if (which == 2) {
#pragma unroll
      for (int c = 0; c < 4; ++c) q[(size_t)(d + c) * ld + b] = v[c];
    } else if (b < B) {
      const int hd = d / dh, dd = d % dh;
      float* dst = (which == 0 ? kc : vc) + (((size_t)b * H + hd) * Smax + t) * dh + dd;
      *(float4*)dst = make_float4(v[0], v[1], v[2], v[3]);
    }
  }
};

// decoder self-attention over the KV cache for one (row, head): last-row causal attention.
// Scores: a lane owns one float4 of the head dimension, 32 / (dh/4) cache rows per warp pass, four passes in
// flight; dh = 128 uses the 4-row transpose-reduction.  Context: attention_context_vec (warp per cache row).
__device__ inline void sa_item(const DecParams& p, const SaHop& h, int b, int hd, int t, float* sm) {
  const int dh = p.Ds / p.heads, n = t + 1, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* sc = sm;             // [n] scores -> probabilities (n <= 8192)
  float* q = sm + 8192;       // [dh]
  float* part = q + 512;      // [NW * dh] context partials / [NT] scalar-path partials
  float* ctxs = part + NW * 128;  // [dh] (vector path needs dh <= 128)
  float* s32 = ctxs + 512;
  const float scale = 1.f / sqrtf((float)dh);
  for (int d = tid; d < dh; d += NT) q[d] = __ldcg(p.sa_q + (size_t)(hd * dh + d) * p.ld + b) * scale;
  __syncthreads();
  const float* kb = h.kcache + ((size_t)b * p.heads + hd) * p.Smax * dh;
  const float* vb = h.vcache + ((size_t)b * p.heads + hd) * p.Smax * dh;
  const int Cq = dh >> 2;
  const bool vec = (dh & 3) == 0 && (Cq & (Cq - 1)) == 0 && Cq <= 32;
  if (vec) {
    const int G = 32 / Cq, c0 = lane % Cq, g = lane / Cq;
    const float4 q4 = *(const float4*)(q + 4 * c0);
    const int stride = NW * G, last = n - 1;
    const float* kl = kb + 4 * c0;
    auto rowp = [&](int r) { return (const float4*)(kl + (size_t)min(warp * G + g + stride * r, last) * dh); };
    float4 kn[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) kn[s] = __ldcg(rowp(s));
    for (int r0 = 0; warp * G + stride * r0 < n; r0 += 4) {
      float sum[4];
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const float4 k4 = kn[s];
        kn[s] = __ldcg(rowp(r0 + 4 + s));
        sum[s] = fmaf(q4.x, k4.x, fmaf(q4.y, k4.y, fmaf(q4.z, k4.z, q4.w * k4.w)));
      }
      if (Cq == 32) {
        const bool b4 = lane & 16, b3 = lane & 8;
        float a0 = b4 ? sum[2] : sum[0], a1 = b4 ? sum[3] : sum[1];
        a0 += __shfl_xor_sync(0xffffffffu, b4 ? sum[0] : sum[2], 16);
        a1 += __shfl_xor_sync(0xffffffffu, b4 ? sum[1] : sum[3], 16);
        float c = b3 ? a1 : a0;
        c += __shfl_xor_sync(0xffffffffu, b3 ? a0 : a1, 8);
        c += __shfl_xor_sync(0xffffffffu, c, 4);
        c += __shfl_xor_sync(0xffffffffu, c, 2);
        c += __shfl_xor_sync(0xffffffffu, c, 1);
        const int jq = warp + stride * (r0 + (lane >> 3));
        if ((lane & 7) == 0 && jq < n) sc[jq] = c;
      } else {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          float c = sum[s];
          for (int o = 1; o < Cq; o <<= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
          const int jq = warp * G + g + stride * (r0 + s);
          if (c0 == 0 && jq < n) sc[jq] = c;
        }
      }
    }
  } else {
    for (int j = warp; j < n; j += NW) {
      float s = 0.f;
      for (int d = lane; d < dh; d += 32) s = fmaf(q[d], __ldcg(kb + (size_t)j * dh + d), s);
      s = warp_sum(s);
      if (lane == 0) sc[j] = s;
    }
  }
  __syncthreads();
  float mx = -INFINITY;
  for (int j = tid; j < n; j += NT) mx = fmaxf(mx, sc[j]);
  mx = block_max(mx, s32);
  float sum = 0.f;
  for (int j = tid; j < n; j += NT) {
    const float e = expf(sc[j] - mx);
    sc[j] = e;
    sum += e;
  }
  sum = block_sum(sum, s32);
  const float inv = 1.f / sum;
  __syncthreads();
  if (h.probs) {
    float* pr = h.probs + (((size_t)b * p.heads + hd) * p.Smax + t) * p.Smax;
    for (int j = tid; j < p.Smax; j += NT) pr[j] = j < n ? sc[j] * inv : 0.f;
  }
  if (vec && ((((size_t)vb) & 15) == 0)) {
    attention_context_vec<1>(vb, sc, n, dh, part, ctxs);
    for (int d = tid; d < dh; d += NT) p.sa_ctx[(size_t)(hd * dh + d) * p.ld + b] = ctxs[d] * inv;
  } else {
    const int G = max(1, NT / dh), d = tid % dh, grp = tid / dh;
    float c = 0.f;
    if (tid < G * dh)
      for (int j = grp; j < n; j += G) c = fmaf(sc[j], __ldcg(vb + (size_t)j * dh + d), c);
    part[tid] = c;
    __syncthreads();
    if (tid < dh) {
      float ctx = 0.f;
      for (int g = 0; g < G; ++g) ctx += part[g * dh + tid];
      p.sa_ctx[(size_t)(hd * dh + tid) * p.ld + b] = ctx * inv;
    }
  }
  __syncthreads();
}

// ---- row-split frame phase -------------------------------------------------------------------------------
// Everything between the last recurrent layer of step t and the attention LSTM of step t+1 is local to an utterance
// row: frame + stop projection, the fed-back frame, and the prenet.  One CTA takes (up to) two rows through all of
// it without a grid barrier in between; the weights (row-major [K][Np]) stream from L2 with 16-byte loads.

// outs[r][n] = act(bias[n] + sum_k xs[r][k] W[k][n]) for RR rows held in shared memory.
// H = false: fp32 weights [K][Np], a lane owns 4 columns; H = true: fp16 weights [K][Np8] (tensor-core mode, where
// the LSTM operands are fp16 as well), a lane owns 8 columns -- half the L2 -> SM bytes, which bound this phase.
template <int RR, bool H>
__device__ inline void row_dense(const DenseDesc& d, const float* xs, int xs_stride, float* part, float* outs, int act) {
  constexpr int CW = H ? 8 : 4;  // columns per lane
  const int tid = threadIdx.x;
  const int Np = H ? d.Np8 : d.Np, Cn = Np / CW, K = d.K;
  const int kparts = min(NT / Cn, K), kper = (K + kparts - 1) / kparts;
  const int kp = tid / Cn, cl = tid - kp * Cn;
  if (kp < kparts) {
    float acc[RR][CW];
#pragma unroll
    for (int r = 0; r < RR; ++r)
#pragma unroll
      for (int c = 0; c < CW; ++c) acc[r][c] = 0.f;
    const int k0 = kp * kper, k1 = min(K, k0 + kper);
    const int rep = d.reps > 1 ? (int)(blockIdx.x % d.reps) : 0;
    const uint4* w = (H ? (const uint4*)((const __half*)d.Wrm16 + rep * d.rep_stride16)
                        : (const uint4*)(d.Wrm + rep * d.rep_stride)) + (size_t)k0 * Cn + cl;
    for (int k = k0; k < k1; k += 8, w += 8 * Cn) {
      uint4 wv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) wv[j] = (k + j < k1) ? __ldg(w + (size_t)j * Cn) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float wf[CW];
        if (H) {
          const float2 a = __half22float2(*(const __half2*)&wv[j].x), b = __half22float2(*(const __half2*)&wv[j].y);
          const float2 c = __half22float2(*(const __half2*)&wv[j].z), e = __half22float2(*(const __half2*)&wv[j].w);
          wf[0] = a.x; wf[1] = a.y; wf[2] = b.x; wf[3] = b.y;
          if (CW == 8) { wf[4 % CW] = c.x; wf[5 % CW] = c.y; wf[6 % CW] = e.x; wf[7 % CW] = e.y; }
        } else {
          wf[0] = __uint_as_float(wv[j].x); wf[1] = __uint_as_float(wv[j].y);
          wf[2] = __uint_as_float(wv[j].z); wf[3] = __uint_as_float(wv[j].w);
        }
#pragma unroll
        for (int r = 0; r < RR; ++r) {
          const float x = (k + j < k1) ? xs[r * xs_stride + k + j] : 0.f;
#pragma unroll
          for (int c = 0; c < CW; ++c) acc[r][c] = fmaf(x, wf[c], acc[r][c]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < RR; ++r)
#pragma unroll
      for (int c = 0; c < CW; c += 4)
        *(float4*)(part + (size_t)(kp * RR + r) * Np + CW * cl + c) =
            make_float4(acc[r][c], acc[r][c + 1], acc[r][c + 2], acc[r][c + 3]);
  }
  __syncthreads();
  const int No = d.Np;  // output row stride (bias is zero-padded to n_tiles * 8 >= Np)
  for (int i = tid; i < RR * No; i += NT) {
    const int r = i / No, n = i - r * No;
    float v = 0.f;
    for (int q = 0; q < kparts; ++q) v += part[(size_t)(q * RR + r) * Np + n];
    v += __ldg(d.bias + n);
    if (act == 1) v = fmaxf(v, 0.f);
    outs[r * No + n] = v;
  }
  __syncthreads();
}

// shared-memory floats the frame phase needs (checked against the phase scratch on the host)
__host__ __device__ inline int al4(int x) { return (x + 3) & ~3; }
__host__ __device__ inline int frame_smem_floats(int Dy, int Np_out, int Mnf, int P0p, int P1p) {
  return al4(2 * Dy) + 2 * Np_out + al4(2 * Mnf) + 4 * P0p + 2 * P1p + 2 * NT * 8 + 64;
}

// rows b[0..RR): (unless `prologue`) project y -> frame + stop token of step t, form the next input frame, and run
// the prenet on it into the attention-LSTM input buffer of step t+1
template <int RR, bool H, bool PROF, bool SPEC = false>
__device__ inline void frame_item(const DecParams& p, const int* b, int t, const float* y, int Dy, Act2D p1_dst,
                                  bool prologue, float* sm) {
  const int tid = threadIdx.x, ld = p.ld;
  const int Npo = p.outp.Np, P0p = p.pre0.Np, P1p = p.pre1.Np;
  float* ys = sm;                   // [RR][Dy]        (every region starts 16-byte aligned)
  float* os = ys + al4(2 * Dy);     // [RR][Npo] projection outputs
  float* fr = os + 2 * Npo;         // [RR][Mnf] next input frame
  float* p0s = fr + al4(2 * p.Mnf); // [RR][P0p]
  float* p0b = p0s + 2 * P0p;       // [RR][P0p] (multi-speaker second dense)
  float* p1s = p0b + 2 * P0p;       // [RR][P1p]
  float* s32 = p1s + 2 * P1p;       // [64]
  float* part = s32 + 64;           // [2 * NT * 8]
  long long tp = 0;
  int ps = 32;
  const bool profiling = PROF && p.prof && blockIdx.x == 0 && tid == 0 && !prologue;
  if (profiling) tp = clock64();
  auto stamp = [&]() {
    if (profiling) {
      const long long now = clock64();
      p.prof[ps] += now - tp;
      tp = now;
    }
    ++ps;
  };
  if (!prologue) {
    for (int i = tid; i < RR * Dy; i += NT) {
      const int r = i / Dy, k = i - r * Dy;
      ys[r * Dy + k] = __ldcg(y + (size_t)k * ld + b[r]);
    }
    __syncthreads();
    stamp();
    row_dense<RR, H>(p.outp, ys, Dy, part, os, 0);
    stamp();
    for (int i = tid; i < RR * (p.Mr + 1); i += NT) {
      const int r = i / (p.Mr + 1), n = i - r * (p.Mr + 1);
      const float v = os[r * Npo + n];
      if (n == p.Mr) p.stop[(size_t)b[r] * p.max_steps + t] = v;
      else p.mel[((size_t)b[r] * p.max_steps + t) * p.Mr + n] = v;
    }
    const int f0 = p.Mr - p.Mnf;
    if (SPEC || p.fused_prenet0) {
      // self feed with prenet layer 0 folded into the projection: its pre-activations are columns Mr + 1 .. of `os`
      for (int i = tid; i < RR * p.P0; i += NT) {
        const int r = i / p.P0, n = i - r * p.P0;
        p0s[r * P0p + n] = fmaxf(os[r * Npo + p.Mr + 1 + n], 0.f);
      }
    } else if (p.feed_mode == SAT_FEED_TEACHER) {
      const int ts = min(t, p.S_target - 1);
      for (int i = tid; i < RR * p.Mnf; i += NT) {
        const int r = i / p.Mnf, f = i - r * p.Mnf;
        fr[r * p.Mnf + f] = __ldg(p.target + ((size_t)b[r] * p.S_target + ts) * p.Mr + f0 + f);
      }
    } else if (p.feed_mode == SAT_FEED_SELF) {
      for (int i = tid; i < RR * p.Mnf; i += NT) {
        const int r = i / p.Mnf, f = i - r * p.Mnf;
        fr[r * p.Mnf + f] = os[r * Npo + f0 + f];
      }
    } else {
      // softmax feedback of the code decoders (modules/helpers.py:98-106): per fed-back frame over its M bins
      for (int r = 0; r < RR; ++r)
        for (int s = 0; s < p.nf; ++s) {
          const float* row = os + r * Npo + f0 + s * p.M;
          float mx = -INFINITY;
          for (int i = tid; i < p.M; i += NT) mx = fmaxf(mx, row[i]);
          mx = block_max(mx, s32);
          float sum = 0.f;
          for (int i = tid; i < p.M; i += NT) sum += expf(row[i] - mx);
          sum = block_sum(sum, s32);
          for (int i = tid; i < p.M; i += NT) fr[r * p.Mnf + s * p.M + i] = expf(row[i] - mx) / sum;
          __syncthreads();
        }
    }
  } else {
    for (int i = tid; i < RR * p.Mnf; i += NT) fr[i] = 0.f;  // <GO> frame
  }
  __syncthreads();
  ps = 34;
  stamp();
  if (prologue || !(SPEC || p.fused_prenet0)) row_dense<RR, H>(p.pre0, fr, p.Mnf, part, p0s, 1);
  stamp();
  const float* pin = p0s;
  if (!SPEC && p.ms) {
    // MultiSpeakerPreNet: relu(dense0(x)) + softsign(speaker projection), then a second relu dense
    for (int i = tid; i < RR * p.P0; i += NT) {
      const int r = i / p.P0, n = i - r * p.P0;
      p0s[r * P0p + n] += __ldg(p.spk_term + (size_t)n * ld + b[r]);
    }
    __syncthreads();
    row_dense<RR, H>(p.msd, p0s, P0p, part, p0b, 1);
    pin = p0b;
  }
  row_dense<RR, H>(p.pre1, pin, P0p, part, p1s, 1);
  ps = 36;
  stamp();
  for (int i = tid; i < RR * p.P1; i += NT) {
    const int r = i / p.P1, n = i - r * p.P1;
    p1_dst.put(n, b[r], p1s[r * P1p + n]);
  }
  __syncthreads();
  stamp();
}

// all rows of the batch through frame_item, two per pass
// live[0 .. n_live): the rows that still decode (RETIRE: the per-row stop rule has retired the others; else all rows)
template <bool RETIRE, bool PROF, bool SPEC>
__device__ inline void frame_phase(const DecParams& p, int t, const float* y, int Dy, Act2D p1_dst, bool prologue, float* sm,
                                   const short* live, int n_live) {
  for (int l0 = blockIdx.x; l0 < n_live; l0 += 2 * gridDim.x) {
    const int l1 = l0 + (int)gridDim.x;
    const bool two = l1 < n_live;
    int bl[2] = {l0, l1};
    const int* b = bl;
    if (RETIRE) {
      // the live rows' indices stay in shared memory (re-read where needed): as two more registers through the whole item
      // they cost the dense layers of this instantiation 17 % at the 128-register limit
      __shared__ int s_fb[2];
      __syncthreads();
      if (threadIdx.x == 0) {
        s_fb[0] = live[l0];
        s_fb[1] = two ? (int)live[l1] : 0;
      }
      __syncthreads();
      b = s_fb;
    }
    if (SPEC || p.tc) {
      if (two) frame_item<2, true, PROF, SPEC>(p, b, t, y, Dy, p1_dst, prologue, sm);
      else frame_item<1, true, PROF, SPEC>(p, b, t, y, Dy, p1_dst, prologue, sm);
    } else {
      if (two) frame_item<2, false, PROF>(p, b, t, y, Dy, p1_dst, prologue, sm);
      else frame_item<1, false, PROF>(p, b, t, y, Dy, p1_dst, prologue, sm);
    }
  }
}

// RETIRE: the per-row stop rule retires stopped rows from the row-split phases (a separate instantiation, so that the
// indirection through the live-row list costs the other stop rules nothing)
// PROF: the per-phase cycle counters of CTA 0 (sat_decoder_profile).  A separate instantiation: compiled in, they cost the
// plain kernel 2 % even when switched off at run time (registers and code at the 128-register limit, measured by an A/B
// of two builds on one box); the profiling instantiation exists for the plain stop rules only.
// SPEC: the instantiation for the free-running baseline Tacotron decoder at B >= 64 (what bench.py times): tensor-core
// form, one attention mechanism, no cluster-split rows, no KV-cached hops, no time-batched tail, self feed with prenet
// layer 0 folded into the projection, no speaker term.  Those run-time switches become constants and the other forms of
// the phases drop out of the kernel, which at the 128-register limit is worth a few per cent.
template <bool RETIRE, bool PROF = false, bool SPEC = false>
__global__ void __launch_bounds__(NT, 1) decoder_persistent_kernel(const __grid_constant__ DecParams p) {
  long long* const kprof = PROF ? p.prof : nullptr;
  const bool k_tc = SPEC ? true : p.tc != 0;
  const int k_cl = SPEC ? 1 : p.cl, k_hops = SPEC ? 0 : p.hops, k_nmech = SPEC ? 1 : p.n_mech;
  const bool k_tail = SPEC ? false : p.batched_tail != 0;
  extern __shared__ __align__(16) float smem[];
  float* scratch = smem;                  // SCRATCH_FLOATS: reduction / row-phase scratch
  float* wres = smem + p.scratch_floats;  // resident LSTM tiles / tensor-core weight images
  __shared__ int s_stopped[MAX_B];
  // Row retirement (per-row stop rule): s_live[0 .. s_nlive) = the rows that have not stopped, in row order.  The
  // row-split phases (query layer + attention, frame phase) hand out live rows only, so a stopped row costs nothing there
  // and the live ones spread over all CTAs again; the LSTM phases work on 128-row tensor tiles and keep running every
  // row (a retired row's state simply stops mattering).  Frames after a row's own stop step are not written.
  __shared__ short s_live[MAX_B];
  __shared__ int s_nlive, s_cnt[MAX_B / 32];
  GridBarrier bar{p.barrier, 0u};
  const int tid = threadIdx.x;
  const int ld = p.ld;

  __shared__ __align__(8) uint64_t s_tcbar[2 * TC_MAX_STAGES + 1];
  __shared__ uint32_t s_tmem;
  TcState tcs;
  tcs.G = p.tc_G;
  tcs.stages = p.tc_stages;
  tcs.ring = (uint8_t*)(((uintptr_t)scratch + 1023) & ~(uintptr_t)1023);
  tcs.wimg = (uint8_t*)(((uintptr_t)wres + 1023) & ~(uintptr_t)1023);
  tcs.full = s_tcbar;
  tcs.empty = s_tcbar + TC_MAX_STAGES;
  tcs.done = s_tcbar + 2 * TC_MAX_STAGES;
  tcs.tmem = 0;
  tcs.n_chunks = 0;
  tcs.n_items = 0;
  tcs.prefilled = 0;
  tcs.prof = kprof ? kprof + 24 : nullptr;
  if (k_tc) {
    if (tid == 0) {
      for (int i = 0; i < 2 * TC_MAX_STAGES + 1; ++i) tc::mbar_init(s_tcbar + i, 1);
      tc::fence_mbar_init();
    }
    if (tid < 32) tc::tmem_alloc(&s_tmem, 32);
    tc_load_resident(p.tc_a, p.RB, tcs);
    tc_load_resident(p.tc_1, p.RB, tcs);
    tc_load_resident(p.tc_2, p.RB, tcs);
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    tcs.tmem = s_tmem;
  } else {
    load_resident(p.lstm_a, wres);
    load_resident(p.lstm_1, wres);
    load_resident(p.lstm_2, wres);
  }
  for (int i = tid; i < MAX_B; i += NT) {
    s_stopped[i] = (p.t_begin > 0 && i < p.B) ? p.stop_steps[i] : -1;  // a continued run keeps its stop decisions
    s_live[i] = (short)i;
  }
  if (tid == 0) s_nlive = p.B;
  __syncthreads();

  if (p.bench_barriers > 0 && kprof) {
    bar.sync();
    const long long c0 = clock64();
    for (int i = 0; i < p.bench_barriers; ++i) bar.sync();
    if (blockIdx.x == 0 && tid == 0) kprof[30] += clock64() - c0;
  }
  long long tprev = 0;
  int ph = 0;
  // next_lstm / next_x: the tensor-core LSTM phase that follows this barrier (null: none) -- the copies of its "old"
  // input columns are requested while the CTA waits for the other CTAs
  auto phase_end = [&](const TcLstmDesc* next_lstm = nullptr, const float* next_x = nullptr) {
    if (next_lstm && next_lstm->n_pre > 0) {
      tc::fence_proxy_async();  // every thread: its generic-proxy writes to the scratch precede the bulk-copy writes
      bar.sync([&] { tc_prefill(*next_lstm, (const __half*)next_x, p.RB, tcs); });
      tcs.prefilled = blockIdx.x < next_lstm->n_tiles * p.RB ? next_lstm->n_pre : 0;
    } else {
      bar.sync();
    }
    if (kprof && blockIdx.x == 0 && tid == 0 && tprev) {  // (tprev == 0: the barrier in front of the first step)
      const long long now = clock64();
      kprof[ph] += now - tprev;
      tprev = now;
    }
    ++ph;
  };

  // prenet outputs of step `step` (precomputed) -> prenet slot of an attention-LSTM input buffer
  auto copy_prenet = [&](int step, Act2D dst) {
    const float* src = p.p1_all + (size_t)step * p.B * p.p1_ld;
    for (int i = blockIdx.x * NT + tid; i < p.B * p.P1; i += gridDim.x * NT) {
      const int b = i / p.P1, n = i - b * p.P1;
      dst.put(n, b, __ldg(src + (size_t)b * p.p1_ld + n));
    }
  };
  // prenet of the <GO> frame -> attention-LSTM input of step 0
  if (p.t_begin > 0) {
    // continued run: the previous launch left the prenet of this step's input frame in place
  } else if (k_tail) copy_prenet(0, Act2D{p.att_in[0], k_tc ? 0 : ld, k_tc ? p.Ka_pad / 64 : 1, k_tc});
  else frame_phase<RETIRE, PROF, SPEC>(p, 0, nullptr, 0, Act2D{p.att_in[0], k_tc ? 0 : ld, k_tc ? p.Ka_pad / 64 : 1, k_tc}, true, scratch, s_live, p.B);
  auto row_of = [&](int li) { return RETIRE ? (int)s_live[li] : li; };
  if (k_tc) phase_end(&p.tc_a, p.att_in[p.t_begin & 1]);
  else bar.sync();

  int t = p.t_begin, fired = 0;
  for (; t < p.t_end; ++t) {
    const int cur = t & 1, nxt = cur ^ 1;
    ph = 0;
    if (kprof && blockIdx.x == 0 && tid == 0) tprev = clock64();
    // ---- C: attention LSTM ----
    if (k_tc) {
      // fp16 row-major buffers: att_in = [prenet | context | zoned h], l1_in = [h_a | context | zoned h1]
      const int h0 = p.P1 + p.Dctx;
      TcLstmIO io{(const __half*)p.att_in[cur], Act2D{p.l1_in[cur], 0, p.K1_pad / 64, 1}, Act2D{p.qfm, ld, 1, 0},
                  Act2D{p.att_in[nxt], h0, p.Ka_pad / 64, 1}, nullptr, 0, 0};
      tc::fence_proxy_async();
      __syncthreads();
      lstm_phase_tc(p.tc_a, io, p.B, p.Bp, p.RB, p.zc, p.zh, tcs);
    } else {
      const size_t hoff = (size_t)(p.P1 + p.Dctx) * ld;
      dense_phase(p.lstm_a, p.att_in[cur], ld, p.Bp, wres, scratch,
                  LstmEpi{p.lstm_a.bias, p.A, ld, p.B, p.c_a, p.att_in[cur] + hoff, p.l1_in[cur],
                          p.att_in[nxt] + hoff, nullptr, 0, 0, p.zc, p.zh});
    }
    phase_end();
    // ---- D+E: query layers and attention mechanisms (row split, no barrier in between) ----
    // The processed query of a row is a row-local dense on the attention-LSTM output, so the CTA that owns the row's
    // attention item computes it itself (two rows per pass share one stream of the query-layer weights from L2).
    if (k_cl > 1) {
      // cluster-split form (fp32 SIMT mode only): item = cluster index, every CTA of the cluster works on it
      const float* qsrc = p.l1_in[cur];  // feature-major [A][ld]
      const int n_items = (RETIRE ? s_nlive : p.B) * k_nmech, ncl = gridDim.x / k_cl, crank = (int)cluster_ctarank();
      for (int item = blockIdx.x / k_cl; item < n_items; item += ncl) {
        const int mi = item % k_nmech, b = row_of(item / k_nmech);
        const bool split = p.cl_split[mi] != 0;
        AttMech m = p.mech[mi];
        const int coff = mi ? p.mech[0].Dv : 0;
        m.ctx_f16 = 0;
        m.ctx_dst[0] = p.att_in[nxt] + (size_t)(p.P1 + coff) * ld;
        m.ctx_dst[1] = p.l1_in[cur] + (size_t)(p.A + coff) * ld;
        m.ctx_sd[0] = m.ctx_sd[1] = ld;
        m.ctx_sb[0] = m.ctx_sb[1] = 1;
        m.prof = kprof ? kprof + 16 : nullptr;
        if (split || crank == 0) {
          // the row's attention state goes in flight together with the query-layer input
          attention_load_state<NT>(m, b, p.T, scratch + p.cl_att_off);
          if (m.kind != SAT_ATT_TEACHER_FORCING) {
            DenseDesc dq = p.dq[mi];
            int c0 = 0;
            if (split) {
              dq.Wrm = p.cl_wq[mi] + (size_t)crank * dq.K * p.cl_sw[mi];
              dq.Np = dq.N = p.cl_sw[mi];
              dq.reps = 1;
              c0 = crank * p.cl_sw[mi];
            }
            float* xs = scratch;            // [A]
            float* outs = xs + al4(p.A);    // [Np]
            float* part = outs + dq.Np;     // [kparts * Np] <= [NT * 4]
            for (int k = tid; k < p.A; k += NT) xs[k] = __ldcg(qsrc + (size_t)k * ld + b);
            __syncthreads();
            row_dense<1, false>(dq, xs, p.A, part, outs, 0);
            for (int n = tid; n < dq.N; n += NT)
              const_cast<float*>(m.pq)[(size_t)(c0 + n) * m.pq_su + (size_t)b * m.pq_sb] = outs[n];
            __syncthreads();
          } else {
            __syncthreads();
          }
          attention_item<NT>(m, b, t, p.T, (int)p.lengths[b], scratch + p.cl_att_off, 0, split ? crank : 0,
                             split ? k_cl : 1, true);
        }
        // the next item's remote partial energies must not land while a CTA of the cluster still reads this item's
        if (item + ncl < n_items) cluster_sync_all();
      }
    } else {
      const float* qsrc = k_tc ? p.qfm : p.l1_in[cur];  // feature-major [A][ld]
      const int n_items = (RETIRE ? s_nlive : p.B) * k_nmech;  // item i = (live row i / n_mech, mechanism i % n_mech)
      for (int i0 = blockIdx.x; i0 < n_items; i0 += 2 * gridDim.x) {
        const int i1 = i0 + gridDim.x;
        const int nit = i1 < n_items ? 2 : 1;
        const int items[2] = {i0, i1};
        const int mi0 = i0 % k_nmech;
        // the items of one CTA share the mechanism whenever gridDim.x is a multiple of n_mech; otherwise one at a time
        const bool pair = nit == 2 && (i1 % k_nmech) == mi0;
        for (int q0 = 0; q0 < nit; q0 += pair ? 2 : 1) {
          const int nq = pair ? 2 : 1;
          const int mi = items[q0] % k_nmech;
          if (p.mech[mi].kind == SAT_ATT_TEACHER_FORCING) continue;
          const DenseDesc& dq = p.dq[mi];
          float* xs = scratch;                 // [2][A]
          float* outs = xs + al4(2 * p.A);     // [2][Np]
          float* part = outs + 2 * dq.Np;      // [2 * NT * 8]
          int rows[2];
          for (int r = 0; r < nq; ++r) rows[r] = row_of(items[q0 + r] / k_nmech);
          for (int i = tid; i < nq * p.A; i += NT) {
            const int r = i / p.A, k = i - r * p.A;
            xs[r * p.A + k] = __ldcg(qsrc + (size_t)k * ld + rows[r]);
          }
          __syncthreads();
          if (k_tc) {
            if (nq == 2) row_dense<2, true>(dq, xs, p.A, part, outs, 0);
            else row_dense<1, true>(dq, xs, p.A, part, outs, 0);
          } else {
            if (nq == 2) row_dense<2, false>(dq, xs, p.A, part, outs, 0);
            else row_dense<1, false>(dq, xs, p.A, part, outs, 0);
          }
          const AttMech& mm = p.mech[mi];
          for (int i = tid; i < nq * dq.N; i += NT) {
            const int r = i / dq.N, n = i - r * dq.N;
            const_cast<float*>(mm.pq)[(size_t)n * mm.pq_su + (size_t)rows[r] * mm.pq_sb] = outs[r * dq.Np + n];
          }
          __syncthreads();
        }
        auto setup = [&](int item, AttMech& m) {
          const int mi = item % k_nmech;
          m = p.mech[mi];
          const int coff = mi ? p.mech[0].Dv : 0;
          if (k_tc) {
            // fp16 operand-tile order (tc_act_index): ctx_sd = feature offset of the context, ctx_sb = chunks per row
            m.ctx_dst[0] = p.att_in[nxt];
            m.ctx_dst[1] = p.l1_in[cur];
            m.ctx_sd[0] = p.P1 + coff;
            m.ctx_sd[1] = p.A + coff;
            m.ctx_sb[0] = p.Ka_pad / 64;
            m.ctx_sb[1] = p.K1_pad / 64;
            m.ctx_f16 = 1;
          } else {
            m.ctx_f16 = 0;
            m.ctx_dst[0] = p.att_in[nxt] + (size_t)(p.P1 + coff) * ld;
            m.ctx_dst[1] = p.l1_in[cur] + (size_t)(p.A + coff) * ld;
            m.ctx_sd[0] = m.ctx_sd[1] = ld;
            m.ctx_sb[0] = m.ctx_sb[1] = 1;
          }
          m.prof = kprof ? kprof + 16 : nullptr;
        };
        if (p.att_pair && pair && att_row_pairable(p.mech[mi0])) {
          // two rows of the same mechanism: one per half of the CTA, concurrently (named barriers 1 / 2)
          const int gid = tid / (NT / 2);
          AttMech m;
          setup(items[gid], m);
          const int b = row_of(items[gid] / k_nmech);
          if (RETIRE || PROF) attention_item_inl<NT / 2>(m, b, t, p.T, (int)p.lengths[b], scratch + gid * p.att_row_floats, gid);
          else attention_item<NT / 2>(m, b, t, p.T, (int)p.lengths[b], scratch + gid * p.att_row_floats, gid);
          __syncthreads();  // both halves are done with the scratch
        } else {
          for (int q = 0; q < nit; ++q) {
            AttMech m;
            setup(items[q], m);
            const int b = row_of(items[q] / k_nmech);
            if (RETIRE || PROF) attention_item_inl<NT>(m, b, t, p.T, (int)p.lengths[b], scratch);
            else attention_item<NT>(m, b, t, p.T, (int)p.lengths[b], scratch);
          }
        }
      }
    }
    if (k_tc) phase_end(&p.tc_1, p.l1_in[cur]);
    else phase_end();
    // ---- F: decoder LSTM 1 ----
    if (k_tail && t + 1 < p.max_steps)
      copy_prenet(t + 1, Act2D{p.att_in[nxt], k_tc ? 0 : ld, k_tc ? p.Ka_pad / 64 : 1, k_tc});
    if (k_tc) {
      const int h0 = p.A + p.Dctx;
      TcLstmIO io{(const __half*)p.l1_in[cur], Act2D{p.l2_in[cur], 0, p.K2_pad / 64, 1}, Act2D{nullptr, 0, 0, 0},
                  Act2D{p.l1_in[nxt], h0, p.K1_pad / 64, 1}, nullptr, 0, 0};
      tc::fence_proxy_async();
      __syncthreads();
      lstm_phase_tc(p.tc_1, io, p.B, p.Bp, p.RB, p.zc, p.zh, tcs);
    } else {
      const size_t hoff = (size_t)(p.A + p.Dctx) * ld;
      dense_phase(p.lstm_1, p.l1_in[cur], ld, p.Bp, wres, scratch,
                  LstmEpi{p.lstm_1.bias, p.Dd, ld, p.B, p.c_1, p.l1_in[cur] + hoff, p.l2_in[cur],
                          p.l1_in[nxt] + hoff, nullptr, 0, 0, p.zc, p.zh});
    }
    if (k_tc) phase_end(&p.tc_2, p.l2_in[cur]);
    else phase_end();
    // ---- G: decoder LSTM 2 ----
    if (k_tc) {
      const int h0 = p.Dd;
      TcLstmIO io{(const __half*)p.l2_in[cur], Act2D{p.dout, ld, 1, 0}, Act2D{nullptr, 0, 0, 0},
                  Act2D{p.l2_in[nxt], h0, p.K2_pad / 64, 1}, p.dec_out, p.max_steps * p.Dd, t * p.Dd};
      tc::fence_proxy_async();
      __syncthreads();
      lstm_phase_tc(p.tc_2, io, p.B, p.Bp, p.RB, p.zc, p.zh, tcs);
    } else {
      const size_t hoff = (size_t)p.Dd * ld;
      dense_phase(p.lstm_2, p.l2_in[cur], ld, p.Bp, wres, scratch,
                  LstmEpi{p.lstm_2.bias, p.Dd, ld, p.B, p.c_2, p.l2_in[cur] + hoff, p.dout,
                          p.l2_in[nxt] + hoff, p.dec_out, p.max_steps * p.Dd, t * p.Dd, p.zc, p.zh});
    }
    if (k_tc && k_tail && t + 1 < p.t_end) phase_end(&p.tc_a, p.att_in[nxt]);
    else phase_end();
    if (!k_tail) {
      // ---- H-K: decoder self-attention hops over the KV cache ----
      const float* y = p.dout;
      for (int hi = 0; hi < k_hops; ++hi) {
        const SaHop& h = p.hop[hi];
        dense_phase(h.qkv, y, ld, p.Bp, wres, scratch,
                    QkvEpi{h.qkv.bias, p.Ds, p.Ds / p.heads, p.heads, p.Smax, t, ld, p.B, h.kcache, h.vcache, p.sa_q});
        phase_end();
        for (int item = blockIdx.x; item < p.B * p.heads; item += gridDim.x)
          sa_item(p, h, item / p.heads, item % p.heads, t, scratch);
        phase_end();
        dense_phase(h.o, p.sa_ctx, ld, p.Bp, wres, scratch,
                    DenseEpi{h.o.bias, h.o.N, ld, 0, Act2D{p.sa_a, ld, 1, 0}, nullptr, nullptr});
        phase_end();
        dense_phase(h.tr, p.sa_a, ld, p.Bp, wres, scratch,
                    DenseEpi{h.tr.bias, h.tr.N, ld, 2, Act2D{p.sa_y[hi & 1], ld, 1, 0}, nullptr, y});
        phase_end();
        y = p.sa_y[hi & 1];
      }
      // ---- L+A+B: frame + stop projection of step t, fed-back frame, prenet of step t+1 (row split) ----
      frame_phase<RETIRE, PROF, SPEC>(p, t, y, p.outp.K, Act2D{p.att_in[nxt], k_tc ? 0 : ld, k_tc ? p.Ka_pad / 64 : 1, k_tc}, false, scratch,
                          s_live, RETIRE ? s_nlive : p.B);
      if (k_tc && t + 1 < p.t_end) phase_end(&p.tc_a, p.att_in[nxt]);
      else phase_end();
    }
    // ---- stop rule (every CTA evaluates the same data -> same decision) ----
    if (p.stop_mode != SAT_STOP_FIXED) {
      bool term = true, stopped = true, fresh = false;
      if (tid < p.B) {
        const float logit = __ldcg(p.stop + (size_t)tid * p.max_steps + t);
        term = (sigmoidf_(logit) > 0.5f) && (t > p.min_iters);
        if (term && s_stopped[tid] < 0) {
          s_stopped[tid] = t;
          fresh = true;
          if (blockIdx.x == 0) p.stop_steps[tid] = t;
        }
        stopped = s_stopped[tid] >= 0;
      }
      // RETIRE runs under the per-row rule only: its first vote asks whether any row stopped at this step
      const int all_term = RETIRE ? __syncthreads_or(fresh ? 1 : 0) : __syncthreads_and(term ? 1 : 0);
      const int all_stopped = __syncthreads_and(stopped ? 1 : 0);
      if ((!RETIRE && p.stop_mode == SAT_STOP_ALL_ROWS) ? all_term : all_stopped) {
        ++t;
        fired = 1;
        break;
      }
      if (RETIRE && all_term) {
        // compact the rows that are still decoding (every CTA derives the same list from the same stop logits)
        const bool alive = tid < p.B && s_stopped[tid] < 0;
        const unsigned bal = __ballot_sync(0xffffffffu, alive);
        if (tid < MAX_B && (tid & 31) == 0) s_cnt[tid >> 5] = __popc(bal);
        __syncthreads();
        if (alive) {
          int off = __popc(bal & ((1u << (tid & 31)) - 1u));
          for (int w = 0; w < (tid >> 5); ++w) off += s_cnt[w];
          s_live[off] = (short)tid;
        }
        if (tid == 0) {
          int n = 0;
          for (int w = 0; w < MAX_B / 32; ++w) n += s_cnt[w];
          s_nlive = n;
        }
        __syncthreads();
      }
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    *p.n_steps = t;
    if (p.finished) *p.finished = fired;
  }
  if (k_tc) {
    if (tid == 0) {
      // a stop decision can leave the requests for the next step in flight: let them land before the CTA exits
      for (uint32_t c = 0; c < tcs.prefilled; ++c) {
        const uint32_t n = tcs.n_chunks + c;
        tc::mbar_wait(tcs.full + n % tcs.stages, (n / tcs.stages) & 1);
      }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (tid < 32) tc::tmem_dealloc(tcs.tmem, 32);
  }
}

}  // namespace dec
}  // namespace sat
