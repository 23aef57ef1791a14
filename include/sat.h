/* sat.h -- C ABI of libsat.so: the B200-native mel-synthesis hot path of
 * self-attention Tacotron.  Plain pointers and sizes only; no torch types.
 *
 * Every entry point replaces a piece of the reference's Python/TensorFlow call
 * surface (citations are into /root/reference):
 *   sat_encoder_*   SelfAttentionCBHGEncoder.call  modules/module.py:425-438
 *                   ZoneoutEncoderV1.call          modules/module.py:336-339
 *                   Embedding lookup               models/models.py:28,54
 *   sat_decoder_*   ExtendedDecoder.call           modules/module.py:562-623
 *                   DualSourceDecoder.call         modules/module.py:1083-1152
 *                   DualSourceTransformerDecoder.call :1499-1562 (+RNNTransformer :726-784)
 *                   attention_mechanism_factory    modules/attentions.py:25-62
 *                   helpers (feed/stop rules)      modules/helpers.py:13-108,156-160
 *   sat_mha_forward SelfAttention / MultiHeadAttention  modules/self_attention.py:89-144
 *   sat_sat_forward SelfAttentionTransformer.call  modules/module.py:363-371
 *   sat_attention_step  ForwardAttention.__call__  modules/forward_attention.py:88-122
 *                       (+ additive / location_sensitive, modules/attentions.py:35-57)
 *   sat_dense_forward   tf.layers.Dense (PreNet, MultiSpeakerPreNet modules/multi_speaker_modules.py:27-32)
 *   sat_postnet_*   tacotron2 PostNetV2 (models/models.py:4)
 *
 * Conventions: all tensors fp32 row-major in TensorFlow layouts ([B,T,C]
 * activations, Dense [in,out], Conv1D [k,in,out], LSTM [in+units,4*units] gate
 * order i,j,f,o); ids/lengths int64.  Activation/result pointers are DEVICE
 * pointers owned by the caller; weight pointers handed to *_create are HOST
 * pointers (copied and repacked once).  All work is enqueued on `stream`
 * (a cudaStream_t passed as void*).  Every function returns 0 or a negative
 * SAT_E* code; sat_last_error() gives the message (thread-local).
 */
#ifndef SAT_H_
#define SAT_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SAT_ABI_VERSION 2

enum { SAT_OK = 0, SAT_EINVAL = -1, SAT_ECUDA = -2, SAT_EMISSING = -3, SAT_EUNSUPPORTED = -4, SAT_ENOMEM = -5 };

/* attention kinds (modules/attentions.py:27-59) */
enum { SAT_ATT_NONE = 0, SAT_ATT_ADDITIVE = 1, SAT_ATT_LOCATION_SENSITIVE = 2, SAT_ATT_FORWARD = 3,
       SAT_ATT_TEACHER_FORCING = 4 };
/* decoder kinds */
enum { SAT_DEC_EXTENDED = 0, SAT_DEC_DUAL_SOURCE = 1, SAT_DEC_DUAL_SOURCE_TRANSFORMER = 2 };
/* encoder kinds */
enum { SAT_ENC_ZONEOUT_V1 = 0, SAT_ENC_SELF_ATTENTION_CBHG = 1 };
/* feed rule = which seq2seq Helper the reference would use (modules/helpers.py) */
enum { SAT_FEED_TEACHER = 0,      /* targets[:, t, -M*n_feed:]        helpers.py:54-58 */
       SAT_FEED_SELF = 1,         /* output[:, -M*n_feed:]            tacotron2 helpers (A.7) */
       SAT_FEED_SELF_SOFTMAX = 2  /* softmax per r-slot of output     helpers.py:98-106 */ };
/* stop rule */
enum { SAT_STOP_FIXED = 0,     /* validation: S = T_dec / r steps          helpers.py:102 */
       SAT_STOP_ALL_ROWS = 1,  /* reduce_all(sigmoid(stop)>.5 & t>10)      helpers.py:156-160 */
       SAT_STOP_PER_ROW = 2    /* same rule per utterance; run ends when every row stopped.  A stopped row is retired:
                                  mel / stop / alignment entries after its own stop step are NOT written */ };

/* named host tensor of the weight blob (names: DESIGN.md "weight blob") */
typedef struct { const char* name; const float* data; int64_t numel; } sat_tensor;

/* the hyper-parameters the path reads; names follow hparams.py */
typedef struct {
  int32_t num_mels, outputs_per_step, n_feed_frame;
  int32_t num_symbols, embedding_dim;
  int32_t encoder_kind;
  int32_t encoder_prenet_units[2];
  int32_t cbhg_out_units, conv_channels, max_filter_width;
  int32_t projection1_out_channels, projection2_out_channels, num_highway;
  int32_t self_attention_out_units, self_attention_num_heads, self_attention_num_hop;
  int32_t decoder_kind;
  int32_t decoder_prenet_units[2];
  int32_t attention_out_units, decoder_out_units;
  int32_t attention_kind[2];   /* per source; [1] unused for SAT_DEC_EXTENDED */
  int32_t attention_units[2];
  int32_t memory_dim[2];
  int32_t attention_kernel, attention_filters;
  int32_t cumulative_weights, use_transition_agent;
  int32_t decoder_self_attention_out_units, decoder_self_attention_num_heads, decoder_self_attention_num_hop;
  int32_t multi_speaker_prenet, speaker_embedding_dim;
  int32_t num_postnet_layers, postnet_kernel_size, postnet_channels;
  int32_t max_iters;
  float zoneout_factor_cell, zoneout_factor_output;
} sat_hparams;

typedef struct sat_encoder sat_encoder;
typedef struct sat_decoder sat_decoder;
typedef struct sat_postnet sat_postnet;

int sat_abi_version(void);
const char* sat_last_error(void);
/* number of kernels this library has launched on the calling thread since the last reset */
int64_t sat_launch_count(void);
void sat_launch_count_reset(void);

/* Host helper of the sharded predict path (no reference counterpart: the reference buckets with
 * tf.data bucket_by_sequence_length, datasets/dataset.py): batch cuts of `n` utterance lengths sorted longest first,
 * minimising sum over batches of longest * (c0 + c1 * batch size) with at most batch_size utterances per batch.
 * cut[i] = size of the batch that starts at i in the optimal cutting of lengths[i..n) (follow it from i = 0).  CPU
 * only; no device work. */
int sat_bucket_cuts(const int64_t* sorted_lengths, int64_t n, int32_t batch_size, double c0, double c1, int32_t* cut);

/* ---- encoder ---------------------------------------------------------- */
int sat_encoder_create(const sat_hparams* hp, const sat_tensor* weights, int32_t n_weights, sat_encoder** out);
/* ids [B,T] int64 (device), lengths [B] int64 (device).
 * lstm_out [B,T,cbhg_out_units]; sa_out [B,T,sa_units] and sa_probs
 * [hops*heads,B,T,T] may be NULL (must be NULL for SAT_ENC_ZONEOUT_V1). */
int sat_encoder_forward(sat_encoder* h, const int64_t* ids, const int64_t* lengths, int32_t B, int32_t T,
                        float* lstm_out, float* sa_out, float* sa_probs, void* stream);
/* same, for callers that did the embedding lookup themselves: embedded [B,T,embedding_dim] (device) */
int sat_encoder_forward_embedded(sat_encoder* h, const float* embedded, const int64_t* lengths, int32_t B, int32_t T,
                                 float* lstm_out, float* sa_out, float* sa_probs, void* stream);
/* 1 if this encoder runs its conv bank / max-pool / proj1 on the tensor cores (tcgen05 implicit-GEMM over fp16
 * activations, fp32 accumulation; needs prenet/conv widths that are multiples of 64 / 128), 0 for the fp32 SIMT form.
 * Environment SAT_ENC_TC=0 at create time forces the fp32 form. */
int sat_encoder_uses_tensor_cores(sat_encoder* h, int32_t* used);
/* Padded frames of a batch.  0 (default) = the reference's batched graph: padding (the embedding of id 0) flows through
 * the conv bank / projections and the encoder self-attention unmasked (modules/module.py:77-110, 425-438; no
 * sequence-length masking anywhere but the BiLSTM).  1 = every utterance is computed as if it were alone in the batch,
 * i.e. what the reference's predict produces with its forced batch_size=1 (predict_mel.py:43-44): frames past an
 * utterance's length read as zeros in every multi-tap convolution and are masked as keys of the self-attention. */
int sat_encoder_set_mask_padding(sat_encoder* h, int32_t on);
/* Synchronises `stream`, then returns SAT_EINVAL if any forward issued so far met a symbol id outside [0, num_symbols)
 * (tacotron2 Embedding asserts the range, models/models.py:28; here such an id embeds as zeros and sets a sticky flag,
 * which the next sat_encoder_forward also reports once the offending forward has finished). */
int sat_encoder_check(sat_encoder* h, void* stream);
/* The same report without synchronising: covers the forwards that have already finished on the device (each ends by
 * copying the flag to pinned host memory) -- for pipelined callers that wait on their own events. */
int sat_encoder_poll_error(sat_encoder* h);
void sat_encoder_destroy(sat_encoder* h);

/* ---- decoder ---------------------------------------------------------- */
typedef struct {
  int32_t B, T_enc;             /* batch, padded memory length */
  const float* memory[2];       /* [B,T_enc,memory_dim[i]] encoder outputs (device) */
  const int64_t* memory_lengths;/* [B] (device) */
  const float* speaker_embed;   /* [B,speaker_embedding_dim] or NULL */
  const float* target;          /* [B,S*r,M] teacher frames or NULL */
  int32_t target_steps;         /* S (= T_dec / r) when target != NULL */
  const float* teacher_alignments[2]; /* [B,S,T_enc] or NULL (SAT_ATT_TEACHER_FORCING replay) */
  int32_t forced_alignment;     /* 1: replace every mechanism by teacher-forcing replay */
  int32_t feed_mode, stop_mode;
  int32_t max_steps;            /* upper bound on steps (min(max_iters, S) for SAT_STOP_FIXED) */
  int32_t min_iters;            /* 10 in the reference */
  /* outputs (device, caller-owned, sized for max_steps) */
  float* mel;                   /* [B,max_steps*r,M] */
  float* stop;                  /* [B,max_steps] stop-token logits */
  float* alignments[2];         /* [B,max_steps,T_enc] or NULL */
  int32_t* sample_ids;          /* [B,max_steps,r] argmax ids or NULL   helpers.py:49-52 */
  float* decoder_outputs;       /* [B,max_steps,decoder_out_units] pre-projection cell outputs or NULL */
  int32_t* stop_steps;          /* [B] first step meeting the stop rule (-1 if none) */
  int32_t* n_steps;             /* [1] steps actually run */
  /* decoder self-attention probabilities per hop, [B,heads,max_steps,max_steps] (row t = step t, causal; zero above
   * the diagonal) or NULL: what models/models.py:110-111,143-144 reads from decoder_state.alignments ("alignment3/4") */
  float* self_attention_probs[2];
} sat_decoder_io;

int sat_decoder_create(const sat_hparams* hp, const sat_tensor* weights, int32_t n_weights, sat_decoder** out);
int sat_decoder_run(sat_decoder* h, const sat_decoder_io* io, void* stream);
/* One decoder step with the recurrent state kept in the handle -- the RNN-cell contract of the reference
 * (`cell(inputs, state) -> (output, new_state)`: modules/rnn_wrappers.py:74-124,210-214, tacotron2 DecoderRNNV2 /
 * AttentionRNN) for debugging and step-by-step parity.  `io` describes the WHOLE run exactly as for sat_decoder_run
 * (buffers sized for max_steps); the call executes step `step` only and writes its frames, stop logit and alignments
 * at position `step` of those buffers.  step 0 initialises the state; step k > 0 must follow step k - 1 of the same
 * run on the same handle (SAT_EINVAL otherwise).  n_steps receives step + 1; `finished` (device int32, may be null)
 * receives 1 if the stop rule (io->stop_mode) ended the run at this step, else 0.  Runs the fp32 form at every batch
 * size (the tensor-core form keeps the cell state on the SMs for the length of a launch). */
int sat_decoder_step(sat_decoder* h, const sat_decoder_io* io, int32_t step, int32_t* finished, void* stream);
/* average algorithmic bytes one decoder step moves for the last run (SURVEY.md 8d formula) */
double sat_decoder_step_bytes(const sat_decoder* h, int32_t B, int32_t T_enc, int32_t avg_cache_len);
void sat_decoder_destroy(sat_decoder* h);

/* ---- stand-alone module entry points (unit parity) -------------------- */
/* SelfAttention: weights = {key,value,query,output}/{kernel,bias} under `prefix`.
 * x [B,T,Din] -> out [B,T,D]; probs [B,H,T,T] or NULL.  lengths != NULL enables the padding mask. */
int sat_mha_forward(const sat_tensor* weights, int32_t n_weights, const char* prefix,
                    const float* x, int32_t B, int32_t T, int32_t Din, int32_t D, int32_t H,
                    int32_t causal, const int64_t* lengths, float* out, float* probs, void* stream);
/* Allocation-free variant of sat_mha_forward.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  It calls no cudaMalloc / cudaFree / allocating library routine and does not synchronise the stream.
 * Weights are DEVICE pointers supplied by the caller: wqkv [Din, 3*D] with the column blocks key | value | query,
 * bqkv [3*D], wo [D, D], bo [D]; x [B,T,Din] -> out [B,T,D]; probs [B,H,T,T] or NULL; lengths != NULL enables the
 * padding mask.  The qkv / context scratch (B*T*4*D floats) lives in one file-scope device array of
 * SAT_MHA_NOALLOC_MAX_FLOATS floats (8 MiB, held by every GPU that loads the library until exit); larger problems are
 * rejected with SAT_EINVAL.  All calls share that array: the region a call uses is cleared first, and calls that
 * overlap on the GPU (other streams / threads) race -- serialise them. */
#define SAT_MHA_NOALLOC_MAX_FLOATS (2 * 1024 * 1024)
int sat_mha_forward_noalloc(const float* wqkv, const float* bqkv, const float* wo, const float* bo,
                            const float* x, int32_t B, int32_t T, int32_t Din, int32_t D, int32_t H,
                            int32_t causal, const int64_t* lengths, float* out, float* probs, void* stream);
int64_t sat_mha_noalloc_max_floats(void);
/* ScaledDotProductAttentionMechanism.__call__ (modules/self_attention.py:45-65, masks :67-86) on already projected
 * heads: qkv device [B,T,3D] = key | value | query with head h in columns h*(D/H) .. of each block; ctx device [B,T,D]
 * (heads merged), probs device [B,H,T,T] or null.  softmax(mask(Q K^T / sqrt(D/H))) V; `lengths` (device [B] or null)
 * enables the padding mask (query rows and key columns past the length), `causal` the subsequent mask.  Asynchronous. */
int sat_sdpa_forward(const float* qkv, int32_t B, int32_t T, int32_t D, int32_t H, int32_t causal,
                     const int64_t* lengths, float* ctx, float* probs, void* stream);
/* SelfAttentionTransformer: prefix/mha/... and prefix/transform/... ; out = x + tanh(Dense(mha(x))) */
int sat_sat_forward(const sat_tensor* weights, int32_t n_weights, const char* prefix,
                    const float* x, int32_t B, int32_t T, int32_t D, int32_t H, int32_t causal,
                    float* out, float* probs, void* stream);
/* one attention-mechanism call: keys [B,T,U], values [B,T,Dv], query [B,A];
 * state_* are updated in place: prev_alignments [B,T], alpha [B,T], u [B];
 * alignments [B,T] and context [B,Dv] are written. */
int sat_attention_step(const sat_tensor* weights, int32_t n_weights, const char* prefix, int32_t kind,
                       int32_t B, int32_t T, int32_t U, int32_t Dv, int32_t A,
                       int32_t attention_kernel, int32_t attention_filters,
                       int32_t cumulative_weights, int32_t use_transition_agent,
                       const float* keys, const float* values, const int64_t* lengths, const float* query,
                       float* state_alignments, float* state_alpha, float* state_u,
                       float* alignments, float* context, void* stream);
/* keys = (memory * mask) @ memory_layer ; values = memory * mask   (TF BahdanauAttention, A.5) */
int sat_attention_prepare(const sat_tensor* weights, int32_t n_weights, const char* prefix,
                          const float* memory, const int64_t* lengths, int32_t B, int32_t T, int32_t Dm, int32_t U,
                          float* keys, float* values, void* stream);
/* y = act(x @ kernel + bias); act: 0 none, 1 relu, 2 tanh, 3 sigmoid, 4 softsign */
int sat_dense_forward(const float* kernel_host, const float* bias_host, int32_t K, int32_t N, int32_t act,
                      const float* x, int64_t rows, float* y, void* stream);
/* table[ids - index_offset]  (tacotron2 Embedding; models/models.py:28,43-50): table host [rows,dim], ids device [n] */
int sat_embedding_lookup(const float* table_host, int64_t table_rows, int32_t dim, int64_t index_offset,
                         const int64_t* ids, int64_t n, float* out, void* stream);
/* cycles accumulated by CTA 0 of the persistent decoder kernel (profiling aid): 48 host slots,
 * [0,16) phases of a step, [16,24) stages of the attention item, [24,27) tensor-core LSTM, [32,40) frame-item stages.
 * The counters cost ~2 % of a step, so they only run from the first call of this function on the handle onwards (or with
 * SAT_PROF=1 in the environment): call it once (reset = 1) before the runs to be profiled. */
int sat_decoder_profile(sat_decoder* h, int64_t* out16, int32_t reset);
/* device time (ms) of the persistent decoder kernel of the last sat_decoder_run, from CUDA events recorded on the
 * launching stream around that launch; blocks until the kernel has finished */
int sat_decoder_kernel_ms(sat_decoder* h, float* ms);
/* 1 if the last sat_decoder_run computed the LSTM gate GEMMs on the tensor cores (tcgen05, fp16 operands, fp32
 * accumulation; batch >= 64), 0 if it used the fp32 SIMT form */
int sat_decoder_used_tensor_cores(sat_decoder* h, int32_t* used);
/* thread-block cluster size of the last sat_decoder_run: 4 or 2 when every attention row (one
 * ForwardAttention.__call__, modules/forward_attention.py:88-122) was split over a cluster of CTAs (small batches:
 * B * mechanisms * size <= SM count, fp32 form), 1 when each row ran on one CTA */
int sat_decoder_cluster_size(sat_decoder* h, int32_t* size);
/* SAT_STOP_PER_ROW only: 1 (default) = a row leaves the row-split phases at its own stop step (the retirement
 * instantiation of the kernel: ~10 % slower per step while no row has stopped, up to ~12 % faster once rows stop far
 * apart), 0 = stopped rows keep running until the last row has stopped (same results up to each row's own stop step).
 * No reference counterpart: the reference synthesises one utterance per run (predict_mel.py:43-44). */
int sat_decoder_set_row_retirement(sat_decoder* h, int32_t on);

/* diagnostic: one-CTA self-test of the tcgen05 / TMA / TMEM building blocks (csrc/tc.cuh):
 * D[128,16] = A[128,K] . Wt[16,K]^T in TF32 with fp32 accumulation; all pointers device, K % 32 == 0, K <= 1024 */
int sat_tc_selftest(const float* A, const float* Wt, int32_t K, float* D, void* stream);

/* ---- post-net --------------------------------------------------------- */
int sat_postnet_create(const sat_hparams* hp, const sat_tensor* weights, int32_t n_weights, sat_postnet** out);
/* mel [B,T,M] -> mel_postnet [B,T,M] = mel + Dense(conv stack(mel)) */
int sat_postnet_forward(sat_postnet* h, const float* mel, int32_t B, int32_t T, float* mel_postnet, void* stream);
void sat_postnet_destroy(sat_postnet* h);
/* 1 if the post-net convolutions run on the tensor cores (tcgen05 implicit-GEMM, fp16 operands, fp32 accumulation);
 * SAT_POSTNET_TC=0 at create time forces the fp32 SIMT form */
int sat_postnet_uses_tensor_cores(sat_postnet* h, int32_t* used);

#ifdef __cplusplus
}
#endif
#endif /* SAT_H_ */
