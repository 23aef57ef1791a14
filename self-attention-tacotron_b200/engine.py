"""Handle objects over the C ABI: thin owners of sat_encoder / sat_decoder /
sat_postnet.  They only marshal torch CUDA tensors' ``data_ptr()`` into
libsat.so; all arithmetic is in the CUDA kernels."""
import ctypes as C

import torch

from . import _lib
from ._lib import lib, check, ptr, stream_ptr


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("libsat operates on CUDA tensors only (no CPU fallback)")


class DecoderHandle:
    """sat_decoder_* (ExtendedDecoder / DualSourceDecoder / DualSourceTransformerDecoder,
    modules/module.py:562-623, 1083-1152, 1499-1562)."""

    def __init__(self, hp, weights):
        self.hp = hp
        self._h = C.c_void_p()
        arr, n, keep = _lib.tensor_list({k: v for k, v in weights.items() if k.startswith("decoder/")})
        s = _lib.make_hparams(hp)
        check(lib().sat_decoder_create(C.byref(s), arr, n, C.byref(self._h)))
        self.n_mech = 1 if hp.decoder == "ExtendedDecoder" else 2

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value and lib is not None:  # (module globals are gone at interpreter exit)
            lib().sat_decoder_destroy(self._h)
            self._h = C.c_void_p()

    def run(self, memories, lengths, *, target=None, teacher_alignments=None, speaker_embed=None,
            feed_mode, stop_mode, max_steps, forced_alignment=False, min_iters=10,
            want_alignments=True, want_sample_ids=False, want_decoder_outputs=False,
            want_self_attention_probs=False, stream=None, step_by_step=False):
        hp = self.hp
        if not isinstance(memories, (tuple, list)):
            memories = (memories,)
        memories = [m.contiguous().float() for m in memories]
        _need_cuda(*memories, lengths, target, speaker_embed)
        dev = memories[0].device
        B, T = memories[0].shape[0], memories[0].shape[1]
        M, r = hp.num_mels, hp.outputs_per_step
        lengths = lengths.to(device=dev, dtype=torch.int64).contiguous()
        io = _lib.SatDecoderIO()
        io.B, io.T_enc = B, T
        io.memory = (C.c_void_p * 2)(*[m.data_ptr() for m in memories] + [0] * (2 - len(memories)))
        io.memory_lengths = lengths.data_ptr()
        if speaker_embed is not None:
            speaker_embed = speaker_embed.contiguous().float()
            io.speaker_embed = speaker_embed.data_ptr()
        S_t = 0
        if target is not None:
            target = target.contiguous().float()
            S_t = target.shape[1] // r
            io.target = target.data_ptr()
        io.target_steps = S_t
        tas = []
        if teacher_alignments is not None:
            tas = [a.contiguous().float() if a is not None else None for a in teacher_alignments]
            io.teacher_alignments = (C.c_void_p * 2)(*[(a.data_ptr() if a is not None else 0) for a in tas]
                                                     + [0] * (2 - len(tas)))
            if S_t == 0:
                S_t = tas[0].shape[1]
                io.target_steps = S_t
        io.forced_alignment = int(forced_alignment)
        io.feed_mode, io.stop_mode, io.max_steps, io.min_iters = feed_mode, stop_mode, max_steps, min_iters
        S = max_steps
        # per-row stop rule: a row leaves the row-split phases at its own stop step, nothing is written for it
        # afterwards -- the buffers start at zero so that those frames are well defined
        new = torch.zeros if stop_mode == _lib.STOP_PER_ROW else torch.empty
        mel = new((B, S * r, M), device=dev, dtype=torch.float32)
        stop = new((B, S), device=dev, dtype=torch.float32)
        aligns = [new((B, S, T), device=dev, dtype=torch.float32) if want_alignments else None
                  for _ in range(self.n_mech)]
        ids = torch.empty((B, S, r), device=dev, dtype=torch.int32) if want_sample_ids else None
        dout = torch.empty((B, S, hp.decoder_out_units), device=dev, dtype=torch.float32) if want_decoder_outputs else None
        stop_steps = torch.empty((B,), device=dev, dtype=torch.int32)
        n_steps = torch.zeros((1,), device=dev, dtype=torch.int32)
        io.mel, io.stop = mel.data_ptr(), stop.data_ptr()
        io.alignments = (C.c_void_p * 2)(*[(a.data_ptr() if a is not None else 0) for a in aligns]
                                         + [0] * (2 - len(aligns)))
        io.sample_ids = ids.data_ptr() if ids is not None else 0
        io.decoder_outputs = dout.data_ptr() if dout is not None else 0
        io.stop_steps, io.n_steps = stop_steps.data_ptr(), n_steps.data_ptr()
        sa_probs = []
        if want_self_attention_probs and hp.decoder == "DualSourceTransformerDecoder":
            sa_probs = [torch.empty((B, hp.decoder_self_attention_num_heads, S, S), device=dev, dtype=torch.float32)
                        for _ in range(hp.decoder_self_attention_num_hop)]
        io.self_attention_probs = (C.c_void_p * 2)(*[a.data_ptr() for a in sa_probs] + [0] * (2 - len(sa_probs)))
        if step_by_step:
            # sat_decoder_step: the same run, one launch per decoder step (state kept in the handle)
            fin = torch.zeros((1,), device=dev, dtype=torch.int32)
            for t in range(S):
                check(lib().sat_decoder_step(self._h, C.byref(io), t, ptr(fin), stream_ptr(stream)))
                if stop_mode != _lib.STOP_FIXED and int(fin.item()):
                    break  # the stop rule ended the run at step t
        else:
            check(lib().sat_decoder_run(self._h, C.byref(io), stream_ptr(stream)))
        return {"mel": mel, "stop": stop, "alignments": aligns, "sample_ids": ids, "decoder_outputs": dout,
                "self_attention_probs": sa_probs,
                "stop_steps": stop_steps, "n_steps": n_steps, "_keep": (memories, lengths, target, tas, speaker_embed)}

    def step_bytes(self, B, T, avg_cache_len=0):
        return lib().sat_decoder_step_bytes(self._h, B, T, avg_cache_len)

    def kernel_ms(self):
        ms = C.c_float()
        check(lib().sat_decoder_kernel_ms(self._h, C.byref(ms)))
        return ms.value

    def used_tensor_cores(self):
        used = C.c_int32()
        check(lib().sat_decoder_used_tensor_cores(self._h, C.byref(used)))
        return bool(used.value)

    def set_row_retirement(self, on):
        """per-row stop rule: retire stopped rows from the row-split phases (default) or let them run on"""
        check(lib().sat_decoder_set_row_retirement(self._h, C.c_int32(1 if on else 0)))

    def cluster_size(self):
        size = C.c_int32()
        check(lib().sat_decoder_cluster_size(self._h, C.byref(size)))
        return int(size.value)

    def profile(self, reset=True):
        out = (C.c_int64 * 48)()
        check(lib().sat_decoder_profile(self._h, out, int(reset)))
        return list(out)


class EncoderHandle:
    """sat_encoder_* (ZoneoutEncoderV1 modules/module.py:336-339, SelfAttentionCBHGEncoder :425-438,
    embedding lookup models/models.py:28,54)."""

    def __init__(self, hp, weights):
        self.hp = hp
        self._h = C.c_void_p()
        keep = {k: v for k, v in weights.items() if k.startswith("encoder/") or k == "embedding"}
        arr, n, _keep = _lib.tensor_list(keep)
        s = _lib.make_hparams(hp)
        check(lib().sat_encoder_create(C.byref(s), arr, n, C.byref(self._h)))
        self.self_attention = hp.encoder == "SelfAttentionCBHGEncoder"

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value and lib is not None:  # (module globals are gone at interpreter exit)
            lib().sat_encoder_destroy(self._h)
            self._h = C.c_void_p()

    def uses_tensor_cores(self):
        used = C.c_int32()
        check(lib().sat_encoder_uses_tensor_cores(self._h, C.byref(used)))
        return bool(used.value)

    def set_mask_padding(self, on):
        """True: every utterance is computed as if alone in the batch (the reference's predict runs batch_size=1,
        predict_mel.py:43-44); False (default): the reference's batched graph, padding unmasked."""
        check(lib().sat_encoder_set_mask_padding(self._h, int(bool(on))))

    def check(self, stream=None):
        """Synchronises, then raises ValueError if a symbol id outside [0, num_symbols) reached the encoder."""
        check(lib().sat_encoder_check(self._h, stream_ptr(stream)))

    def poll_error(self):
        """check() without the synchronisation: reports bad ids of the forwards that have finished on the device"""
        check(lib().sat_encoder_poll_error(self._h))

    def forward(self, ids_or_embedded, lengths, want_probs=False, stream=None):
        hp = self.hp
        x = ids_or_embedded
        _need_cuda(x, lengths)
        dev = x.device
        B, T = x.shape[0], x.shape[1]
        lengths = lengths.to(device=dev, dtype=torch.int64).contiguous()
        lstm_out = torch.empty((B, T, hp.cbhg_out_units), device=dev, dtype=torch.float32)
        sa_out = probs = None
        if self.self_attention:
            sa_out = torch.empty((B, T, hp.self_attention_out_units), device=dev, dtype=torch.float32)
            if want_probs:
                probs = torch.empty((hp.self_attention_num_hop * hp.self_attention_num_heads, B, T, T),
                                    device=dev, dtype=torch.float32)
        if x.dtype in (torch.int64, torch.int32):
            x = x.to(torch.int64).contiguous()
            check(lib().sat_encoder_forward(self._h, ptr(x), ptr(lengths), B, T, ptr(lstm_out), ptr(sa_out), ptr(probs),
                                            stream_ptr(stream)))
        else:
            x = x.contiguous().float()
            if x.shape[2] != hp.embedding_dim:
                raise ValueError(f"encoder input depth {x.shape[2]} != embedding_dim {hp.embedding_dim}")
            check(lib().sat_encoder_forward_embedded(self._h, ptr(x), ptr(lengths), B, T, ptr(lstm_out), ptr(sa_out),
                                                     ptr(probs), stream_ptr(stream)))
        self._keep = (x, lengths)
        return lstm_out, sa_out, probs


class PostnetHandle:
    """sat_postnet_* (tacotron2 PostNetV2, models/models.py:4)."""

    def __init__(self, hp, weights):
        self.hp = hp
        self._h = C.c_void_p()
        arr, n, _keep = _lib.tensor_list({k: v for k, v in weights.items() if k.startswith("postnet/")})
        s = _lib.make_hparams(hp)
        check(lib().sat_postnet_create(C.byref(s), arr, n, C.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value and lib is not None:  # (module globals are gone at interpreter exit)
            lib().sat_postnet_destroy(self._h)
            self._h = C.c_void_p()

    def uses_tensor_cores(self):
        used = C.c_int32()
        check(lib().sat_postnet_uses_tensor_cores(self._h, C.byref(used)))
        return bool(used.value)

    def forward(self, mel, stream=None):
        _need_cuda(mel)
        mel = mel.contiguous().float()
        out = torch.empty_like(mel)
        check(lib().sat_postnet_forward(self._h, ptr(mel), mel.shape[0], mel.shape[1], ptr(out), stream_ptr(stream)))
        self._keep = mel
        return out


# ---- stand-alone module entry points (unit parity; each call uploads its weights) ----

def dense_forward(x, kernel, bias=None, act=None, stream=None):
    """tf.layers.Dense on the last axis through sat_dense_forward."""
    import numpy as np
    _need_cuda(x)
    x = x.contiguous().float()
    K, N = kernel.shape
    if x.shape[-1] != K:
        raise ValueError(f"Dense: input depth {x.shape[-1]} != kernel rows {K}")
    k = np.ascontiguousarray(kernel, dtype=np.float32)
    b = np.ascontiguousarray(bias, dtype=np.float32) if bias is not None else None
    y = torch.empty(x.shape[:-1] + (N,), device=x.device, dtype=torch.float32)
    rows = x.numel() // K
    check(lib().sat_dense_forward(C.c_void_p(k.ctypes.data), C.c_void_p(b.ctypes.data if b is not None else 0),
                                  K, N, _lib.ACT[act], ptr(x), C.c_int64(rows), ptr(y), stream_ptr(stream)))
    return y


def embedding_lookup(table, ids, index_offset=0, stream=None):
    import numpy as np
    _need_cuda(ids)
    t = np.ascontiguousarray(table, dtype=np.float32)
    ids = ids.to(torch.int64).contiguous()
    out = torch.empty(tuple(ids.shape) + (t.shape[1],), device=ids.device, dtype=torch.float32)
    check(lib().sat_embedding_lookup(C.c_void_p(t.ctypes.data), C.c_int64(t.shape[0]), t.shape[1],
                                     C.c_int64(index_offset), ptr(ids), C.c_int64(ids.numel()), ptr(out),
                                     stream_ptr(stream)))
    return out


def mha_forward(weights, prefix, x, num_heads, model_dim, causal=False, lengths=None, want_probs=True, stream=None):
    """MultiHeadAttention with K=V=Q=x (SelfAttention, modules/self_attention.py:131-144)."""
    _need_cuda(x, lengths)
    x = x.contiguous().float()
    B, T, Din = x.shape
    arr, n, _keep = _lib.tensor_list({k: v for k, v in weights.items() if k.startswith(prefix + "/")})
    out = torch.empty((B, T, model_dim), device=x.device, dtype=torch.float32)
    probs = torch.empty((B, num_heads, T, T), device=x.device, dtype=torch.float32) if want_probs else None
    if lengths is not None:
        lengths = lengths.to(device=x.device, dtype=torch.int64).contiguous()
    check(lib().sat_mha_forward(arr, n, prefix.encode(), ptr(x), B, T, Din, model_dim, num_heads, int(causal),
                                ptr(lengths), ptr(out), ptr(probs), stream_ptr(stream)))
    return out, probs


def sdpa_forward(query, keys, values, causal=False, lengths=None, want_probs=True, stream=None):
    """ScaledDotProductAttentionMechanism.__call__ (modules/self_attention.py:45-65) on projected heads
    [B, H, T, d_head] -> (context [B, H, T, d_head], probs [B, H, T, T]).  The three tensors are interleaved into the
    kernel's [B, T, 3D] layout (tensor plumbing); the arithmetic is sat_sdpa_forward."""
    _need_cuda(query, keys, values, lengths)
    B, H, T, dh = query.shape
    D = H * dh
    merge = lambda a: a.float().permute(0, 2, 1, 3).reshape(B, T, D)  # noqa: E731
    qkv = torch.cat((merge(keys), merge(values), merge(query)), dim=-1).contiguous()
    ctx = torch.empty((B, T, D), device=query.device, dtype=torch.float32)
    probs = torch.empty((B, H, T, T), device=query.device, dtype=torch.float32) if want_probs else None
    if lengths is not None:
        lengths = lengths.to(device=query.device, dtype=torch.int64).contiguous()
    check(lib().sat_sdpa_forward(ptr(qkv), B, T, D, H, int(causal), ptr(lengths), ptr(ctx), ptr(probs), stream_ptr(stream)))
    return ctx.reshape(B, T, H, dh).permute(0, 2, 1, 3), probs


def sat_forward(weights, prefix, x, num_heads, causal=False, want_probs=True, stream=None):
    """SelfAttentionTransformer (modules/module.py:363-371): x + tanh(Dense(SelfAttention(x)))."""
    _need_cuda(x)
    x = x.contiguous().float()
    B, T, D = x.shape
    arr, n, _keep = _lib.tensor_list({k: v for k, v in weights.items() if k.startswith(prefix + "/")})
    out = torch.empty_like(x)
    probs = torch.empty((B, num_heads, T, T), device=x.device, dtype=torch.float32) if want_probs else None
    check(lib().sat_sat_forward(arr, n, prefix.encode(), ptr(x), B, T, D, num_heads, int(causal), ptr(out), ptr(probs),
                                stream_ptr(stream)))
    return out, probs


def attention_prepare(weights, prefix, memory, lengths, num_units, stream=None):
    """keys = (memory * mask) @ memory_layer, values = memory * mask (TF BahdanauAttention)."""
    _need_cuda(memory, lengths)
    memory = memory.contiguous().float()
    B, T, Dm = memory.shape
    lengths = lengths.to(device=memory.device, dtype=torch.int64).contiguous()
    arr, n, _keep = _lib.tensor_list({k: v for k, v in weights.items() if k.startswith(prefix + "/")})
    keys = torch.empty((B, T, num_units), device=memory.device, dtype=torch.float32)
    values = torch.empty_like(memory)
    check(lib().sat_attention_prepare(arr, n, prefix.encode(), ptr(memory), ptr(lengths), B, T, Dm, num_units,
                                      ptr(keys), ptr(values), stream_ptr(stream)))
    return keys, values


def attention_step(weights, prefix, kind, options, keys, values, lengths, query, state, stream=None):
    """One AttentionMechanism.__call__ (+ context) through sat_attention_step.
    state = dict(alignments [B,T], alpha [B,T] | None, u [B] | None) -- updated IN PLACE."""
    _need_cuda(keys, values, lengths, query)
    B, T, U = keys.shape
    Dv, A = values.shape[2], query.shape[1]
    query = query.contiguous().float()
    lengths = lengths.to(device=keys.device, dtype=torch.int64).contiguous()
    arr, n, _keep = _lib.tensor_list({k: v for k, v in weights.items() if k.startswith(prefix + "/")})
    align = torch.empty((B, T), device=keys.device, dtype=torch.float32)
    ctx = torch.empty((B, Dv), device=keys.device, dtype=torch.float32)
    check(lib().sat_attention_step(arr, n, prefix.encode(), _lib.SAT_ATT[kind], B, T, U, Dv, A,
                                   int(options.get("attention_kernel", 1)), int(options.get("attention_filters", 1)),
                                   int(bool(options.get("cumulative_weights", False))),
                                   int(bool(options.get("use_transition_agent", False))),
                                   ptr(keys), ptr(values), ptr(lengths), ptr(query),
                                   ptr(state.get("alignments")), ptr(state.get("alpha")), ptr(state.get("u")),
                                   ptr(align), ptr(ctx), stream_ptr(stream)))
    return align, ctx
