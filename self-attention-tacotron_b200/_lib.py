"""ctypes binding of libsat.so (include/sat.h).  No CPU fallback: a missing
library or a failing call raises."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SAT_LIB", os.path.join(_HERE, "libsat.so"))  # SAT_LIB: A/B timing of another build

SAT_ATT = {"none": 0, "additive": 1, "location_sensitive": 2, "forward": 3,
           "teacher_forcing_forward": 4, "teacher_forcing_additive": 4}
SAT_DEC = {"ExtendedDecoder": 0, "DualSourceDecoder": 1, "DualSourceTransformerDecoder": 2}
SAT_ENC = {"ZoneoutEncoderV1": 0, "SelfAttentionCBHGEncoder": 1}
FEED_TEACHER, FEED_SELF, FEED_SELF_SOFTMAX = 0, 1, 2
STOP_FIXED, STOP_ALL_ROWS, STOP_PER_ROW = 0, 1, 2
ACT = {None: 0, "relu": 1, "tanh": 2, "sigmoid": 3, "softsign": 4}


class SatTensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("numel", C.c_int64)]


class SatHParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "num_mels", "outputs_per_step", "n_feed_frame", "num_symbols", "embedding_dim", "encoder_kind")] + [
        ("encoder_prenet_units", C.c_int32 * 2)] + [(n, C.c_int32) for n in (
            "cbhg_out_units", "conv_channels", "max_filter_width", "projection1_out_channels",
            "projection2_out_channels", "num_highway", "self_attention_out_units", "self_attention_num_heads",
            "self_attention_num_hop", "decoder_kind")] + [
        ("decoder_prenet_units", C.c_int32 * 2), ("attention_out_units", C.c_int32), ("decoder_out_units", C.c_int32),
        ("attention_kind", C.c_int32 * 2), ("attention_units", C.c_int32 * 2), ("memory_dim", C.c_int32 * 2)] + [
        (n, C.c_int32) for n in (
            "attention_kernel", "attention_filters", "cumulative_weights", "use_transition_agent",
            "decoder_self_attention_out_units", "decoder_self_attention_num_heads", "decoder_self_attention_num_hop",
            "multi_speaker_prenet", "speaker_embedding_dim", "num_postnet_layers", "postnet_kernel_size",
            "postnet_channels", "max_iters")] + [
        ("zoneout_factor_cell", C.c_float), ("zoneout_factor_output", C.c_float)]


class SatDecoderIO(C.Structure):
    _fields_ = [("B", C.c_int32), ("T_enc", C.c_int32), ("memory", C.c_void_p * 2), ("memory_lengths", C.c_void_p),
                ("speaker_embed", C.c_void_p), ("target", C.c_void_p), ("target_steps", C.c_int32),
                ("teacher_alignments", C.c_void_p * 2), ("forced_alignment", C.c_int32),
                ("feed_mode", C.c_int32), ("stop_mode", C.c_int32), ("max_steps", C.c_int32), ("min_iters", C.c_int32),
                ("mel", C.c_void_p), ("stop", C.c_void_p), ("alignments", C.c_void_p * 2), ("sample_ids", C.c_void_p),
                ("decoder_outputs", C.c_void_p), ("stop_steps", C.c_void_p), ("n_steps", C.c_void_p),
                ("self_attention_probs", C.c_void_p * 2)]


_lib = None


def lib():
    """Load libsat.so (built by __graft_entry__.build()); raise loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`. "
                               "There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.sat_last_error.restype = C.c_char_p
        L.sat_launch_count.restype = C.c_int64
        L.sat_decoder_step_bytes.restype = C.c_double
        L.sat_decoder_step_bytes.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
        for n in ("sat_encoder_destroy", "sat_decoder_destroy", "sat_postnet_destroy", "sat_launch_count_reset"):
            getattr(L, n).argtypes = [C.c_void_p] if n != "sat_launch_count_reset" else []
            if hasattr(L, n):
                getattr(L, n).restype = None
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().sat_last_error().decode("utf-8", "replace")
        raise (ValueError if rc in (-1, -3, -4) else RuntimeError)(f"libsat error {rc}: {msg}")


def tensor_list(weights):
    """{name: ndarray} -> (SatTensor array, n, keep-alive list)."""
    keep, arr = [], (SatTensor * len(weights))()
    for i, (k, v) in enumerate(weights.items()):
        a = np.ascontiguousarray(v, dtype=np.float32)
        keep.append(a)
        arr[i] = SatTensor(k.encode(), a.ctypes.data, a.size)
    return arr, len(weights), keep


def _mem_dims(hp):
    from .weights import _mem_dims as md
    d = md(hp)
    return d if len(d) == 2 else d + [0]


def make_hparams(hp):
    s = SatHParams()
    for n in ("num_mels", "outputs_per_step", "n_feed_frame", "num_symbols", "embedding_dim", "cbhg_out_units",
              "conv_channels", "max_filter_width", "projection1_out_channels", "projection2_out_channels",
              "num_highway", "self_attention_out_units", "self_attention_num_heads", "self_attention_num_hop",
              "attention_out_units", "decoder_out_units", "attention_kernel", "attention_filters",
              "decoder_self_attention_out_units", "decoder_self_attention_num_heads",
              "decoder_self_attention_num_hop", "max_iters"):
        setattr(s, n, int(getattr(hp, n)))
    from .weights import speaker_dim
    s.speaker_embedding_dim = int(speaker_dim(hp))
    if hp.encoder not in SAT_ENC:
        raise ValueError(f"Unknown encoder: {hp.encoder}")
    if hp.decoder not in SAT_DEC:
        raise ValueError(f"Unknown decoder: {hp.decoder}")
    if hp.decoder_version != "v2":
        raise ValueError("only decoder_version='v2' (DecoderRNNV2) is implemented")
    s.encoder_kind = SAT_ENC[hp.encoder]
    s.decoder_kind = SAT_DEC[hp.decoder]
    s.encoder_prenet_units = (C.c_int32 * 2)(*hp.encoder_prenet_out_units)
    s.decoder_prenet_units = (C.c_int32 * 2)(*hp.decoder_prenet_out_units)
    for k in (hp.attention, hp.attention2):
        if k not in SAT_ATT:
            raise ValueError(f"Unknown attention mechanism: {k}")
    md = _mem_dims(hp)
    if hp.decoder == "ExtendedDecoder":
        s.attention_kind = (C.c_int32 * 2)(SAT_ATT[hp.attention], 0)
        s.attention_units = (C.c_int32 * 2)(hp.attention_out_units, 0)
        s.memory_dim = (C.c_int32 * 2)(md[0], 0)
    else:
        s.attention_kind = (C.c_int32 * 2)(SAT_ATT[hp.attention], SAT_ATT[hp.attention2])
        s.attention_units = (C.c_int32 * 2)(hp.attention1_out_units, hp.attention2_out_units)
        s.memory_dim = (C.c_int32 * 2)(*md)
    s.cumulative_weights = int(hp.cumulative_weights)
    s.use_transition_agent = int(hp.use_forward_attention_transition_agent)
    s.multi_speaker_prenet = int(bool(hp.use_speaker_embedding and hp.speaker_embedd_to_prenet))
    s.num_postnet_layers = int(hp.num_postnet_v2_layers)
    s.postnet_kernel_size = int(hp.postnet_v2_kernel_size)
    s.postnet_channels = int(hp.postnet_v2_out_channels)
    s.zoneout_factor_cell = float(hp.zoneout_factor_cell)
    s.zoneout_factor_output = float(hp.zoneout_factor_output)
    return s


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def ptr(t):
    """device pointer of a torch tensor (or None)."""
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
