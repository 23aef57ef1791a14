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
        if getattr(self, "_h", None) and self._h.value:
            lib().sat_decoder_destroy(self._h)
            self._h = C.c_void_p()

    def run(self, memories, lengths, *, target=None, teacher_alignments=None, speaker_embed=None,
            feed_mode, stop_mode, max_steps, forced_alignment=False, min_iters=10,
            want_alignments=True, want_sample_ids=False, want_decoder_outputs=False, stream=None):
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
        mel = torch.empty((B, S * r, M), device=dev, dtype=torch.float32)
        stop = torch.empty((B, S), device=dev, dtype=torch.float32)
        aligns = [torch.empty((B, S, T), device=dev, dtype=torch.float32) if want_alignments else None
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
        check(lib().sat_decoder_run(self._h, C.byref(io), stream_ptr(stream)))
        return {"mel": mel, "stop": stop, "alignments": aligns, "sample_ids": ids, "decoder_outputs": dout,
                "stop_steps": stop_steps, "n_steps": n_steps, "_keep": (memories, lengths, target, tas, speaker_embed)}

    def step_bytes(self, B, T, avg_cache_len=0):
        return lib().sat_decoder_step_bytes(self._h, B, T, avg_cache_len)

    def profile(self, reset=True):
        out = (C.c_int64 * 16)()
        check(lib().sat_decoder_profile(self._h, out, int(reset)))
        return list(out)
