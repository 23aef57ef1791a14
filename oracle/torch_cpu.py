"""torch-CPU float32 port of the oracle, used ONLY as the timed CPU baseline (bench.py ``cpu_baseline`` and
``--impl reference``) -- test infrastructure like the rest of oracle/, never on the product path.

Same graph as oracle/{encoder,decoder,attention,nn}.py (each function cites the same reference lines), written
with torch ops so that every stage uses all host threads (``torch.set_num_threads``); tests/test_oracle_cpu.py
pins it to the NumPy oracle.  The reference's own TF-CPU path cannot run here (TensorFlow 1.x absent), so this
is what "the reference's CPU implementation on the box's host cores" means in every report: kind = "port"."""
import numpy as np
import torch
import torch.nn.functional as F


def _t(W):
    return {k: torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)) for k, v in W.items()}


def _conv1d_bn_act(x, W, p, act):
    """tacotron2 Conv1d (modules/module.py:46-68): SAME conv (no bias) -> BN(inference) -> act."""
    k = W[p + "/kernel"]  # [k, in, out]
    ks = k.shape[0]
    pl = (ks - 1) // 2
    xp = F.pad(x.transpose(1, 2), (pl, ks - 1 - pl))
    y = F.conv1d(xp, k.permute(2, 1, 0)).transpose(1, 2)
    scale = W[p + "/bn/gamma"] / torch.sqrt(W[p + "/bn/var"] + 1e-3)
    y = y * scale + (W[p + "/bn/beta"] - W[p + "/bn/mean"] * scale)
    return act(y) if act is not None else y


def _zoneout_lstm(x, state, k, b, zc, zh):
    """tacotron2 ZoneoutLSTMCell, inference form (A.4): output = un-zoned h, state = zoned (c, h)."""
    c_prev, h_prev = state
    z = torch.cat([x, h_prev], dim=-1) @ k + b
    i, j, f, o = z.chunk(4, dim=-1)
    c = torch.sigmoid(f + 1.0) * c_prev + torch.sigmoid(i) * torch.tanh(j)
    h = torch.sigmoid(o) * torch.tanh(c)
    return h, (c - zc * (c - c_prev), h - zh * (h - h_prev))


def _bilstm(x, lengths, W, p, zc, zh):
    """tf.nn.bidirectional_dynamic_rnn with sequence_length (modules/module.py:93-108; A.12)."""
    B, T, _ = x.shape
    U = W[p + "/lstm_fw/bias"].shape[0] // 4
    lens = torch.as_tensor(lengths)
    idx = torch.arange(T)[None, :]
    rev = torch.where(idx < lens[:, None], lens[:, None] - 1 - idx, idx)  # reverse each row by its own length

    def run(xs, name):
        k, b = W[f"{p}/{name}/kernel"], W[f"{p}/{name}/bias"]
        c = torch.zeros(B, U)
        h = torch.zeros(B, U)
        outs = []
        for t in range(T):
            out, (c2, h2) = _zoneout_lstm(xs[:, t], (c, h), k, b, zc, zh)
            live = (t < lens)[:, None]
            outs.append(torch.where(live, out, torch.zeros_like(out)))
            c = torch.where(live, c2, c)
            h = torch.where(live, h2, h)
        return torch.stack(outs, dim=1)

    gather = lambda a: torch.gather(a, 1, rev[:, :, None].expand(-1, -1, a.shape[2]))  # noqa: E731
    return torch.cat([run(x, "lstm_fw"), gather(run(gather(x), "lstm_bw"))], dim=-1)


def _mha(x, W, p, H, causal):
    """modules/self_attention.py:108-128, 45-65 (padding mask off, as every caller leaves it)."""
    B, T, _ = x.shape
    k = x @ W[p + "/key/kernel"] + W[p + "/key/bias"]
    v = x @ W[p + "/value/kernel"] + W[p + "/value/bias"]
    q = x @ W[p + "/query/kernel"] + W[p + "/query/bias"]
    D = k.shape[-1]
    dh = D // H
    hd = lambda a: a.reshape(B, T, H, dh).transpose(1, 2)  # noqa: E731
    s = hd(q) @ hd(k).transpose(2, 3) / np.sqrt(dh)
    if causal:
        s = s.masked_fill(~torch.tril(torch.ones(T, T, dtype=torch.bool)), float("-inf"))
    pr = torch.softmax(s, dim=-1)
    ctx = (pr @ hd(v)).transpose(1, 2).reshape(B, T, D)
    return ctx @ W[p + "/output/kernel"] + W[p + "/output/bias"]


def encoder(ids, lengths, W, hp):
    """Embedding (models/models.py:28,54) -> ZoneoutEncoderV1 (modules/module.py:336-339) or
    SelfAttentionCBHGEncoder (:425-438); ZoneoutCBHG :77-110."""
    x = W["embedding"][torch.as_tensor(ids)]
    for i in range(len(hp.encoder_prenet_out_units)):
        x = torch.relu(x @ W[f"encoder/prenet/dense_{i}/kernel"] + W[f"encoder/prenet/dense_{i}/bias"])
    c = "encoder/cbhg"
    bank = torch.cat([_conv1d_bn_act(x, W, f"{c}/conv1d_K{k}", torch.relu) for k in range(1, hp.max_filter_width + 1)], -1)
    pooled = bank.clone()
    pooled[:, :-1] = torch.maximum(bank[:, :-1], bank[:, 1:])
    y = _conv1d_bn_act(_conv1d_bn_act(pooled, W, c + "/proj1", torch.relu), W, c + "/proj2", None) + x
    if y.shape[2] != hp.cbhg_out_units // 2:
        y = y @ W[c + "/adjustment/kernel"] + W[c + "/adjustment/bias"]
    for i in range(hp.num_highway):
        h = torch.relu(y @ W[f"{c}/highway_{i}/H/kernel"] + W[f"{c}/highway_{i}/H/bias"])
        t = torch.sigmoid(y @ W[f"{c}/highway_{i}/T/kernel"] + W[f"{c}/highway_{i}/T/bias"])
        y = h * t + y * (1.0 - t)
    lstm_out = _bilstm(y, lengths, W, c, hp.zoneout_factor_cell, hp.zoneout_factor_output)
    if hp.encoder != "SelfAttentionCBHGEncoder":
        return (lstm_out,)
    sa = lstm_out @ W["encoder/sa_projection/kernel"] + W["encoder/sa_projection/bias"]
    for i in range(hp.self_attention_num_hop):
        a = _mha(sa, W, f"encoder/sa_{i}/mha", hp.self_attention_num_heads, False)
        sa = sa + torch.tanh(a @ W[f"encoder/sa_{i}/transform/kernel"] + W[f"encoder/sa_{i}/transform/bias"])
    return lstm_out, sa


class _Mech:
    """One attention mechanism: forward (modules/forward_attention.py:88-122), location_sensitive (A.10) or
    additive (TF BahdanauAttention, A.5)."""

    def __init__(self, W, p, kind, memory, lengths, hp):
        self.W, self.p, self.kind, self.hp = W, p, kind, hp
        B, T, _ = memory.shape
        self.mask = torch.arange(T)[None, :] < torch.as_tensor(lengths)[:, None]
        self.values = memory * self.mask[:, :, None]
        self.keys = self.values @ W[p + "/memory_layer/kernel"]
        self.prev = torch.zeros(B, T)
        self.alpha = torch.zeros(B, T)
        self.alpha[:, 0] = 1.0
        self.u = torch.full((B, 1), 0.5)

    def __call__(self, query):
        W, p = self.W, self.p
        x = self.keys + (query @ W[p + "/query_layer/kernel"])[:, None, :]
        if self.kind != "additive":
            k = W[p + "/location_conv/kernel"]  # [K, 1, F]
            K = k.shape[0]
            pl = (K - 1) // 2
            f = F.conv1d(F.pad(self.prev[:, None, :], (pl, K - 1 - pl)), k.permute(2, 1, 0), W[p + "/location_conv/bias"])
            x = x + f.transpose(1, 2) @ W[p + "/location_layer/kernel"] + W[p + "/attention_bias"]
        e = (W[p + "/attention_v"] * torch.tanh(x)).sum(-1)
        a = torch.softmax(e.masked_fill(~self.mask, float("-inf")), dim=-1)
        out = a
        if self.kind == "forward":
            shifted = F.pad(self.alpha[:, :-1], (1, 0))
            al = ((1 - self.u) * self.alpha + self.u * shifted + 1e-7) * a
            out = self.alpha = al / al.sum(1, keepdim=True)
        if self.kind != "additive":
            self.prev = a + self.prev if self.hp.cumulative_weights else a
        return out


def decoder_free_run(sources, lengths, W, hp, n_steps):
    """ExtendedDecoder / DualSourceDecoder / DualSourceTransformerDecoder free run for exactly ``n_steps`` steps
    (modules/module.py:562-623, 1083-1152, 1499-1562; step order SURVEY.md 3.2; KV-cached self-attention)."""
    B = sources[0].shape[0]
    M, r, nf = hp.num_mels, hp.outputs_per_step, hp.n_feed_frame
    zc, zh = hp.zoneout_factor_cell, hp.zoneout_factor_output
    kinds = [hp.attention] if hp.decoder == "ExtendedDecoder" else [hp.attention, hp.attention2]
    mechs = [_Mech(W, f"decoder/attention{i + 1}", k, s, lengths, hp) for i, (k, s) in enumerate(zip(kinds, sources))]
    A, Dd = hp.attention_out_units, hp.decoder_out_units
    z = lambda n: torch.zeros(B, n)  # noqa: E731
    st_a, st_1, st_2 = (z(A), z(A)), (z(Dd), z(Dd)), (z(Dd), z(Dd))
    attention = z(sum(m.values.shape[-1] for m in mechs))
    hops = hp.decoder_self_attention_num_hop if hp.decoder == "DualSourceTransformerDecoder" else 0
    H = hp.decoder_self_attention_num_heads
    cache = [([], []) for _ in range(hops)]
    x = z(M * nf)
    outs = []
    kb = lambda n: (W[n + "/kernel"], W[n + "/bias"])  # noqa: E731
    for _ in range(n_steps):
        p = x
        for i in range(len(hp.decoder_prenet_out_units)):
            k, b = kb(f"decoder/prenet/dense_{i}")
            p = torch.relu(p @ k + b)
        query, st_a = _zoneout_lstm(torch.cat([p, attention], -1), st_a, *kb("decoder/attention_lstm"), zc, zh)
        attention = torch.cat([torch.einsum("bt,btd->bd", m(query), m.values) for m in mechs], -1)
        o1, st_1 = _zoneout_lstm(torch.cat([query, attention], -1), st_1, *kb("decoder/lstm_1"), zc, zh)
        y, st_2 = _zoneout_lstm(o1, st_2, *kb("decoder/lstm_2"), zc, zh)
        for i in range(hops):
            pm = f"decoder/sa_{i}/mha"
            cache[i][0].append(y @ W[pm + "/key/kernel"] + W[pm + "/key/bias"])
            cache[i][1].append(y @ W[pm + "/value/kernel"] + W[pm + "/value/bias"])
            q = y @ W[pm + "/query/kernel"] + W[pm + "/query/bias"]
            Kc, Vc = torch.stack(cache[i][0], 1), torch.stack(cache[i][1], 1)
            t, D = Kc.shape[1], Kc.shape[2]
            dh = D // H
            s = torch.einsum("bhd,bthd->bht", q.reshape(B, H, dh), Kc.reshape(B, t, H, dh)) / np.sqrt(dh)
            ctx = torch.einsum("bht,bthd->bhd", torch.softmax(s, -1), Vc.reshape(B, t, H, dh)).reshape(B, D)
            a = ctx @ W[pm + "/output/kernel"] + W[pm + "/output/bias"]
            k, b = kb(f"decoder/sa_{i}/transform")
            y = y + torch.tanh(a @ k + b)
        out = y @ W["decoder/out_projection/kernel"] + W["decoder/out_projection/bias"]
        outs.append(out)
        x = out[:, -M * nf:]
    return torch.stack(outs, 1).reshape(B, n_steps * r, M)


def synthesize_free_run(ids, lengths, W, hp, n_steps, threads=None):
    """Whole path (encoder + n_steps decoder steps) in torch-CPU fp32; W is the float32 numpy blob."""
    if threads:
        torch.set_num_threads(threads)
    with torch.no_grad():
        Wt = _t(W) if not isinstance(next(iter(W.values())), torch.Tensor) else W
        src = encoder(ids, lengths, Wt, hp)
        return decoder_free_run(src, lengths, Wt, hp, n_steps)
