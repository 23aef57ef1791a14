"""Parity checkers at the BASELINE.json shapes (test infrastructure, like the rest of ``oracle/``).

* ``synthesize_rows_parallel`` runs ``oracle.postnet.synthesize`` over row chunks of a batch in forked worker
  processes.  Every utterance row of the synthesis path is independent of the others (batch norm is in inference
  mode; the only cross-row operation is the ``reduce_all`` of the stop rule, modules/helpers.py:156-160, which
  ``stop_rule`` below applies afterwards to the gathered stop logits), so the chunked run is the same arithmetic.
* ``free_run_report`` compares a free-running CUDA result with the oracle by the gates of BASELINE.json's
  north_star: identical number of steps, identical per-row stop step, identical alignment argmax over the first
  N = min(stop, 200) steps, mel within 1e-3 relative on those frames -- and reports the decision margins (distance
  of the deciding stop logit from 0, top-1 minus top-2 alignment weight) so that an "identical" verdict is not luck.
"""
import multiprocessing as mp
import os

import numpy as np

MIN_ITERS = 10  # StopTokenBasedInferenceHelper(min_iters=10), modules/helpers.py:113
CHECK_STEPS = (10, 50, 100, 250, 500, 1000)


def _shrink(x):
    """large float64 arrays travel back to the parent as float32 (relative 6e-8, far below the 1e-3 bar)"""
    if isinstance(x, np.ndarray):
        return x.astype(np.float32) if x.dtype == np.float64 and x.size > (1 << 16) else x
    if isinstance(x, dict):
        return {k: _shrink(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_shrink(v) for v in x)
    return x


def _merge(parts, sizes):
    """concatenate per-chunk results along the batch axis (arrays whose leading axis is the chunk's batch)"""
    p0 = parts[0]
    if isinstance(p0, np.ndarray):
        if p0.ndim >= 1 and all(p.shape[0] == n for p, n in zip(parts, sizes)):
            return np.concatenate(parts, axis=0)
        return p0
    if isinstance(p0, dict):
        return {k: _merge([p[k] for p in parts], sizes) for k in p0}
    if isinstance(p0, (list, tuple)):
        return type(p0)(_merge([p[i] for p in parts], sizes) for i in range(len(p0)))
    return p0


def _worker(conn, fn, kwargs):
    try:
        # one BLAS / OpenMP thread per worker: the parallelism is over rows, and N workers x N BLAS threads thrash
        # (setting OMP_NUM_THREADS here would be too late -- numpy is already loaded in the forked child)
        try:
            from threadpoolctl import threadpool_limits
            limit = threadpool_limits(limits=1)
        except ImportError:  # pragma: no cover
            limit = None
        conn.send(("ok", _shrink(fn(**kwargs))))
        del limit
    except BaseException as e:  # noqa: BLE001 -- reported to the parent
        import traceback
        conn.send(("err", f"{e!r}\n{traceback.format_exc()}"))
    finally:
        conn.close()


def run_rows_parallel(fn, row_kwargs, shared_kwargs, B, nproc=None, drop=("sa_probs_steps", "sources")):
    """fn(**row_kwargs[rows], **shared_kwargs) over contiguous row chunks in forked processes; results merged on
    axis 0.  row_kwargs: {name: array with leading axis B (or None)}."""
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        cores = os.cpu_count() or 1
    nproc = max(1, min(nproc or cores, B))
    bounds = np.linspace(0, B, nproc + 1).astype(int)
    chunks = [(int(a), int(b)) for a, b in zip(bounds[:-1], bounds[1:]) if b > a]
    ctx = mp.get_context("fork")
    procs = []
    for a, b in chunks:
        kw = dict(shared_kwargs)
        kw.update({k: (None if v is None else v[a:b]) for k, v in row_kwargs.items()})
        parent, child = ctx.Pipe(duplex=False)
        pr = ctx.Process(target=_worker, args=(child, fn, kw), daemon=True)
        pr.start()
        child.close()
        procs.append((pr, parent))
    parts = []
    for pr, parent in procs:
        status, payload = parent.recv()
        pr.join()
        if status != "ok":
            raise RuntimeError("oracle worker failed: " + payload)
        for k in drop:
            payload.pop(k, None)
        parts.append(payload)
    return _merge(parts, [b - a for a, b in chunks])


def _dense_sa_probs(steps, S):
    """list over steps of list over hops of [B,H,t+1] -> list over hops of [B,H,S,S] (zero above the diagonal)"""
    hops = len(steps[0])
    out = []
    for h in range(hops):
        B, H = steps[0][h].shape[:2]
        p = np.zeros((B, H, S, S), dtype=np.float32)
        for t, st in enumerate(steps):
            p[:, :, t, :t + 1] = st[h]
        out.append(p)
    return out


def synthesize_rows_parallel(ids, lengths, W64, hp, *, target=None, speaker_ids=None, nproc=None,
                             dense_sa_probs=False, **kw):
    """oracle.postnet.synthesize on row chunks.  Free-running modes must use stop_mode='never' (the cross-row stop
    rule is applied afterwards by ``stop_rule``); teacher-forced / validation / forced-alignment modes need equal
    target lengths for all rows, which every BASELINE config has."""
    from .postnet import synthesize
    if kw.get("mode", "free_run") == "free_run":
        assert kw.get("stop_mode") == "never", "row-parallel free run needs stop_mode='never' + stop_rule()"
    T = ids.shape[1]

    def fn(ids, lengths, target, speaker_ids):
        res = synthesize(ids, lengths, W64, hp, target=target, speaker_ids=speaker_ids, **kw)
        if dense_sa_probs and res.get("sa_probs_steps"):
            res["decoder_sa_probs"] = _dense_sa_probs(res["sa_probs_steps"], res["n_steps"])
        return res

    res = run_rows_parallel(fn, dict(ids=ids, lengths=lengths, target=target, speaker_ids=speaker_ids), {}, ids.shape[0],
                            nproc)
    assert res["mel"].shape[0] == ids.shape[0] and T == ids.shape[1]
    return res


def stop_rule(stop_logits, mode, max_iters, min_iters=MIN_ITERS):
    """modules/helpers.py:156-160 on gathered logits [B,S]: -> (n_steps, stop_steps[B], margin[B]).
    stop_steps[b] = first t > min_iters with sigmoid(logit) > 0.5 (-1: never); mode 'all_rows' ends the batch at the
    first t where every row meets the rule at that t, 'per_row' when every row has met it once.  margin[b] = the
    smallest |logit| among the steps that decide the row's stop step (all steps min_iters < t <= stop step)."""
    logit = np.asarray(stop_logits, dtype=np.float64).reshape(stop_logits.shape[0], -1)
    B, S = logit.shape
    t = np.arange(S)[None, :]
    term = (logit > 0) & (t > min_iters)
    stop_steps = np.where(term.any(1), term.argmax(1), -1)
    if mode == "all_rows":
        allt = term.all(0)
        n_steps = int(allt.argmax()) + 1 if allt.any() else min(S, max_iters)
    elif mode == "per_row":
        n_steps = int(stop_steps.max()) + 1 if (stop_steps >= 0).all() else min(S, max_iters)
    else:
        n_steps = min(S, max_iters)
    stop_steps = np.where(stop_steps >= n_steps, -1, stop_steps)
    margin = np.empty(B)
    for b in range(B):
        last = stop_steps[b] if stop_steps[b] >= 0 else n_steps - 1
        seg = np.abs(logit[b, min_iters + 1:last + 1])
        margin[b] = seg.min() if seg.size else np.inf
    return n_steps, stop_steps.astype(np.int64), margin


def rel_err(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-30))


def free_run_report(got, ref, stop_mode, r, max_iters, first_n=200, tol=1e-3):
    """got: dict(n_steps, stop_steps [B], mel [B,n*r,M], stop [B,n], alignment [B,n,T]) from the CUDA path;
    ref: oracle run with stop_mode='never' on the same weights (mel [B,S*r,M], stop [B,S,1], alignments[0] [B,S,T]).
    Returns a JSON-able report; report['ok'] is the verdict of the north_star free-run gate."""
    ref_stop = np.asarray(ref["stop"]).reshape(ref["stop"].shape[0], -1)
    n_ref, ss_ref, margin = stop_rule(ref_stop, stop_mode, max_iters)
    B = ref_stop.shape[0]
    scale = float(np.max(np.abs(ref_stop[:, :n_ref])))
    decidable = margin > tol * scale  # a stop decision inside the fp tolerance of the logit cannot be pinned
    ss_got = np.asarray(got["stop_steps"]).astype(np.int64)
    rep = {"stop_mode": stop_mode, "rows": int(B), "n_steps": int(got["n_steps"]), "n_steps_oracle": int(n_ref),
           "stop_step_mismatches": int(np.sum(ss_got != ss_ref)),
           "stop_step_mismatches_decidable": int(np.sum((ss_got != ss_ref) & decidable)),
           "rows_stopped": int(np.sum(ss_ref >= 0)),
           "stop_logit_margin_min": float(np.min(margin)) if np.isfinite(np.min(margin)) else None,
           "stop_logit_scale": scale}
    n = min(int(got["n_steps"]), n_ref)
    # per-row number of compared steps: N = min(stop step + 1, first_n, n)
    per_row = np.where(ss_ref >= 0, ss_ref + 1, n)
    if stop_mode == "per_row":
        # a row whose stop decision is inside the fp tolerance may stop a step apart in the two runs; a retired row emits
        # nothing after ITS stop, so the compared window ends at the earlier of the two
        per_row = np.minimum(per_row, np.where(ss_got >= 0, ss_got + 1, n))
    per_row = np.minimum(np.minimum(per_row, first_n), n)
    al_g, al_r = np.asarray(got["alignment"])[:, :n], np.asarray(ref["alignments"][0])[:, :n]
    valid = np.arange(n)[None, :] < per_row[:, None]
    am_g, am_r = al_g.argmax(-1), al_r.argmax(-1)
    part = np.partition(al_r, -2, axis=-1)
    gap = part[..., -1] - part[..., -2]
    mism = (am_g != am_r) & valid
    rep["alignment_argmax_compared"] = int(valid.sum())
    rep["alignment_argmax_mismatches"] = int(mism.sum())
    rep["alignment_argmax_mismatches_decidable"] = int((mism & (gap > tol)).sum())
    rep["alignment_top1_top2_gap_min"] = float(gap[valid].min()) if valid.any() else None
    rep["alignment_top1_top2_gap_median"] = float(np.median(gap[valid])) if valid.any() else None
    rep["alignment_rel_err"] = rel_err(np.where(valid[..., None], al_g, 0), np.where(valid[..., None], al_r, 0))
    M = ref["mel"].shape[-1]
    mel_g = np.asarray(got["mel"])[:, :n * r].reshape(B, n, r * M)
    mel_r = np.asarray(ref["mel"])[:, :n * r].reshape(B, n, r * M)
    rep["mel_rel_err_first_n"] = rel_err(np.where(valid[..., None], mel_g, 0), np.where(valid[..., None], mel_r, 0))
    # every step a row emits: all n under the all_rows / never rules, up to its own stop step under the per-row rule
    # (a retired row writes nothing afterwards)
    own = (np.minimum(np.where(ss_ref >= 0, ss_ref + 1, n), np.where(ss_got >= 0, ss_got + 1, n))
           if stop_mode == "per_row" else np.full(B, n))
    emitted = np.arange(n)[None, :] < np.minimum(own, n)[:, None]
    rep["mel_rel_err_all_steps"] = rel_err(np.where(emitted[..., None], mel_g, 0), np.where(emitted[..., None], mel_r, 0))
    rep["stop_rel_err"] = rel_err(np.where(emitted, np.asarray(got["stop"])[:, :n], 0), np.where(emitted, ref_stop[:, :n], 0))
    ref_max = float(np.max(np.abs(mel_r)))
    rep["mel_rel_err_at_step"] = {str(s): float(np.max(np.abs(mel_g[:, s - 1] - mel_r[:, s - 1])) / ref_max)
                                  for s in CHECK_STEPS if s <= n}
    rep["tolerance"] = tol
    rep["ok"] = bool(rep["n_steps"] == rep["n_steps_oracle"] and rep["stop_step_mismatches_decidable"] == 0
                     and rep["alignment_argmax_mismatches_decidable"] == 0 and rep["mel_rel_err_first_n"] <= tol)
    return rep


def spread_stop_bias(ref_stop_logits, target_median, min_iters=MIN_ITERS):
    """The stop logit does not feed back, so one never-stopping run predicts every row's stop step for any shift of
    the stop-projection bias.  -> the shift (to ADD to the bias) whose median first crossing is nearest
    `target_median` and whose decision margins are widest among near-equal candidates."""
    logit = np.asarray(ref_stop_logits, dtype=np.float64).reshape(ref_stop_logits.shape[0], -1)
    S = logit.shape[1]
    best, best_key = 0.0, None
    for shift in np.quantile(-logit[:, min_iters + 1:], np.linspace(0.0, 0.3, 121)):
        n, ss, margin = stop_rule(logit + shift, "per_row", S, min_iters)
        med = np.median(np.where(ss < 0, S, ss))
        key = (abs(med - target_median) > 25, -float(np.min(margin)))
        if best_key is None or key < best_key:
            best, best_key = float(shift), key
    return best
