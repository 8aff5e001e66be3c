"""oracle/restate.py — TEST INFRASTRUCTURE, NOT PRODUCT.

numpy-facing wrapper over oracle/restate.c (the plain-C CPU restatement of the reference's
learner hot path).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module.  Build: `make -C oracle restate` (done by __graft_entry__.build()).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "librestate.so")
_lib = None

f32p = C.POINTER(C.c_float)
u8p = C.POINTER(C.c_uint8)
i64p = C.POINTER(C.c_int64)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "restate"])


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "restate.c")
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
            build()
        _lib = C.CDLL(_SO)
        _lib.rs_clip_adam.restype = C.c_float
        _lib.rs_param_count.restype = C.c_int64
        _lib.rs_param_count.argtypes = [C.c_int64, C.c_int64]
    return _lib


def _f(a):
    return None if a is None else a.ctypes.data_as(f32p)


def _u(a):
    return None if a is None else a.ctypes.data_as(u8p)


def _i(a):
    return None if a is None else a.ctypes.data_as(i64p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


class InvalidArgument(ValueError):
    pass


def _check(rc):
    if rc == 1:
        raise InvalidArgument("invalid argument")
    if rc == 2:
        raise RuntimeError("runtime error")
    if rc:
        raise RuntimeError(f"oracle error {rc}")


def gae(rewards, values, next_values, terminals, truncations, starts, gamma, lam):
    r, v, nv = _c(rewards, np.float32), _c(values, np.float32), _c(next_values, np.float32)
    te, tr, st = (_c(x, np.uint8) for x in (terminals, truncations, starts))
    E, T = r.shape
    adv = np.zeros((E, T), np.float32)
    _check(lib().rs_gae(_f(adv), _f(r), _f(v), _f(nv), _u(te), _u(tr), _u(st), C.c_int64(E),
                        C.c_int64(T), C.c_float(gamma), C.c_float(lam)))
    return adv


def buffer_get(rewards, values, next_values, terminals, truncations, starts, gamma, lam,
               write_index=0):
    r = _c(rewards, np.float32).copy()
    v, nv = _c(values, np.float32), _c(next_values, np.float32)
    te, tr, st = (_c(x, np.uint8) for x in (terminals, truncations, starts))
    E, T = r.shape
    adv, ret = np.zeros((E, T), np.float32), np.zeros((E, T), np.float32)
    masks = np.zeros((E, T), np.uint8)
    _check(lib().rs_buffer_get(_f(r), _f(v), _f(nv), _u(te), _u(tr), _u(st), C.c_int64(E),
                               C.c_int64(T), C.c_float(gamma), C.c_float(lam),
                               C.c_int64(write_index), _f(adv), _f(ret), _u(masks)))
    return dict(rewards=r, advantages=adv, returns=ret, masks=masks)


def preprocess(frames_a, frames_b):
    fa, fb = _c(frames_a, np.uint8), _c(frames_b, np.uint8)
    B = fa.shape[0]
    g = np.zeros((B, 84, 84), np.float32)
    q = np.zeros((B, 84, 84), np.uint8)
    _check(lib().rs_preprocess(_u(fa), _u(fb), C.c_int64(B), _f(g), _u(q)))
    return g, q


def stack_update(stack, frames, starts):
    s = _c(stack, np.uint8).copy()
    f, st = _c(frames, np.uint8), _c(starts, np.uint8)
    _check(lib().rs_stack_update(_u(s), _u(f), _u(st), C.c_int64(s.shape[0])))
    return s


def log_softmax(logits):
    z = _c(logits, np.float32)
    out = np.zeros_like(z)
    lib().rs_log_softmax(_f(z), C.c_int64(z.shape[0]), C.c_int64(z.shape[1]), _f(out))
    return out


def loss(logits, old_log_probs, actions, advantages, values, returns, masks, clip, vcoef, ecoef):
    z, ol = _c(logits, np.float32), _c(old_log_probs, np.float32)
    a = _c(actions, np.int64)
    adv, v, ret = (_c(x, np.float32) for x in (advantages, values, returns))
    mk = _c(masks, np.uint8)
    n, A = z.shape
    out = {k: np.zeros(n, np.float32) for k in
           ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dvalues")}
    out["dlogits"] = np.zeros((n, A), np.float32)
    ls = np.zeros(1, np.float32)
    _check(lib().rs_loss(_f(z), _f(ol), _i(a), _f(adv), _f(v), _f(ret), _u(mk), C.c_int64(n),
                         C.c_int64(A), C.c_float(clip), C.c_float(vcoef), C.c_float(ecoef),
                         _f(ls), _f(out["clipped_losses"]), _f(out["value_losses"]),
                         _f(out["entropies"]), _f(out["total_losses"]), _f(out["ratio"]),
                         _f(out["dlogits"]), _f(out["dvalues"])))
    out["loss"] = float(ls[0])
    return out


ACT_STRIDE = 4 * 84 * 84 + 32 * 400 + 64 * 81 + 64 * 49


def param_count(hidden, actions=4):
    return int(lib().rs_param_count(hidden, actions))


def net_forward(params, hidden, obs, actions=4, keep_acts=False):
    p, o = _c(params, np.float32), _c(obs, np.uint8)
    n = o.shape[0]
    logits, values = np.zeros((n, actions), np.float32), np.zeros(n, np.float32)
    acts = np.zeros((n, ACT_STRIDE + hidden), np.float32) if keep_acts else None
    _check(lib().rs_net_forward(_f(p), C.c_int64(hidden), C.c_int64(actions), _u(o), C.c_int64(n),
                                _f(logits), _f(values), _f(acts)))
    return (logits, values, acts) if keep_acts else (logits, values)


def net_backward(params, hidden, acts, dlogits, dvalues, actions=4):
    p = _c(params, np.float32)
    dl, dv = _c(dlogits, np.float32), _c(dvalues, np.float32)
    grads = np.zeros(p.shape[0], np.float32)
    _check(lib().rs_net_backward(_f(p), C.c_int64(hidden), C.c_int64(actions), _f(acts),
                                 C.c_int64(dl.shape[0]), _f(dl), _f(dv), _f(grads)))
    return grads


def clip_adam(params, grads, m, v, hidden, step, lr, max_norm, actions=4):
    """In-place on copies; returns (params, clipped grads, m, v, pre-clip total norm)."""
    p, g, m_, v_ = (_c(x, np.float32).copy() for x in (params, grads, m, v))
    tn = lib().rs_clip_adam(_f(p), _f(g), _f(m_), _f(v_), C.c_int64(hidden), C.c_int64(actions),
                            C.c_int64(step), C.c_double(lr), C.c_double(max_norm))
    return p, g, m_, v_, float(tn)


def train(params, m, v, step, hidden, obs, actions_taken, log_probs, advantages, returns, masks,
          epochs, nmb, clip, vcoef, ecoef, max_norm, lr, actions=4):
    p, m_, v_ = (_c(x, np.float32).copy() for x in (params, m, v))
    o, a = _c(obs, np.uint8), _c(actions_taken, np.int64)
    lp, adv, ret = (_c(x, np.float32) for x in (log_probs, advantages, returns))
    mk = _c(masks, np.uint8)
    N = o.shape[0]
    if nmb <= 0 or N % nmb:
        raise RuntimeError("Batch size must be divisible by num_mini_batches")
    mb = N // nmb
    met = {k: np.zeros((epochs, nmb, mb), np.float32) for k in
           ("loss", "clipped_losses", "value_losses", "entropies", "ratio", "total_losses",
            "advantages", "returns")}
    met["masks"] = np.zeros((epochs, nmb, mb), np.uint8)
    met["clipped_gradients"] = np.zeros((epochs, nmb), np.float32)
    st = C.c_int64(step)
    _check(lib().rs_train(_f(p), _f(m_), _f(v_), C.byref(st), C.c_int64(hidden), C.c_int64(actions),
                          _u(o), _i(a), _f(lp), _f(adv), _f(ret), _u(mk), C.c_int64(N),
                          C.c_int64(epochs), C.c_int64(nmb), C.c_float(clip), C.c_float(vcoef),
                          C.c_float(ecoef), C.c_float(max_norm), C.c_double(lr), _f(met["loss"]),
                          _f(met["clipped_losses"]), _f(met["value_losses"]), _f(met["entropies"]),
                          _f(met["ratio"]), _f(met["total_losses"]), _f(met["advantages"]),
                          _f(met["returns"]), _u(met["masks"]), _f(met["clipped_gradients"])))
    return p, m_, v_, int(st.value), met


def act_sample(logits, uniforms):
    """The action selector of src/bin/train.cc:373-375 — probabilities = softmax(logits),
    actions = multinomial(probabilities, 1) — restated as the inverse CDF of a supplied u in
    [0,1): action = first a with cumsum(p)[a] > u (torch::multinomial's own Philox stream is not
    reproducible, so the draw is parameterised by u).  Returns (actions, cumsum(p)) in fp32."""
    z = np.asarray(logits, np.float32)
    u = np.asarray(uniforms, np.float32)
    e = np.exp(z - z.max(axis=1, keepdims=True), dtype=np.float32)
    p = (e / e.sum(axis=1, keepdims=True, dtype=np.float32)).astype(np.float32)
    cum = np.cumsum(p, axis=1, dtype=np.float32)
    hit = cum > u[:, None]
    actions = np.where(hit.any(axis=1), hit.argmax(axis=1), z.shape[1] - 1).astype(np.int64)
    return actions, cum


def normalize_advantages(advantages, masks):
    """Opt-in extra (not in the reference; BASELINE.json north_star): per-minibatch
    A' = (A - mean) / (std + 1e-8), mean and UNBIASED std (torch .std()) over the unmasked rows."""
    a = np.asarray(advantages, np.float64)
    m = np.asarray(masks).astype(bool)
    mean, std = a[m].mean(), a[m].std(ddof=1)
    return ((a - mean) / (std + 1e-8)).astype(np.float32)


def gather_minibatches(indices, *arrays):
    """Opt-in extra: the shuffled row order train() consumes (rows indices[...] of every array)."""
    idx = np.asarray(indices, np.int64)
    return tuple(np.ascontiguousarray(np.asarray(x)[idx]) for x in arrays)

