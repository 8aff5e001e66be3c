"""oracle/ref.py — TEST INFRASTRUCTURE, NOT PRODUCT.

ctypes binding of oracle/_ref/libref_oracle.so: the reference's OWN sources (gae.cc, buffer.cc,
vision.cc, ppo/losses.cc, ppo/train.{h,cc}) compiled unmodified against libtorch-CPU, behind
oracle/ref_harness.cc.  Used to (a) validate oracle/restate.c, (b) generate tests/golden/, and
(c) time the reference's CPU path (`bench.py --impl reference`, cpu_baseline kind "reference").
The library is built by `make -C oracle ref` in the container that has /root/reference; on the
GPU box only the prebuilt .so exists.  `available()` tells which.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libref_oracle.so")
_lib = None

f32p, u8p, i64p = C.POINTER(C.c_float), C.POINTER(C.c_uint8), C.POINTER(C.c_int64)


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        import torch  # noqa: F401  (loads libtorch so the rpath'd deps resolve identically)
        _lib = C.CDLL(SO)
        _lib.ref_last_error.restype = C.c_char_p
        _lib.ref_learner_create.restype = C.c_void_p
        _lib.ref_learner_create.argtypes = [C.c_int64, C.c_int64, C.c_uint64, C.c_double]
        _lib.ref_learner_param_count.restype = C.c_int64
        _lib.ref_learner_tensor_count.restype = C.c_int64
        for name in ("ref_learner_destroy", "ref_learner_param_count", "ref_learner_tensor_count"):
            getattr(_lib, name).argtypes = [C.c_void_p]
    return _lib


def _f(a):
    return None if a is None else a.ctypes.data_as(f32p)


def _u(a):
    return None if a is None else a.ctypes.data_as(u8p)


def _i(a):
    return None if a is None else a.ctypes.data_as(i64p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def _check(rc):
    if rc == 0:
        return
    msg = lib().ref_last_error().decode()
    if rc == 1:
        raise ValueError(msg)  # std::invalid_argument
    raise RuntimeError(msg)    # std::runtime_error / other


def set_num_threads(n):
    lib().ref_set_num_threads(int(n))


def get_num_threads():
    return int(lib().ref_get_num_threads())


def gae(rewards, values, next_values, terminals, truncations, starts, gamma, lam):
    r, v, nv = _c(rewards, np.float32), _c(values, np.float32), _c(next_values, np.float32)
    te, tr, st = (_c(x, np.uint8) for x in (terminals, truncations, starts))
    E, T = r.shape
    adv = np.zeros((E, T), np.float32)
    _check(lib().ref_gae(_f(adv), _f(r), _f(v), _f(nv), _u(te), _u(tr), _u(st), C.c_int64(E),
                         C.c_int64(T), C.c_float(gamma), C.c_float(lam)))
    return adv


def buffer_roundtrip(actions, rewards, terminals, truncations, starts, logits, values,
                     next_values, gamma, lam, obs=None, steps_added=None):
    a = _c(actions, np.int64)
    r, lg, v, nv = (_c(x, np.float32) for x in (rewards, logits, values, next_values))
    te, tr, st = (_c(x, np.uint8) for x in (terminals, truncations, starts))
    E, T = r.shape
    A = lg.shape[2]
    o = None if obs is None else _c(obs, np.uint8)
    out = dict(rewards=np.zeros((E, T), np.float32), masks=np.zeros((E, T), np.uint8),
               advantages=np.zeros((E, T), np.float32), returns=np.zeros((E, T), np.float32))
    if o is not None:
        out["observations"] = np.zeros_like(o)
    _check(lib().ref_buffer_roundtrip(_u(o), _i(a), _f(r), _u(te), _u(tr), _u(st), _f(lg), _f(v),
                                      _f(nv), C.c_int64(E), C.c_int64(T), C.c_int64(A),
                                      C.c_float(gamma), C.c_float(lam),
                                      C.c_int64(T if steps_added is None else steps_added),
                                      _u(out.get("observations")), _f(out["rewards"]),
                                      _u(out["masks"]), _f(out["advantages"]), _f(out["returns"])))
    return out


def preprocess(frames_a, frames_b):
    fa, fb = _c(frames_a, np.uint8), _c(frames_b, np.uint8)
    B = fa.shape[0]
    g, q = np.zeros((B, 84, 84), np.float32), np.zeros((B, 84, 84), np.uint8)
    _check(lib().ref_preprocess(_u(fa), _u(fb), C.c_int64(B), _f(g), _u(q)))
    return g, q


def resize_gray_stacked(images):
    x = _c(images, np.float32)
    B, F = x.shape[:2]
    out = np.zeros((B, F, 84, 84), np.float32)
    _check(lib().ref_resize_gray_stacked(_f(x), C.c_int64(B), C.c_int64(F), _f(out)))
    return out


def stack_update(stack, frames, starts):
    s = _c(stack, np.uint8).copy()
    f, st = _c(frames, np.uint8), _c(starts, np.uint8)
    _check(lib().ref_stack_update(_u(s), _u(f), _u(st), C.c_int64(s.shape[0])))
    return s


def normalize_logits(logits):
    z = _c(logits, np.float32)
    out = np.zeros_like(z)
    _check(lib().ref_normalize_logits(_f(z), C.c_int64(z.shape[0]), C.c_int64(z.shape[1]), _f(out)))
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
    _check(lib().ref_loss(_f(z), _f(ol), _i(a), _f(adv), _f(v), _f(ret), _u(mk), C.c_int64(n),
                          C.c_int64(A), C.c_float(clip), C.c_float(vcoef), C.c_float(ecoef), _f(ls),
                          _f(out["clipped_losses"]), _f(out["value_losses"]), _f(out["entropies"]),
                          _f(out["total_losses"]), _f(out["ratio"]), _f(out["dlogits"]),
                          _f(out["dvalues"])))
    out["loss"] = float(ls[0])
    return out


class Learner:
    """The reference Network (restated) + torch::optim::Adam(lr, eps=1e-5) + reference train()."""

    def __init__(self, hidden=512, actions=4, seed=42, lr=2.5e-4):
        self.h = lib().ref_learner_create(hidden, actions, seed, lr)
        if not self.h:
            raise RuntimeError(lib().ref_last_error().decode())
        self.h = C.c_void_p(self.h)
        self.hidden, self.actions = hidden, actions
        self.P = int(lib().ref_learner_param_count(self.h))

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_learner_destroy(self.h)
            self.h = None

    def tensor_sizes(self):
        n = int(lib().ref_learner_tensor_count(self.h))
        out = np.zeros(n, np.int64)
        lib().ref_learner_tensor_sizes(self.h, _i(out))
        return out

    def get_params(self):
        out = np.zeros(self.P, np.float32)
        _check(lib().ref_learner_get_params(self.h, _f(out)))
        return out

    def set_params(self, p):
        _check(lib().ref_learner_set_params(self.h, _f(_c(p, np.float32))))

    def set_lr(self, lr):
        lib().ref_learner_set_lr(self.h, C.c_double(lr))

    def get_adam(self):
        m, v = np.zeros(self.P, np.float32), np.zeros(self.P, np.float32)
        st = C.c_int64(0)
        _check(lib().ref_learner_get_adam(self.h, _f(m), _f(v), C.byref(st)))
        return m, v, int(st.value)

    def forward(self, obs):
        o = _c(obs, np.uint8)
        n = o.shape[0]
        lg, v = np.zeros((n, self.actions), np.float32), np.zeros(n, np.float32)
        _check(lib().ref_learner_forward(self.h, _u(o), C.c_int64(n), _f(lg), _f(v)))
        return lg, v

    def activations(self, obs):
        o = _c(obs, np.uint8)
        n = o.shape[0]
        c1 = np.zeros((n, 32, 20, 20), np.float32)
        c2 = np.zeros((n, 64, 9, 9), np.float32)
        c3 = np.zeros((n, 64, 7, 7), np.float32)
        fc = np.zeros((n, self.hidden), np.float32)
        _check(lib().ref_learner_activations(self.h, _u(o), C.c_int64(n), _f(c1), _f(c2), _f(c3), _f(fc)))
        return c1, c2, c3, fc

    def grads(self, obs, actions, old_log_probs, advantages, returns, masks, clip, vcoef, ecoef):
        o, a = _c(obs, np.uint8), _c(actions, np.int64)
        ol, adv, ret = (_c(x, np.float32) for x in (old_log_probs, advantages, returns))
        mk = _c(masks, np.uint8)
        n = o.shape[0]
        g = np.zeros(self.P, np.float32)
        ls = np.zeros(1, np.float32)
        lg, v = np.zeros((n, self.actions), np.float32), np.zeros(n, np.float32)
        _check(lib().ref_learner_grads(self.h, _u(o), _i(a), _f(ol), _f(adv), _f(ret), _u(mk),
                                       C.c_int64(n), C.c_int64(self.actions), C.c_float(clip),
                                       C.c_float(vcoef), C.c_float(ecoef), _f(g), _f(ls), _f(lg), _f(v)))
        return g, float(ls[0]), lg, v

    def apply_grads(self, grads, max_norm):
        tn = np.zeros(1, np.float32)
        _check(lib().ref_learner_apply_grads(self.h, _f(_c(grads, np.float32)), C.c_double(max_norm), _f(tn)))
        return float(tn[0])

    def train(self, obs, actions, log_probs, advantages, returns, masks, epochs, nmb, clip, vcoef,
              ecoef, max_norm):
        o, a = _c(obs, np.uint8), _c(actions, np.int64)
        lp, adv, ret = (_c(x, np.float32) for x in (log_probs, advantages, returns))
        mk = _c(masks, np.uint8)
        N = o.shape[0]
        mb = N // nmb if nmb > 0 and N % nmb == 0 else 1
        met = {k: np.zeros((epochs, nmb, mb), np.float32) for k in
               ("loss", "clipped_losses", "value_losses", "entropies", "ratio", "total_losses",
                "advantages", "returns")}
        met["masks"] = np.zeros((epochs, nmb, mb), np.uint8)
        met["clipped_gradients"] = np.zeros((epochs, nmb), np.float32)
        _check(lib().ref_learner_train(self.h, _u(o), _i(a), _f(lp), _f(adv), _f(ret), _u(mk),
                                       C.c_int64(N), C.c_int64(self.actions), C.c_int64(epochs),
                                       C.c_int64(nmb), C.c_float(clip), C.c_float(vcoef),
                                       C.c_float(ecoef), C.c_float(max_norm), _f(met["loss"]),
                                       _f(met["clipped_losses"]), _f(met["value_losses"]),
                                       _f(met["entropies"]), _f(met["ratio"]), _f(met["total_losses"]),
                                       _f(met["advantages"]), _f(met["returns"]), _u(met["masks"]),
                                       _f(met["clipped_gradients"])))
        return met

    def iteration(self, frame_pool, actions, rewards, terminals, truncations, starts, logits,
                  values, next_values, epochs, nmb, gamma, lam, clip, vcoef, ecoef, max_norm,
                  preprocess_steps=None):
        fp = _c(frame_pool, np.uint8)
        a = _c(actions, np.int64)
        r, lg, v, nv = (_c(x, np.float32) for x in (rewards, logits, values, next_values))
        te, tr, st = (_c(x, np.uint8) for x in (terminals, truncations, starts))
        E, T = r.shape
        secs = (C.c_double * 3)()
        ml = C.c_float(0)
        _check(lib().ref_iteration(self.h, _u(fp), C.c_int64(fp.shape[0]), _i(a), _f(r), _u(te),
                                   _u(tr), _u(st), _f(lg), _f(v), _f(nv), C.c_int64(E),
                                   C.c_int64(T), C.c_int64(self.actions), C.c_int64(epochs),
                                   C.c_int64(nmb), C.c_float(gamma), C.c_float(lam), C.c_float(clip),
                                   C.c_float(vcoef), C.c_float(ecoef), C.c_float(max_norm),
                                   C.c_int64(T if preprocess_steps is None else preprocess_steps),
                                   secs, C.byref(ml)))
        return list(secs), float(ml.value)
