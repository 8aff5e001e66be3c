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


def resize_area(planes):
    """interpolate(size=84x84, mode=area) of src/ai/vision.cc:8-10 over [..., 210, 160] float planes:
    output (oy, ox) = mean of input rows [floor(oy*210/84), ceil((oy+1)*210/84)) x columns
    [floor(ox*160/84), ceil((ox+1)*160/84)), fp32 sum (rows, then columns) / count."""
    x = np.asarray(planes, np.float32)
    lead = x.shape[:-2]
    x = x.reshape(-1, 210, 160)
    out = np.zeros((x.shape[0], 84, 84), np.float32)
    for oy in range(84):
        y0, y1 = (oy * 210) // 84, -((-(oy + 1) * 210) // 84)
        for ox in range(84):
            x0, x1 = (ox * 160) // 84, -((-(ox + 1) * 160) // 84)
            s = np.zeros(x.shape[0], np.float32)
            for y in range(y0, y1):
                for xx in range(x0, x1):
                    s = s + x[:, y, xx]
            out[:, oy, ox] = s / np.float32((y1 - y0) * (x1 - x0))
    return out.reshape(lead + (84, 84))


def rgb_to_grayscale(images):
    """rgb_to_grayscale_frame_stacked_images (src/ai/vision.cc:71-84): [..., 3, 84, 84] -> [..., 84, 84],
    matmul with (0.2125, 0.7154, 0.0721) over the channel axis."""
    x = np.asarray(images, np.float32)
    f = np.float32
    return ((x[..., 0, :, :] * f(0.2125) + x[..., 1, :, :] * f(0.7154)) + x[..., 2, :, :] * f(0.0721)).astype(f)


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


def loss(logits, old_log_probs, actions, advantages, values, returns, masks, clip, vcoef, ecoef,
         old_values=None):
    """old_values: the opt-in clipped value loss (not in the reference; see restate.c rs_loss_ex)."""
    ov = _c(old_values, np.float32) if old_values is not None else None
    z, ol = _c(logits, np.float32), _c(old_log_probs, np.float32)
    a = _c(actions, np.int64)
    adv, v, ret = (_c(x, np.float32) for x in (advantages, values, returns))
    mk = _c(masks, np.uint8)
    n, A = z.shape
    out = {k: np.zeros(n, np.float32) for k in
           ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dvalues")}
    out["dlogits"] = np.zeros((n, A), np.float32)
    ls = np.zeros(1, np.float32)
    _check(lib().rs_loss_ex(_f(z), _f(ol), _i(a), _f(adv), _f(v), _f(ret), _u(mk), C.c_int64(n),
                            C.c_int64(A), C.c_float(clip), C.c_float(vcoef), C.c_float(ecoef),
                            _f(ls), _f(out["clipped_losses"]), _f(out["value_losses"]),
                            _f(out["entropies"]), _f(out["total_losses"]), _f(out["ratio"]),
                            _f(out["dlogits"]), _f(out["dvalues"]),
                            _f(ov) if ov is not None else None))
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
          epochs, nmb, clip, vcoef, ecoef, max_norm, lr, actions=4, old_values=None):
    ov = _c(old_values, np.float32) if old_values is not None else None
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
    _check(lib().rs_train_ex(_f(p), _f(m_), _f(v_), C.byref(st), C.c_int64(hidden), C.c_int64(actions),
                          _u(o), _i(a), _f(lp), _f(adv), _f(ret), _u(mk), C.c_int64(N),
                          C.c_int64(epochs), C.c_int64(nmb), C.c_float(clip), C.c_float(vcoef),
                          C.c_float(ecoef), C.c_float(max_norm), C.c_double(lr), _f(met["loss"]),
                          _f(met["clipped_losses"]), _f(met["value_losses"]), _f(met["entropies"]),
                          _f(met["ratio"]), _f(met["total_losses"]), _f(met["advantages"]),
                          _f(met["returns"]), _u(met["masks"]), _f(met["clipped_gradients"]),
                          _f(ov) if ov is not None else None))
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



# ---- live path: stb_image_resize2's stbir_resize_uint8_linear, restated (PARITY UNPINNED) ------------
# src/ai/vision.cc:86-95 calls the third-party header (stb@0.0.0-20241109-5c20573, MODULE.bazel:38),
# which is not under /root/reference; the reference holds no test of this overload.  The rules below
# are the published algorithm with stb's defaults: per axis Mitchell-Netravali (B = C = 1/3) when the
# axis shrinks — evaluated in output space, support 2 output pixels — Catmull-Rom when it grows,
# clamp edges (outside taps fold onto the edge pixel), weights of each output pixel normalised to 1,
# linear light, u8 <-> float unscaled, result = (u8)clamp(x + .5, 0, 255).
def _mitchell(x):
    x = np.abs(x).astype(np.float32)
    one, two = np.float32(1), np.float32(2)
    near = (np.float32(16) + x * x * (np.float32(21) * x - np.float32(36))) / np.float32(18)
    far = (np.float32(32) + x * (np.float32(-60) + x * (np.float32(36) - np.float32(7) * x))) / np.float32(18)
    return np.where(x < one, near, np.where(x < two, far, np.float32(0))).astype(np.float32)


def _catmull_rom(x):
    x = np.abs(x).astype(np.float32)
    one, two = np.float32(1), np.float32(2)
    near = one - x * x * (np.float32(2.5) - np.float32(1.5) * x)
    far = two - x * (np.float32(4) + x * (np.float32(0.5) * x - np.float32(2.5)))
    return np.where(x < one, near, np.where(x < two, far, np.float32(0))).astype(np.float32)


def stb_axis_matrix(n_in, n_out):
    """Dense [n_out, n_in] fp32 weight matrix of one axis, built in SCATTER form for shrinking axes
    (each input pixel spreads onto the output pixels within 2 of its centre — the way stb forms
    its downsampling coefficients), gather form for growing ones."""
    scale = np.float32(n_out) / np.float32(n_in)
    W = np.zeros((n_out, n_in), np.float32)
    if scale < 1:
        margin = int(np.ceil(2.0 / float(scale))) + 2
        for i in range(-margin, n_in + margin):
            centre = (np.float32(i) + np.float32(0.5)) * scale        # in output space
            lo = int(np.floor(float(centre) - 2.0 - 0.5)) - 1
            hi = int(np.ceil(float(centre) + 2.0 - 0.5)) + 1
            o = np.arange(max(lo, 0), min(hi, n_out - 1) + 1)
            if o.size == 0:
                continue
            x = (o.astype(np.float32) + np.float32(0.5)) - centre
            W[o, min(max(i, 0), n_in - 1)] += _mitchell(x) * scale     # clamp: fold onto the edge
    else:
        for o in range(n_out):
            centre = (np.float32(o) + np.float32(0.5)) / scale        # in input space
            for i in range(int(np.floor(float(centre))) - 3, int(np.floor(float(centre))) + 4):
                w = _catmull_rom(np.float32(centre - (np.float32(i) + np.float32(0.5))))
                W[o, min(max(i, 0), n_in - 1)] += w
    W /= W.sum(axis=1, dtype=np.float32, keepdims=True)
    return W.astype(np.float32)


def resize_grayscale_stb(images, new_width, new_height):
    """images [n,h,w] u8 -> (u8 [n,nh,nw], fp32 value before rounding)."""
    a = np.asarray(images, np.uint8)
    n, h, w = a.shape
    Wv, Wh = stb_axis_matrix(h, new_height), stb_axis_matrix(w, new_width)
    # float64 accumulation of the fp32 weights: the reference value the fp32 kernel is held to
    f = np.einsum("oh,nhw,pw->nop", Wv.astype(np.float64), a.astype(np.float64), Wh.astype(np.float64))
    q = np.clip(f + 0.5, 0.0, 255.0).astype(np.int64).astype(np.uint8)
    return q, f.astype(np.float32)


def preprocess_gray(frames_a, frames_b):
    """live order (rollout.cc:146-161): resize each screen, then the 2-frame max."""
    qa, _ = resize_grayscale_stb(frames_a, 84, 84)
    qb, _ = resize_grayscale_stb(frames_b, 84, 84)
    return np.maximum(qa, qb)
