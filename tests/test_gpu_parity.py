"""GPU (-m gpu): the CUDA path, called through the C ABI (alegpu.py -> libalegpu.so), against
the oracle restatement (oracle/restate.c) on the same seeded inputs, against the committed
golden fixtures produced by the reference's own code, and through size-independent properties
at the benchmark sizes.  Tolerances are the ones SURVEY.md §8c states."""
import os

import numpy as np
import pytest

import alegpu
import cases
import synth
from oracle import restate

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
load = lambda n: np.load(os.path.join(G, n))
DA = alegpu.DeviceArray


def hp(clip=cases.CLIP, v=cases.VCOEF, e=cases.ECOEF, g=cases.MAX_NORM):
    return alegpu.Hyperparameters(clip, v, e, g)


def rel_l2(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


# ---- A5: GAE / Buffer::get ---------------------------------------------------------------------
@pytest.mark.parametrize("name", list(cases.GAE_KATS))
def test_gae_known_answers(name):
    c = cases.GAE_KATS[name]
    got = alegpu.gae_numpy(c["rewards"], c["values"], c["next_values"], c["terminals"],
                           c["truncations"], c["episode_starts"], cases.GAMMA, cases.LAMBDA)
    np.testing.assert_allclose(got, cases.gae_kat_expected(c), atol=1e-5, rtol=0)
    np.testing.assert_allclose(got, load("gae_kats.npz")[name], atol=1e-6, rtol=0)


@pytest.mark.parametrize("E,T", [(1, 1), (5, 37), (8, 128), (33, 65), (64, 200), (4096, 5)])
def test_buffer_get_matches_oracle_bitwise(E, T):
    s = synth.rollout_scalars(E, T, seed=E * 1000 + T)
    want = restate.buffer_get(s["rewards"], s["values"], s["next_values"], s["terminals"],
                              s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    got = alegpu.buffer_get_numpy(s["rewards"], s["values"], s["next_values"], s["terminals"],
                                  s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    for k in ("rewards", "masks", "advantages", "returns"):  # same op order, no fma: bit-exact
        assert np.array_equal(got[k], want[k]), k


def test_buffer_get_golden_and_errors():
    s, g = cases.gae_case(), load("buffer_get.npz")
    got = alegpu.buffer_get_numpy(s["rewards"], s["values"], s["next_values"], s["terminals"],
                                  s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    assert np.array_equal(got["rewards"], g["rewards"]) and np.array_equal(got["masks"], g["masks"])
    np.testing.assert_allclose(got["advantages"], g["advantages"], atol=1e-5, rtol=1e-5)
    np.testing.assert_allclose(got["returns"], g["returns"], atol=1e-5, rtol=1e-5)
    with pytest.raises(RuntimeError):  # buffer.cc:64-65
        alegpu.buffer_get_numpy(s["rewards"], s["values"], s["next_values"], s["terminals"],
                                s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA, 2)
    bad = s["terminals"].copy()
    bad[2, 5] = 1
    st = s["episode_starts"].copy()
    st[2, 5] = 1
    with pytest.raises(ValueError):  # gae.cc:49-53
        alegpu.gae_numpy(s["rewards"], s["values"], s["next_values"], bad, s["truncations"], st,
                         cases.GAMMA, cases.LAMBDA)


def test_gae_full_size_property():
    # sweep shape: with lambda = 0 and no flags the scan degenerates to the one-step TD error
    E, T = 4096, 128
    rng = np.random.default_rng(0)
    r, v = rng.standard_normal((E, T)).astype(np.float32), rng.standard_normal((E, T)).astype(np.float32)
    nv = rng.standard_normal(E).astype(np.float32)
    z = np.zeros((E, T), np.uint8)
    got = alegpu.gae_numpy(r, v, nv, z, z, z, 0.99, 0.0)
    nxt = np.concatenate([v[:, 1:], nv[:, None]], 1)
    want = (r + np.float32(0.99) * nxt) - v
    assert np.array_equal(got, want.astype(np.float32))


# ---- A1-A3: preprocess ----------------------------------------------------------------------------
def _preprocess_step(fa, fb, starts, stack):
    E = fa.shape[0]
    dst, gray = DA.zeros((E, 4, 84, 84), np.uint8), DA.zeros((E, 84, 84), np.float32)
    alegpu.preprocess_step(DA.from_numpy(fa), DA.from_numpy(fb), DA.from_numpy(starts, np.uint8),
                           DA.from_numpy(stack), dst, gray)
    alegpu.synchronize()
    return dst.numpy(), gray.numpy()


def test_preprocess_constant_images():
    # test/ai/vision-test.cc:5-30: constant in -> the same constant out, exactly
    for val in (0, 1, 2, 3, 255):
        f = np.full((1, 210, 160, 3), val, np.uint8)
        st, _ = _preprocess_step(f, f, np.ones(1, np.uint8), np.zeros((1, 4, 84, 84), np.uint8))
        assert np.all(st == val)


def test_preprocess_matches_oracle_and_golden():
    fr, g = cases.frames_case(), load("preprocess.npz")
    E = fr.shape[0]
    prev = synth.observations(E, seed=5)
    starts = np.array([1, 0, 0, 1, 0], np.uint8)
    got, gray = _preprocess_step(fr[:, 0], fr[:, 1], starts, prev)
    want_gray, want_q = restate.preprocess(fr[:, 0], fr[:, 1])
    assert np.array_equal(gray, want_gray)                  # same fp32 op order: bit-exact
    assert np.array_equal(got, restate.stack_update(prev, want_q, starts))  # max/stack: bit-exact
    np.testing.assert_allclose(gray, g["gray_f32"], atol=1e-3, rtol=0)   # vs libtorch area+matmul
    diff = got[:, 0].astype(int) - g["gray_u8"].astype(int)
    assert np.abs(diff).max() <= 1 and (diff != 0).mean() <= 1e-3


def test_vision_operators_match_oracle():
    """The stand-alone vision operators of src/ai/vision.h:6-13 (alegpu_vision_*): bit-exact against the
    restatement (same fp32 order), which is pinned to libtorch in test_oracle_golden.py; chained they
    reproduce the fused preprocess kernel's un-quantised grayscale to 1e-3."""
    rng = np.random.default_rng(8)
    x = (rng.integers(0, 256, (2, 2, 3, 210, 160))).astype(np.float32)
    small = alegpu.vision_resize_area(x)
    assert small.shape == (2, 2, 3, 84, 84)
    assert np.array_equal(small, restate.resize_area(x))
    gray = alegpu.vision_rgb_to_grayscale(small)
    assert np.array_equal(gray, restate.rgb_to_grayscale(small))
    for v in (0.0, 17.0, 255.0):  # test/ai/vision-test.cc:5-30
        assert np.all(alegpu.vision_resize_area(np.full((3, 210, 160), v, np.float32)) == v)
    hwc = np.ascontiguousarray(x[:, 0].transpose(0, 2, 3, 1)).astype(np.uint8)  # [2,210,160,3]
    fused_gray, _ = restate.preprocess(hwc, hwc)
    np.testing.assert_allclose(gray[:, 0], fused_gray, atol=1e-3, rtol=0)


def test_preprocess_rollout_equals_step_by_step():
    E, T = 3, 9
    s = synth.rollout_scalars(E, T, seed=4)
    starts = s["episode_starts"].copy()
    starts[1, 3] = 1; starts[1, 4] = 1; starts[2, T - 1] = 1  # back-to-back resets, reset at the end
    frames = synth.frames(T * E, seed=8, kind="ale").reshape(T, E, 2, 210, 160, 3)
    frames[2] = synth.frames(E, seed=9)
    carry = synth.observations(E, seed=6)
    obs, cout = DA.zeros((E, T, 4, 84, 84), np.uint8), DA.zeros((E, 4, 84, 84), np.uint8)
    alegpu.preprocess_rollout(DA.from_numpy(frames), DA.from_numpy(starts), DA.from_numpy(carry),
                              cout, obs, E, T)
    alegpu.synchronize()
    got, got_carry = obs.numpy(), cout.numpy()
    stack = carry.copy()
    for t in range(T):  # oracle: Buffer::add(obs) then update_observations (rollout.cc:229-237)
        assert np.array_equal(got[:, t], stack), t
        _, q = restate.preprocess(frames[t, :, 0], frames[t, :, 1])
        stack = restate.stack_update(stack, q, starts[:, t])
    assert np.array_equal(got_carry, stack)


@pytest.mark.parametrize("ring_slots", [None, 2, 4])
def test_rollout_ingest_equals_step_by_step(ring_slots, monkeypatch):
    """alegpu_rollout_ingest (the hand-off the e2e bench number runs through): pinned host frames ->
    copy stream -> device ring -> preprocess.  Both branches — the rollout-deep ring (one copy, one
    fused kernel) and the per-step ring with slot reuse (ALEGPU_RING_SLOTS < T) — over THREE
    consecutive rollouts with the carry stack alternating, bit-exact against T calls of
    alegpu_preprocess_step and against the oracle restatement."""
    import ctypes as C
    if ring_slots is not None:
        monkeypatch.setenv("ALEGPU_RING_SLOTS", str(ring_slots))
    E, T = 5, 7
    lib = alegpu.lib()
    buf = alegpu.Buffer(E, T, 4)
    slab = E * 2 * alegpu.FRAME_BYTES
    hptr = C.c_void_p()
    alegpu._check(lib.alegpu_host_alloc(C.byref(hptr), C.c_size_t(T * slab)))
    host = np.ctypeslib.as_array(C.cast(hptr, C.POINTER(C.c_uint8)), shape=(T * slab,))
    host = host.reshape(T, E, 2, 210, 160, 3)
    stream = C.c_void_p()
    alegpu._check(lib.alegpu_stream_create(C.byref(stream)))
    carry0 = synth.observations(E, seed=61)
    buf.carry.view(0, (E, 4, 84, 84)).copy_from(carry0)
    stack = carry0.copy()          # oracle-side carried stack
    step_src = DA.from_numpy(carry0)
    for it in range(3):
        s = synth.rollout_scalars(E, T, seed=70 + it)
        starts = s["episode_starts"].copy()
        starts[1, 2] = 1; starts[1, 3] = 1; starts[2, T - 1] = 1
        frames = synth.frames(T * E, seed=80 + it, kind="ale" if it % 2 else "uniform")
        host[...] = frames.reshape(T, E, 2, 210, 160, 3)
        buf.episode_starts.copy_from(starts)
        alegpu._check(lib.alegpu_rollout_ingest(buf.h, hptr, C.c_int(it % 2), stream))
        alegpu.synchronize(stream.value)
        got = buf.observations.numpy()
        got_carry = buf.carry.numpy()[1 - it % 2]
        # (a) the oracle restatement
        for t in range(T):
            assert np.array_equal(got[:, t], stack), (it, t)
            _, q = restate.preprocess(host[t, :, 0], host[t, :, 1])
            stack = restate.stack_update(stack, q, starts[:, t])
        assert np.array_equal(got_carry, stack), it
        # (b) T calls of alegpu_preprocess_step on device-resident frames
        dframes, dstarts = DA.from_numpy(host), DA.from_numpy(starts)
        for t in range(T):
            dst = DA.zeros((E, 4, 84, 84), np.uint8)
            fa = dframes.view(t * slab, (E, 210, 160, 3))
            alegpu._check(lib.alegpu_preprocess_step_strided(  # frames interleaved [E,2,...]
                C.c_void_p(fa.ptr), C.c_void_p(fa.ptr + alegpu.FRAME_BYTES),
                C.c_int64(2 * alegpu.FRAME_BYTES), C.c_void_p(dstarts.ptr + t), C.c_int64(T),
                C.c_void_p(step_src.ptr), C.c_int64(alegpu.STACK_BYTES), C.c_void_p(dst.ptr),
                C.c_int64(alegpu.STACK_BYTES), C.c_int64(E), None))
            alegpu.synchronize()
            step_src = dst
            want_next = got[:, t + 1] if t + 1 < T else got_carry
            assert np.array_equal(dst.numpy(), want_next), (it, t)
    alegpu._check(lib.alegpu_stream_destroy(stream))
    del buf
    alegpu._check(lib.alegpu_host_free(hptr))


def test_preprocess_full_size_properties():
    # v0 step count (1024 frame pairs): max is commutative, and a pair (f,f) is idempotent
    E, T = 8, 16
    frames = synth.frames(E * T, seed=3).reshape(T, E, 2, 210, 160, 3)
    swapped = np.ascontiguousarray(frames[:, :, ::-1])
    starts = synth.rollout_scalars(E, T, seed=1)["episode_starts"]
    outs = []
    for f in (frames, swapped):
        obs, cout = DA.zeros((E, T, 4, 84, 84), np.uint8), DA.zeros((E, 4, 84, 84), np.uint8)
        alegpu.preprocess_rollout(DA.from_numpy(f), DA.from_numpy(starts),
                                  DA.zeros((E, 4, 84, 84), np.uint8), cout, obs, E, T)
        alegpu.synchronize()
        outs.append((obs.numpy(), cout.numpy()))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    # plane p of slot s equals plane 0 of slot s-p when no reset lies between
    o = outs[0][0]
    for e in range(E):
        for s_ in range(4, T):
            if not starts[e, s_ - 3:s_].any():
                for p in range(1, 4):
                    assert np.array_equal(o[e, s_, p], o[e, s_ - p, 0])


# ---- A8: loss ---------------------------------------------------------------------------------------
def test_loss_matches_oracle_and_golden():
    lc, g = cases.loss_case(), load("loss.npz")
    got = alegpu.losses.compute(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"],
                                lc["values"], lc["returns"], lc["masks"], cases.CLIP, cases.VCOEF,
                                cases.ECOEF)
    want = restate.loss(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"],
                        lc["values"], lc["returns"], lc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    for ref_ in (want, g):
        for k in ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dlogits",
                  "dvalues"):
            np.testing.assert_allclose(got[k], ref_[k], rtol=1e-5, atol=1e-6, err_msg=k)
        assert abs(got["loss"] - float(ref_["loss"])) <= 1e-5 * abs(float(ref_["loss"])) + 1e-6
    np.testing.assert_allclose(alegpu.losses.normalize_logits(lc["logits"]), g["log_softmax"], atol=1e-6)


@pytest.mark.parametrize("n,A", [(1, 4), (257, 4), (16384, 4), (100, 18)])
def test_loss_sizes(n, A):
    lc = cases.loss_case(n, A, seed=n + A)
    got = alegpu.losses.compute(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"],
                                lc["values"], lc["returns"], lc["masks"] | (np.arange(n) == 0), 0.2,
                                0.5, 0.01)
    want = restate.loss(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"],
                        lc["values"], lc["returns"], lc["masks"] | (np.arange(n) == 0), 0.2, 0.5, 0.01)
    for k in ("total_losses", "ratio", "dlogits", "dvalues"):
        np.testing.assert_allclose(got[k], want[k], rtol=1e-5, atol=1e-6, err_msg=k)
    assert abs(got["loss"] - want["loss"]) <= 1e-5 * abs(want["loss"]) + 1e-6


# ---- A7/A9: network, gradients, clip + Adam (fp32-faithful mode) -----------------------------------
def _batch(tc, n=None):
    sl = slice(0, n)
    return alegpu.Batch(tc["obs"][sl], tc["actions"][sl], tc["log_probs"][sl], tc["advantages"][sl],
                        tc["returns"][sl], tc["masks"][sl])


def test_network_forward_backward_fp32_golden():
    H, g = 32, load("net_h32.npz")
    params, tc = cases.net_params(H), cases.train_case()
    L = alegpu.Learner(H, 4, 16, alegpu.MATH_FP32, cases.LR)
    L.set_params(params)
    assert np.array_equal(L.get_params(), params)  # layout permutation round-trips exactly
    logits, values = L.forward(tc["obs"][:6])
    np.testing.assert_allclose(logits, g["logits"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(values, g["values"], rtol=1e-4, atol=1e-5)
    c1, c2, c3, fc = L.activations(2)
    np.testing.assert_allclose(c1[:, :, ::5, ::5], g["conv1"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(c2, g["conv2"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(c3, g["conv3"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(fc, g["fc"], rtol=1e-4, atol=1e-5)
    grads, loss, _, _ = L.grads(_batch(tc), hp())
    assert abs(loss - float(g["loss"])) <= 1e-5 * abs(float(g["loss"])) + 1e-6
    offs, idx = cases.param_offsets(H), cases.sample_idx(params.size)
    norms = [np.linalg.norm(grads[offs[i]:offs[i + 1]].astype(np.float64)) for i in range(12)]
    np.testing.assert_allclose(norms, g["grad_norms"], rtol=1e-4)
    np.testing.assert_allclose(grads[idx], g["grad_samples"], rtol=1e-3, atol=1e-6)
    # per-tensor rel-L2 against the oracle's full gradient
    lg, v, acts = restate.net_forward(params, H, tc["obs"], keep_acts=True)
    lo = restate.loss(lg, tc["log_probs"], tc["actions"], tc["advantages"], v, tc["returns"],
                      tc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    want = restate.net_backward(params, H, acts, lo["dlogits"], lo["dvalues"])
    for i in range(12):
        assert rel_l2(grads[offs[i]:offs[i + 1]], want[offs[i]:offs[i + 1]]) <= 1e-4, i
    # clip + Adam: two steps from the golden's own gradient sequence
    tn1 = L.apply_grads(grads, cases.MAX_NORM)
    p1 = L.get_params()
    tn2 = L.apply_grads(grads * np.float32(0.01), cases.MAX_NORM)
    p2 = L.get_params()
    m2, v2, st = L.get_adam()
    np.testing.assert_allclose([tn1, tn2], g["total_norm"], rtol=1e-5)
    np.testing.assert_allclose(p1[idx], g["p1"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(p2[idx], g["p2"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(m2[idx], g["m2"], rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(v2[idx], g["v2"], rtol=1e-4, atol=1e-12)
    assert st == 2


@pytest.mark.parametrize("graph", [False, True])
def test_train_fp32_golden(graph):
    H, g = 32, load("train_h32.npz")
    params, tc = cases.net_params(H), cases.train_case()
    L = alegpu.Learner(H, 4, 8, alegpu.MATH_FP32, cases.LR)
    L.set_params(params)
    met = alegpu.Metrics(2, 2, 8)
    L.train(_batch(tc), met, 2, 2, hp(), cuda_graph=graph)
    assert L.last_launch_count() > 0
    idx = cases.sample_idx(params.size)
    p = L.get_params()
    np.testing.assert_allclose(p[idx], g["params"], rtol=1e-4, atol=1e-6)
    m = met.numpy()
    for k in ("loss", "clipped_losses", "value_losses", "entropies", "ratio", "total_losses",
              "clipped_gradients"):
        np.testing.assert_allclose(m[k], g[k], rtol=1e-3, atol=1e-5, err_msg=k)
    for k in ("advantages", "returns", "masks"):
        assert np.array_equal(m[k], g[k]), k
    assert L.get_adam()[2] == 4
    with pytest.raises(RuntimeError):  # train.h:140-143
        L.train(_batch(tc), None, 1, 5, hp())


def test_train_fp32_v0_minibatch_vs_oracle():
    # one v0-sized minibatch (mb=256, H=512): forward + gradient vs the oracle, then 2 graph replays
    H, n = 512, 256
    params = cases.net_params(H, seed=13)
    tc = cases.train_case(n, seed=31)
    L = alegpu.Learner(H, 4, n, alegpu.MATH_FP32, cases.LR)
    L.set_params(params)
    grads, loss, logits, values = L.grads(_batch(tc), hp())
    lg, v, acts = restate.net_forward(params, H, tc["obs"], keep_acts=True)
    lo = restate.loss(lg, tc["log_probs"], tc["actions"], tc["advantages"], v, tc["returns"],
                      tc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    want = restate.net_backward(params, H, acts, lo["dlogits"], lo["dvalues"])
    assert rel_l2(logits, lg) <= 1e-5 and rel_l2(values, v) <= 1e-5
    assert abs(loss - lo["loss"]) <= 1e-5 * abs(lo["loss"]) + 1e-6
    offs = cases.param_offsets(H)
    for i in range(12):
        # 3.1M ReLU pre-activations at this size: a few sit within fp32 summation noise of zero and
        # may flip (each flip moves the conv1/conv2 gradients by ~1e-3), so the bound is 5e-3 there
        tol = 5e-3 if i < 4 else 1e-4
        a, b = grads[offs[i]:offs[i + 1]], want[offs[i]:offs[i + 1]]
        assert rel_l2(a, b) <= tol, (i, rel_l2(a, b))
        assert np.dot(a.astype(np.float64), b.astype(np.float64)) / (
            np.linalg.norm(a.astype(np.float64)) * np.linalg.norm(b.astype(np.float64))) >= 0.9999
    # determinism: the same train() twice from the same weights gives identical bits
    outs = []
    for _ in range(2):
        L.set_params(params)
        L2 = alegpu.Learner(H, 4, n // 2, alegpu.MATH_FP32, cases.LR)
        L2.set_params(params)
        L2.train(_batch(tc), None, 1, 2, hp(), cuda_graph=True)
        outs.append(L2.get_params())
    assert np.array_equal(outs[0], outs[1])


# ---- A4/A6: rollout buffer ---------------------------------------------------------------------------
def test_buffer_add_get_roundtrip():
    E, T, A = 6, 11, 4
    s = synth.rollout_scalars(E, T, seed=17)
    obs = synth.observations(E * T, seed=18).reshape(E, T, 4, 84, 84)
    buf = alegpu.Buffer(E, T, A)
    for t in range(T):
        buf.add(obs[:, t], s["actions"][:, t], s["rewards"][:, t], s["terminals"][:, t],
                s["truncations"][:, t], s["episode_starts"][:, t], s["logits"][:, t],
                s["values"][:, t])
    batch = buf.get(s["next_values"], cases.GAMMA, cases.LAMBDA)
    want = restate.buffer_get(s["rewards"], s["values"], s["next_values"], s["terminals"],
                              s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    assert np.array_equal(batch.observations.numpy().reshape(E, T, 4, 84, 84), obs)
    assert np.array_equal(batch.actions.numpy().reshape(E, T), s["actions"])
    assert np.array_equal(batch.advantages.numpy().reshape(E, T), want["advantages"])
    assert np.array_equal(batch.returns.numpy().reshape(E, T), want["returns"])
    assert np.array_equal(batch.masks.numpy().reshape(E, T), want["masks"])
    assert np.array_equal(buf.rewards.numpy(), want["rewards"])
    np.testing.assert_allclose(batch.log_probabilities.numpy(),
                               restate.log_softmax(s["logits"].reshape(-1, A)), atol=1e-6)
    buf.add(obs[:, 0], s["actions"][:, 0], s["rewards"][:, 0], s["terminals"][:, 0],
            s["truncations"][:, 0], s["episode_starts"][:, 0], s["logits"][:, 0], s["values"][:, 0])
    with pytest.raises(RuntimeError):  # not full: buffer.cc:64-65
        buf.get(s["next_values"], cases.GAMMA, cases.LAMBDA)


# ---- N1: rollout-side policy call (action selector, src/bin/train.cc:367-379) -------------------
def test_act_matches_oracle_given_uniforms():
    H, E = 32, 24
    params = cases.net_params(H)
    obs = cases.train_case(E, seed=77)["obs"]
    L = alegpu.Learner(H, 4, 16, alegpu.MATH_FP32, cases.LR)  # E > max_minibatch: chunked
    L.set_params(params)
    u = np.random.default_rng(3).random(E).astype(np.float32)
    actions, logits, values = L.act(obs, uniforms=u)
    want_logits, want_values = restate.net_forward(params, H, obs)[:2]
    np.testing.assert_allclose(logits, want_logits, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(values, want_values, rtol=1e-4, atol=1e-5)
    want_actions, cum = restate.act_sample(logits, u)
    # identical except where u sits within float noise of a CDF step
    near = (np.abs(cum - u[:, None]) < 1e-6).any(axis=1)
    assert np.array_equal(actions[~near], want_actions[~near])
    assert actions.min() >= 0 and actions.max() < 4


def test_act_writes_rollout_column_and_is_categorical():
    H, E, T, t = 32, 512, 3, 1
    params = cases.net_params(H)
    obs = np.repeat(cases.train_case(1, seed=5)["obs"], E, axis=0)  # same state in every env
    L = alegpu.Learner(H, 4, 256, alegpu.MATH_FP32, cases.LR)
    L.set_params(params)
    buf = alegpu.Buffer(E, T, 4)
    counts = np.zeros(4)
    for step in range(8):
        actions, logits, values = L.act(obs, seed=11, step=step, buffer=buf, t=t)
        assert np.array_equal(buf.actions.numpy()[:, t], actions)  # column t, nothing else
        assert np.all(buf.actions.numpy()[:, [0, 2]] == 0)
        np.testing.assert_array_equal(buf.logits.numpy()[:, t], logits)
        np.testing.assert_array_equal(buf.values.numpy()[:, t], values)
        counts += np.bincount(actions, minlength=4)
    p = np.exp(restate.log_softmax(logits[:1]))[0].astype(np.float64)
    n = counts.sum()
    chi2 = float(((counts - n * p) ** 2 / (n * p)).sum())
    assert chi2 < 21.1, (chi2, counts, p)  # 3 dof, p ~ 1e-4
    # a different step gives a different draw, the same (seed, step) the same draw
    a0 = L.act(obs, seed=11, step=0)[0]
    assert np.array_equal(a0, L.act(obs, seed=11, step=0)[0])
    assert not np.array_equal(a0, L.act(obs, seed=11, step=1)[0])


# ---- north_star extras: minibatch shuffle/gather and advantage normalisation (opt-in) ------------
@pytest.mark.parametrize("shuffle,norm", [(True, False), (False, True), (True, True)])
def test_train_shuffle_and_advantage_normalisation(shuffle, norm):
    """train_ex(indices, normalize) on the original rows == the oracle's contiguous train() on rows
    that were gathered / normalised on the host first (one epoch, so one permutation)."""
    H, N, nmb = 32, 64, 4
    mb = N // nmb
    params = cases.net_params(H)
    tc = cases.train_case(N, seed=41)
    perm = np.random.default_rng(9).permutation(N).astype(np.int64) if shuffle else np.arange(N)
    obs, act, lp, adv, ret, msk = restate.gather_minibatches(
        perm, tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"], tc["masks"])
    adv_used = adv.copy()
    if norm:
        for k in range(nmb):
            sl = slice(k * mb, (k + 1) * mb)
            adv_used[sl] = restate.normalize_advantages(adv[sl], msk[sl])
    z = np.zeros_like(params)
    want_p, _, _, _, want_m = restate.train(params, z, z, 0, H, obs, act, lp, adv_used, ret, msk, 1,
                                            nmb, cases.CLIP, cases.VCOEF, cases.ECOEF,
                                            cases.MAX_NORM, cases.LR)
    met = alegpu.Metrics(1, nmb, mb)
    batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                         tc["masks"])
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    for graph in (False, True):
        L = alegpu.Learner(H, 4, mb, alegpu.MATH_FP32, cases.LR)  # fresh Adam state per mode
        L.set_params(params)
        L.train(batch, met, 1, nmb, hp, cuda_graph=graph, indices=perm if shuffle else None,
                normalize_advantages=norm)
        got = L.get_params()
        err = np.linalg.norm((got - want_p).astype(np.float64)) / np.linalg.norm(want_p)
        assert err < 1e-4, (err, graph)  # fp32-faithful mode, same bound as the plain train() test
        m = met.numpy()
        np.testing.assert_allclose(m["loss"], want_m["loss"], rtol=1e-3, atol=1e-5)
        np.testing.assert_array_equal(m["advantages"].ravel(), adv)  # raw, in gathered order
        np.testing.assert_array_equal(m["masks"].ravel().astype(bool), msk.astype(bool))


@pytest.mark.parametrize("shuffle", [False, True])
def test_train_clipped_value_loss(shuffle):
    """alegpu_train_options.clip_value_loss (needs batch.values) == the oracle's train() with the
    clipped value loss restated (pinned against torch autograd in test_oracle_golden.py)."""
    H, N, nmb = 32, 64, 4
    mb = N // nmb
    params = cases.net_params(H)
    tc = cases.train_case(N, seed=43)
    old_v = (tc["returns"] + 0.3 * np.random.default_rng(2).standard_normal(N)).astype(np.float32)
    perm = np.random.default_rng(5).permutation(N).astype(np.int64) if shuffle else np.arange(N)
    obs, act, lp, adv, ret, msk, ov = restate.gather_minibatches(
        perm, tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"], tc["masks"],
        old_v)
    z = np.zeros_like(params)
    want_p, _, _, _, want_m = restate.train(params, z, z, 0, H, obs, act, lp, adv, ret, msk, 1, nmb,
                                            cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM,
                                            cases.LR, old_values=ov)
    plain_p = restate.train(params, z, z, 0, H, obs, act, lp, adv, ret, msk, 1, nmb, cases.CLIP,
                            cases.VCOEF, cases.ECOEF, cases.MAX_NORM, cases.LR)[0]
    assert not np.array_equal(want_p, plain_p)  # the option changes the update
    met = alegpu.Metrics(1, nmb, mb)
    batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                         tc["masks"], values=old_v)
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    for graph in (False, True):
        L = alegpu.Learner(H, 4, mb, alegpu.MATH_FP32, cases.LR)
        L.set_params(params)
        L.train(batch, met, 1, nmb, hp, cuda_graph=graph, indices=perm if shuffle else None,
                clip_value_loss=True)
        got = L.get_params()
        err = np.linalg.norm((got - want_p).astype(np.float64)) / np.linalg.norm(want_p)
        assert err < 1e-4, (err, graph)
        m = met.numpy()
        # minibatches 1.. see weights that already took Adam steps (sign-like for near-zero gradients):
        # exact for minibatch 0, 5e-3 after
        np.testing.assert_allclose(m["value_losses"][0, 0], want_m["value_losses"][0, 0], rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(m["value_losses"], want_m["value_losses"], rtol=5e-3, atol=1e-5)
        np.testing.assert_allclose(m["loss"], want_m["loss"], rtol=1e-3, atol=1e-5)
    with pytest.raises(ValueError):  # the option without the stored values
        L.train(alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"],
                             tc["returns"], tc["masks"]), met, 1, nmb, hp, clip_value_loss=True)


def test_train_action_out_of_range_is_reported():
    """An action index outside [0, A) (the reference's gather at losses.cc:12 would throw) is flagged
    on the device through a mapped host flag and reported by the next learner call."""
    H, N = 32, 16
    tc = cases.train_case(N, seed=3)
    L = alegpu.Learner(H, 4, N, alegpu.MATH_FP32, cases.LR)
    L.set_params(cases.net_params(H))
    acts = tc["actions"].copy()
    acts[5] = 4
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    bad = alegpu.Batch(tc["obs"], acts, tc["log_probs"], tc["advantages"], tc["returns"], tc["masks"])
    good = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                        tc["masks"])
    L.train(bad, None, 1, 1, hp)
    with pytest.raises(IndexError):
        L.train(good, None, 1, 1, hp)
    L.train(good, None, 1, 1, hp)  # the flag was cleared


def test_train_shuffle_index_out_of_range():
    H, N = 32, 16
    tc = cases.train_case(N, seed=2)
    L = alegpu.Learner(H, 4, N, alegpu.MATH_FP32, cases.LR)
    L.set_params(cases.net_params(H))
    batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                         tc["masks"])
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    bad = np.arange(N, dtype=np.int64)
    bad[3] = N  # torch index_select would throw (std::out_of_range family)
    before = L.get_params()
    with pytest.raises(IndexError):  # std::out_of_range, from the SAME call, before anything trains
        L.train(batch, None, 1, 1, hp, indices=bad)
    assert np.array_equal(L.get_params(), before) and L.get_adam()[2] == 0
    L.train(batch, None, 1, 1, hp, indices=np.arange(N, dtype=np.int64))  # flag was cleared
    assert L.get_adam()[2] == 1


# ---- N3: on-device reduction of Metrics for the logging feed (src/bin/train.cc:163-210) ---------
@pytest.mark.parametrize("shape", [(2, 4, 64), (1, 3, 5000)])
def test_metrics_summary_and_masked_select(shape):
    rng = np.random.default_rng(sum(shape))
    met = alegpu.Metrics(*shape)
    host = {}
    for k in alegpu.Metrics.FIELDS:
        host[k] = rng.standard_normal(shape).astype(np.float32)
        getattr(met, k).copy_from(host[k])
    host["masks"] = (rng.random(shape) > 0.2).astype(np.uint8)
    met.masks.copy_from(host["masks"])
    host["clipped_gradients"] = rng.random(shape[:2]).astype(np.float32)
    met.clipped_gradients.copy_from(host["clipped_gradients"])
    m = host["masks"].astype(bool)
    s = met.summary()
    # the reference's expressions, in float64 (torch's fp32 mean differs by summation order only)
    want = {"mean_clipped_gradient": host["clipped_gradients"].mean(dtype=np.float64),
            "mean_loss": host["loss"].mean(dtype=np.float64),
            "mean_clipped_loss": host["clipped_losses"][m].mean(dtype=np.float64),
            "mean_value_loss": host["value_losses"][m].mean(dtype=np.float64),
            "mean_entropy": host["entropies"][m].mean(dtype=np.float64),
            "mean_ratio": host["ratio"][m].mean(dtype=np.float64),
            "mean_total_loss": host["total_losses"][m].mean(dtype=np.float64),
            "unmasked": float(m.sum())}
    for k, w in want.items():
        assert abs(s[k] - w) <= 1e-6 + 1e-5 * abs(w), (k, s[k], w)
    for k in ("total_losses", "advantages"):  # bit-exact, order preserved
        assert np.array_equal(met.masked(k), host[k][m])



# ---- rollout collection loop (Rollout::rollout, src/ai/rollout.cc:198-278) -----------------------
def _collect(E, T, H, workers, seed):
    params = cases.net_params(H)
    L = alegpu.Learner(H, 4, max(E, 8), alegpu.MATH_FP32, cases.LR)
    L.set_params(params)
    buf = alegpu.Buffer(E, T, 4)
    col = alegpu.Collector(E, T, 4, num_workers=workers, seed=seed)
    out = []
    for _ in range(2):  # two rollouts: the second starts from the carried stacks / episode state
        batch, log = col.rollout(buf, L, cases.GAMMA, cases.LAMBDA)
        out.append({"frames": col.frames().copy(), "log": log,
                    **{k: getattr(buf, k).numpy().copy() for k in
                       ("observations", "actions", "rewards", "values", "logits", "terminals",
                        "truncations", "episode_starts", "advantages", "returns", "masks")}})
    return params, L, out


def test_collector_rollout_replays_through_the_oracle():
    E, T, H = 6, 40, 32
    params, L, outs = _collect(E, T, H, workers=3, seed=5)
    stack = np.zeros((E, 4, 84, 84), np.uint8)
    prev_end = None
    for r, o in enumerate(outs):
        st, te, tr = o["episode_starts"], o["terminals"], o["truncations"]
        # flag invariants of rollout.cc:239-265 / gae.cc:49-53
        assert not np.any((st & te) | (st & tr) | (te & tr))
        assert np.array_equal(st[:, 1:] != 0, (te[:, :-1] | tr[:, :-1]) != 0)
        if r == 0:
            assert np.all(st[:, 0] == 1)
        else:
            assert np.array_equal(st[:, 0] != 0, prev_end)
        prev_end = (te[:, -1] | tr[:, -1]) != 0
        # observations: bit-exact replay of max -> area resize -> gray -> stack on the recorded frames
        for t in range(T):
            assert np.array_equal(o["observations"][:, t], stack), (r, t)
            gray = restate.preprocess(o["frames"][t, :, 0], o["frames"][t, :, 1])[1]
            stack = restate.stack_update(stack, gray, st[:, t])
        # the policy outputs stored in column t are the network's outputs on slot t
        lg, vl = restate.net_forward(params, H, o["observations"].reshape(-1, 4, 84, 84))
        np.testing.assert_allclose(o["logits"].reshape(-1, 4), lg, rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(o["values"].ravel(), vl, rtol=1e-4, atol=1e-5)
        assert o["actions"].min() >= 0 and o["actions"].max() < 4
        # Buffer::get on the collected scalars (rewards were clamped in place, buffer.cc:67)
        nv = restate.net_forward(params, H, stack)[1]
        assert np.all(np.abs(o["rewards"]) <= 1.0)
        want = restate.buffer_get(o["rewards"], o["values"], nv, te, tr, st, cases.GAMMA, cases.LAMBDA)
        np.testing.assert_allclose(o["advantages"], want["advantages"], rtol=1e-4, atol=1e-5)
        np.testing.assert_array_equal(o["masks"] != 0, want["masks"] != 0)
    assert outs[1]["log"]["steps"] > outs[0]["log"]["steps"] > 0


def test_collector_is_deterministic_across_worker_counts():
    a = _collect(5, 12, 32, workers=1, seed=9)[2]
    b = _collect(5, 12, 32, workers=4, seed=9)[2]
    for x, y in zip(a, b):
        assert np.array_equal(x["frames"], y["frames"])
        for k in ("actions", "rewards", "terminals", "truncations", "episode_starts", "observations"):
            assert np.array_equal(x[k], y[k]), k


# ---- N4: checkpoint / resume ------------------------------------------------------------------------
@pytest.mark.parametrize("math", [alegpu.MATH_FP32, alegpu.MATH_BF16])
def test_checkpoint_resume_is_bit_identical(tmp_path, math):
    H, N, nmb = 32, 64, 2
    tc = cases.train_case(N, seed=8)
    batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                         tc["masks"])
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    a = alegpu.Learner(H, 4, N // nmb, math, cases.LR)
    a.set_params(cases.net_params(H))
    a.train(batch, None, 2, nmb, hp)
    path = str(tmp_path / "ckpt.npz")
    a.save(path)
    a.train(batch, None, 1, nmb, hp)
    b = alegpu.Learner(H, 4, N // nmb, math, cases.LR)
    b.load(path)
    assert b.get_adam()[2] == 2 * nmb
    b.train(batch, None, 1, nmb, hp)
    assert np.array_equal(a.get_params(), b.get_params())
    ma, va, sa = a.get_adam()
    mb_, vb, sb = b.get_adam()
    assert sa == sb == 3 * nmb and np.array_equal(ma, mb_) and np.array_equal(va, vb)
