"""GPU (-m gpu): the reference's LIVE observation path — per-frame grayscale resize
(ai::vision::resize_grayscale_image, src/ai/vision.cc:86-95 = stbir_resize_uint8_linear), then the
2-frame max (max_and_skip.cc:33-42), then the frame stack (rollout.cc:184-196) — through the C ABI.

PARITY UNPINNED: stb_image_resize2 is absent offline and the reference has no test of this overload,
so the kernel is held to oracle/restate.py's independent statement of the published algorithm and to
properties every such filter has; max / stack / index work is bit-exact."""
import numpy as np
import pytest

import alegpu
import synth
from oracle import restate

pytestmark = pytest.mark.gpu
DA = alegpu.DeviceArray
F32_TOL = 2e-3  # |pre-rounding value - float64 oracle|: <= 18 fp32 fma of magnitudes <= 255


def _check_against_oracle(img, nw, nh):
    n, h, w = img.shape
    got, got_f = alegpu.vision.resize_grayscale_image(img, w, h, nw, nh, return_f32=True)
    want, want_f = restate.resize_grayscale_stb(img, nw, nh)
    assert got.shape == (n, nh, nw) and got.dtype == np.uint8
    np.testing.assert_allclose(got_f, want_f, atol=F32_TOL, rtol=0)
    # u8 equal except where the oracle's value sits within the fp32 tolerance of a rounding boundary
    diff = got.astype(int) != want.astype(int)
    frac = np.abs((want_f.astype(np.float64) + 0.5) - np.round(want_f.astype(np.float64) + 0.5))
    assert np.all(frac[diff] <= 2 * F32_TOL), "u8 differs away from a rounding boundary"
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    return got


def test_resize_atari_screens_match_restated_stb():
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (5, 210, 160), dtype=np.uint8)          # worst case: white noise
    _check_against_oracle(img, 84, 84)
    ale = synth.frames(3, seed=2, kind="ale")[:, 0, :, :, 1].copy()    # ALE-shaped rectangles
    _check_against_oracle(ale, 84, 84)


@pytest.mark.parametrize("w,h,nw,nh", [(37, 23, 11, 7), (20, 10, 33, 17), (84, 84, 84, 84),
                                       (160, 210, 84, 110), (1, 1, 3, 2), (64, 48, 64, 7),
                                       (4, 1, 2, 1)])
def test_resize_ragged_shapes(w, h, nw, nh):
    rng = np.random.default_rng(w * 1000 + h)
    img = rng.integers(0, 256, (3, h, w), dtype=np.uint8)
    got = _check_against_oracle(img, nw, nh)
    if (w, h) == (nw, nh):
        assert np.array_equal(got, img)                                 # same size: identity


def test_resize_hand_derived_and_constants():
    # Mitchell at scale 1/2: o=0 collects M(.25)+M(.75)+M(1.25)+M(1.75) (x2 by symmetry = 2, so the
    # weights normalise by 1/2); with [0,0,255,255] -> 255*(M(.75)+M(1.25)+M(1.75))/2 = 27.78 -> 28
    got = alegpu.vision.resize_grayscale_image(np.array([0, 0, 255, 255], np.uint8), 4, 1, 2, 1)
    assert got.tolist() == [28, 227]
    for val in (0, 1, 127, 254, 255):                                    # as test/ai/vision-test.cc:5-30
        img = np.full((2, 210, 160), val, np.uint8)
        assert np.all(alegpu.vision.resize_grayscale_image(img, 160, 210, 84, 84) == val)
    # negative lobes overshoot an edge; the encode clamps to [0, 255]
    edge = np.zeros((1, 210, 160), np.uint8)
    edge[:, :, 80:] = 255
    out, f = alegpu.vision.resize_grayscale_image(edge, 160, 210, 84, 84, return_f32=True)
    assert f.min() < 0.0 and f.max() > 255.0 and out.min() == 0 and out.max() == 255


def test_resize_errors_and_empty():
    img = np.zeros((1, 8, 8), np.uint8)
    for nw, nh in ((0, 4), (4, 0), (-1, 4)):                             # resize.cc:15-16
        with pytest.raises(ValueError):
            alegpu.vision.resize_grayscale_image(img, 8, 8, nw, nh)
    with pytest.raises(ValueError):                                     # vision.cc:89 (size mismatch)
        alegpu.vision.resize_grayscale_image(np.zeros(63, np.uint8), 8, 8, 4, 4)
    out = alegpu.vision.resize_grayscale_image(np.zeros((0, 8, 8), np.uint8), 8, 8, 4, 4)
    assert out.shape == (0, 4, 4)


def _gray_step(fa, fb, starts, prev, **kw):
    E = fa.shape[0]
    dst = DA.zeros((E, 4, 84, 84), np.uint8)
    alegpu.preprocess_gray_step(DA.from_numpy(fa), DA.from_numpy(fb), DA.from_numpy(starts),
                                DA.from_numpy(prev), dst, **kw)
    alegpu.synchronize()
    return dst.numpy()


def test_live_step_is_resize_max_stack():
    E = 7
    rng = np.random.default_rng(3)
    fa = rng.integers(0, 256, (E, 210, 160), dtype=np.uint8)
    fb = rng.integers(0, 256, (E, 210, 160), dtype=np.uint8)
    fb[2] = fa[2]                                                        # window ended after one frame
    starts = np.array([1, 0, 0, 1, 0, 0, 0], np.uint8)
    prev = synth.observations(E, seed=5)
    got = _gray_step(fa, fb, starts, prev)
    # bit-exact against the stand-alone kernel + the oracle's max / stack update
    ra = alegpu.vision.resize_grayscale_image(fa, 160, 210, 84, 84)
    rb = alegpu.vision.resize_grayscale_image(fb, 160, 210, 84, 84)
    assert np.array_equal(got, restate.stack_update(prev, np.maximum(ra, rb), starts))
    assert np.array_equal(got[2, 0], ra[2])
    # and against the fully independent oracle chain, up to rounding-boundary pixels
    want = restate.stack_update(prev, restate.preprocess_gray(fa, fb), starts)
    d = got.astype(int) - want.astype(int)
    assert np.abs(d).max() <= 1 and (d != 0).mean() <= 2e-3
    assert np.array_equal(got[1, 1:], prev[1, :3]) and np.array_equal(got[0, 1], got[0, 0])
    # commutative in (a, b)
    assert np.array_equal(got, _gray_step(fb, fa, starts, prev))


def test_live_step_into_rollout_storage():
    # dst = slot t+1 of [E,T,4,84,84] storage, src = slot t, starts read with stride T
    E, T, t = 4, 6, 2
    rng = np.random.default_rng(8)
    fr = rng.integers(0, 256, (E, 2, 210, 160), dtype=np.uint8)         # interleaved [E,2,...] slab
    starts = np.zeros((E, T), np.uint8)
    starts[3, t] = 1
    obs = synth.observations(E * T, seed=1).reshape(E, T, 4, 84, 84)
    d_obs, d_fr, d_st = DA.from_numpy(obs), DA.from_numpy(fr), DA.from_numpy(starts)
    slot = 4 * 84 * 84
    alegpu.preprocess_gray_step(d_fr.ptr, d_fr.ptr + 210 * 160, d_st.ptr + t,
                                d_obs.ptr + t * slot, d_obs.ptr + (t + 1) * slot,
                                frame_env_stride=2 * 210 * 160, starts_stride=T,
                                src_env_stride=T * slot, dst_env_stride=T * slot, E=E)
    alegpu.synchronize()
    got = d_obs.numpy()
    q = np.maximum(alegpu.vision.resize_grayscale_image(fr[:, 0].copy(), 160, 210, 84, 84),
                   alegpu.vision.resize_grayscale_image(fr[:, 1].copy(), 160, 210, 84, 84))
    want = obs.copy()
    want[:, t + 1] = restate.stack_update(obs[:, t], q, starts[:, t])
    assert np.array_equal(got, want)                                      # other slots untouched


def test_live_step_full_size_properties():
    # v1's env count: idempotence under (f, f) and monotonicity of the max in either argument
    E = 4096
    rng = np.random.default_rng(1)
    base = rng.integers(0, 256, (64, 210, 160), dtype=np.uint8)
    fa = base[rng.integers(0, 64, E)]
    fb = base[rng.integers(0, 64, E)]
    starts = np.ones(E, np.uint8)
    prev = np.zeros((E, 4, 84, 84), np.uint8)
    ab, aa, bb = _gray_step(fa, fb, starts, prev), _gray_step(fa, fa, starts, prev), \
        _gray_step(fb, fb, starts, prev)
    assert np.array_equal(ab, np.maximum(aa, bb))
    assert np.array_equal(ab[:, 0], ab[:, 3])
