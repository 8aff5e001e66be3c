"""CPU: pins the oracle restatement (oracle/restate.c) against
  (1) the reference's own known-answer tests (test/ai/gae-test.cc x6, test/ai/vision-test.cc x2),
  (2) fixtures produced by the reference's own code (tools/make_golden.py -> tests/golden/).
If oracle/_ref (the reference compiled unmodified) is present it is cross-checked too."""
import os

import numpy as np
import pytest

import cases
from oracle import ref, restate

G = os.path.join(os.path.dirname(__file__), "golden")
load = lambda n: np.load(os.path.join(G, n))


@pytest.mark.parametrize("name", list(cases.GAE_KATS))
def test_gae_known_answers(name):
    c = cases.GAE_KATS[name]
    got = restate.gae(c["rewards"], c["values"], c["next_values"], c["terminals"], c["truncations"],
                      c["episode_starts"], cases.GAMMA, cases.LAMBDA)
    np.testing.assert_allclose(got, cases.gae_kat_expected(c), atol=1e-5, rtol=0)  # EXPECT_NEAR 1e-5
    np.testing.assert_allclose(got, load("gae_kats.npz")[name], atol=1e-6, rtol=0)


def test_gae_simple_values():
    # the reference's gae() reproduced (2.8109, 1.9308, 0.9950) on the first KAT
    c = cases.GAE_KATS["SimpleNoTerminalsNoTruncations"]
    got = restate.gae(c["rewards"], c["values"], c["next_values"], c["terminals"], c["truncations"],
                      c["episode_starts"], cases.GAMMA, cases.LAMBDA)
    np.testing.assert_allclose(got[0], [2.8109, 1.9308, 0.9950], atol=1e-4)


def test_gae_flag_exclusivity_throws():
    c = dict(cases.GAE_KATS["TerminalHandling"])
    c["episode_starts"] = c["terminals"].copy()
    with pytest.raises(ValueError):  # std::invalid_argument, gae.cc:49-53
        restate.gae(c["rewards"], c["values"], c["next_values"], c["terminals"], c["truncations"],
                    c["episode_starts"], cases.GAMMA, cases.LAMBDA)


def test_buffer_get_golden():
    s, g = cases.gae_case(), load("buffer_get.npz")
    got = restate.buffer_get(s["rewards"], s["values"], s["next_values"], s["terminals"],
                             s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    assert np.array_equal(got["rewards"], g["rewards"])  # clamp: exact
    assert np.array_equal(got["masks"], g["masks"])
    np.testing.assert_allclose(got["advantages"], g["advantages"], atol=1e-5, rtol=1e-5)
    np.testing.assert_allclose(got["returns"], g["returns"], atol=1e-5, rtol=1e-5)
    with pytest.raises(RuntimeError):  # buffer.cc:64-65
        restate.buffer_get(s["rewards"], s["values"], s["next_values"], s["terminals"],
                           s["truncations"], s["episode_starts"], cases.GAMMA, cases.LAMBDA, 3)


def test_vision_constant_images():
    # vision-test.cc:5-30: constant in -> the same constant out, exactly
    for val in (0, 1, 2, 3, 255):
        f = np.full((1, 210, 160, 3), val, np.uint8)
        g, q = restate.preprocess(f, f)
        assert np.all(q == val)
        np.testing.assert_allclose(g, val, atol=2e-4 * max(val, 1))  # weights sum to 1 +- ulp


def test_preprocess_golden():
    fr, g = cases.frames_case(), load("preprocess.npz")
    gray, q = restate.preprocess(fr[:, 0], fr[:, 1])
    np.testing.assert_allclose(gray, g["gray_f32"], atol=1e-3, rtol=0)
    diff = q.astype(int) - g["gray_u8"].astype(int)
    assert np.abs(diff).max() <= 1 and (diff != 0).mean() <= 1e-3  # +-1 LSB at .5 boundaries only
    stack = np.zeros((1, 4, 84, 84), np.uint8)
    for i, st in enumerate([1, 0, 0, 1, 0]):
        stack = restate.stack_update(stack, g["gray_u8"][i:i + 1], np.array([st], np.uint8))
        assert np.array_equal(stack[0], g["stacks"][i])  # frame-stack: bit-exact


def test_loss_golden():
    lc, g = cases.loss_case(), load("loss.npz")
    got = restate.loss(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"],
                       lc["values"], lc["returns"], lc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    for k in ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dlogits",
              "dvalues"):
        np.testing.assert_allclose(got[k], g[k], rtol=1e-5, atol=1e-6, err_msg=k)
    assert abs(got["loss"] - float(g["loss"])) <= 1e-5 * abs(float(g["loss"])) + 1e-6
    np.testing.assert_allclose(restate.log_softmax(lc["logits"]), g["log_softmax"], atol=1e-6)


def _restate_grads(params, tc, H=32):
    logits, values, acts = restate.net_forward(params, H, tc["obs"], keep_acts=True)
    lo = restate.loss(logits, tc["log_probs"], tc["actions"], tc["advantages"], values,
                      tc["returns"], tc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    return restate.net_backward(params, H, acts, lo["dlogits"], lo["dvalues"]), lo["loss"]


def test_network_golden():
    H, g = 32, load("net_h32.npz")
    params, tc = cases.net_params(H), cases.train_case()
    logits, values = restate.net_forward(params, H, tc["obs"][:6])
    np.testing.assert_allclose(logits, g["logits"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(values, g["values"], rtol=1e-4, atol=1e-5)
    grads, loss = _restate_grads(params, tc)
    assert abs(loss - float(g["loss"])) <= 1e-5 * abs(float(g["loss"])) + 1e-6
    offs, idx = cases.param_offsets(H), cases.sample_idx(params.size)
    norms = [np.linalg.norm(grads[offs[i]:offs[i + 1]].astype(np.float64)) for i in range(12)]
    np.testing.assert_allclose(norms, g["grad_norms"], rtol=1e-4)
    np.testing.assert_allclose(grads[idx], g["grad_samples"], rtol=1e-3, atol=1e-6)
    # clip + Adam, two steps (clipped, then un-clipped)
    m, v = np.zeros_like(params), np.zeros_like(params)
    p1, _, m, v, tn1 = restate.clip_adam(params, grads, m, v, H, 1, cases.LR, cases.MAX_NORM)
    p2, _, m, v, tn2 = restate.clip_adam(p1, grads * np.float32(0.01), m, v, H, 2, cases.LR, cases.MAX_NORM)
    np.testing.assert_allclose([tn1, tn2], g["total_norm"], rtol=1e-5)
    np.testing.assert_allclose(p1[idx], g["p1"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(p2[idx], g["p2"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(m[idx], g["m2"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(v[idx], g["v2"], rtol=1e-5, atol=1e-12)
    assert int(g["step"]) == 2


def test_train_golden():
    H, g = 32, load("train_h32.npz")
    params, tc = cases.net_params(H), cases.train_case()
    z = np.zeros_like(params)
    p, m, v, step, met = restate.train(params, z, z, 0, H, tc["obs"], tc["actions"], tc["log_probs"],
                                       tc["advantages"], tc["returns"], tc["masks"], 2, 2,
                                       cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM, cases.LR)
    assert step == 4
    idx = cases.sample_idx(params.size)
    np.testing.assert_allclose(p[idx], g["params"], rtol=1e-4, atol=1e-6)
    for k in ("loss", "clipped_losses", "value_losses", "entropies", "ratio", "total_losses",
              "clipped_gradients"):
        np.testing.assert_allclose(met[k], g[k], rtol=1e-3, atol=1e-5, err_msg=k)
    for k in ("advantages", "returns", "masks"):
        assert np.array_equal(met[k], g[k]), k
    with pytest.raises(RuntimeError):  # train.h:140-143
        restate.train(params, z, z, 0, H, tc["obs"], tc["actions"], tc["log_probs"],
                      tc["advantages"], tc["returns"], tc["masks"], 1, 5, cases.CLIP, cases.VCOEF,
                      cases.ECOEF, cases.MAX_NORM, cases.LR)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_restatement_matches_compiled_reference():
    s = cases.gae_case(8, 128, seed=21)
    a = ref.gae(s["rewards"], s["values"], s["next_values"], s["terminals"], s["truncations"],
                s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    b = restate.gae(s["rewards"], s["values"], s["next_values"], s["terminals"], s["truncations"],
                    s["episode_starts"], cases.GAMMA, cases.LAMBDA)
    np.testing.assert_allclose(a, b, atol=1e-6, rtol=1e-6)
    lc = cases.loss_case(257, seed=77)
    ra = ref.loss(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"], lc["values"],
                  lc["returns"], lc["masks"], 0.2, 0.5, 0.01)
    rb = restate.loss(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"],
                      lc["values"], lc["returns"], lc["masks"], 0.2, 0.5, 0.01)
    for k in ("total_losses", "ratio", "dlogits", "dvalues"):
        np.testing.assert_allclose(ra[k], rb[k], rtol=1e-5, atol=1e-6, err_msg=k)


# ---- restatements of the steps that have no reference test: pinned against libtorch here ----------
def test_act_sample_is_the_inverse_cdf_of_torch_softmax():
    """src/bin/train.cc:373-375: multinomial(softmax(logits), 1).  torch's Philox draw cannot be
    replayed, so the oracle draws by inverse CDF of a supplied u; its CDF must be torch's softmax
    cumsum and its empirical frequencies must follow torch.multinomial's distribution."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(0)
    logits = rng.standard_normal((4096, 4)).astype(np.float32)
    u = rng.random(4096).astype(np.float32)
    actions, cum = restate.act_sample(logits, u)
    want_cum = torch.cumsum(torch.softmax(torch.from_numpy(logits), -1), -1).numpy()
    np.testing.assert_allclose(cum, want_cum, rtol=0, atol=5e-7)  # a few fp32 ulps
    want = (want_cum > u[:, None]).argmax(1)
    near = (np.abs(want_cum - u[:, None]) < 1e-6).any(1)
    assert np.array_equal(actions[~near], want[~near])
    # distribution: one state, many draws, against torch.multinomial's frequencies
    one = np.repeat(logits[:1], 200000, axis=0)
    a, _ = restate.act_sample(one, rng.random(200000).astype(np.float32))
    p = torch.softmax(torch.from_numpy(logits[:1]), -1)[0].numpy().astype(np.float64)
    torch.manual_seed(1)
    t = torch.multinomial(torch.from_numpy(p), 200000, replacement=True).numpy()
    f_ours, f_torch = np.bincount(a, minlength=4) / 2e5, np.bincount(t, minlength=4) / 2e5
    assert np.abs(f_ours - p).max() < 5e-3 and np.abs(f_torch - p).max() < 5e-3


def test_normalize_advantages_matches_torch_std():
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(1)
    adv = rng.standard_normal(257).astype(np.float32) * 3 + 1
    masks = rng.random(257) > 0.1
    got = restate.normalize_advantages(adv, masks)
    sel = torch.from_numpy(adv)[torch.from_numpy(masks)]
    want = (torch.from_numpy(adv) - sel.mean()) / (sel.std() + 1e-8)  # unbiased std, the PPO idiom
    np.testing.assert_allclose(got, want.numpy(), rtol=1e-5, atol=1e-6)


def test_vision_restatements_match_libtorch():
    """restate.resize_area / rgb_to_grayscale against the expressions of src/ai/vision.cc:8-84 in torch
    (interpolate(mode=area), matmul with the grayscale weights) and the constant-image cases of
    test/ai/vision-test.cc:5-30."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(3)
    x = (rng.random((2, 3, 210, 160)) * 255).astype(np.float32)
    want = torch.nn.functional.interpolate(torch.from_numpy(x), size=(84, 84), mode="area").numpy()
    got = restate.resize_area(x)
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-4)
    w = torch.tensor([0.2125, 0.7154, 0.0721])
    g = torch.matmul(torch.from_numpy(want).permute(0, 2, 3, 1), w).numpy()
    np.testing.assert_allclose(restate.rgb_to_grayscale(want), g, rtol=1e-5, atol=1e-4)
    for v in (0.0, 1.0, 255.0):
        assert np.all(restate.resize_area(np.full((1, 210, 160), v, np.float32)) == v)


def test_clipped_value_loss_matches_torch_autograd():
    """The opt-in clipped value loss (north_star / SURVEY D3; NOT in the reference) pinned against the
    CleanRL expression under torch autograd: value, total loss, d/dlogits and d/dvalues."""
    torch = pytest.importorskip("torch")
    c = cases.loss_case(n=193, seed=9)
    rng = np.random.default_rng(4)
    old_v = (c["values"] + 0.15 * rng.standard_normal(193)).astype(np.float32)
    old_v[:5] = c["values"][:5]  # inside the clamp range: both branches equal (a tie)
    clip, vc, ec = 0.1, 0.5, 0.01
    got = restate.loss(c["logits"], c["old_log_probs"], c["actions"], c["advantages"], c["values"],
                       c["returns"], c["masks"], clip, vc, ec, old_values=old_v)
    t = lambda k, **kw: torch.tensor(c[k], **kw)
    z, v = t("logits", requires_grad=True), t("values", requires_grad=True)
    lp = torch.log_softmax(z, -1)
    a = t("actions")
    ratio = torch.exp(lp.gather(-1, a[:, None])[:, 0] - t("old_log_probs").gather(-1, a[:, None])[:, 0])
    adv, ret, ov = t("advantages"), t("returns"), torch.tensor(old_v)
    surr = torch.min(ratio * adv, torch.clamp(ratio, 1 - clip, 1 + clip) * adv)
    v_clipped = ov + torch.clamp(v - ov, -clip, clip)
    vl = 0.5 * torch.max((v - ret) ** 2, (v_clipped - ret) ** 2)
    ent = -(lp.exp() * lp).sum(-1)
    total = -surr + vc * vl - ec * ent
    mask = t("masks").bool()
    loss = (total * mask).sum() / mask.sum()
    loss.backward()
    np.testing.assert_allclose(got["value_losses"], vl.detach().numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(got["loss"], float(loss), rtol=1e-5)
    np.testing.assert_allclose(got["dvalues"], v.grad.numpy(), rtol=1e-4, atol=1e-8)
    np.testing.assert_allclose(got["dlogits"], z.grad.numpy(), rtol=1e-4, atol=1e-8)
    # and off by default == the reference form
    ref = restate.loss(c["logits"], c["old_log_probs"], c["actions"], c["advantages"], c["values"],
                       c["returns"], c["masks"], clip, vc, ec)
    np.testing.assert_array_equal(ref["value_losses"],
                                  (0.5 * (c["values"] - c["returns"]) ** 2).astype(np.float32))


# ---- live-path resize (stb restated; PARITY UNPINNED — no reference vector exists) ---------------
def test_stb_resize_restatement_properties():
    # hand-derived known answer of the stated rules: Mitchell at scale 1/2 on [0,0,255,255]:
    # o=0 weighs in[3] (and its clamp-folded neighbour) by M(1.25)+M(1.75) and in[2] by M(.75), over 2
    M = lambda x: (16 + x * x * (21 * x - 36)) / 18 if x < 1 else (32 + x * (-60 + x * (36 - 7 * x))) / 18
    want0 = 255 * (M(.75) + M(1.25) + M(1.75)) / (2 * (M(.25) + M(.75) + M(1.25) + M(1.75)))
    q, f = restate.resize_grayscale_stb(np.array([[[0, 0, 255, 255]]], np.uint8), 2, 1)
    assert abs(float(f[0, 0, 0]) - want0) < 1e-4 and q[0, 0].tolist() == [28, 227]
    for n_in, n_out in ((210, 84), (160, 84), (84, 84), (10, 33), (7, 3)):
        W = restate.stb_axis_matrix(n_in, n_out)
        np.testing.assert_allclose(W.sum(axis=1), 1.0, atol=1e-6)       # normalised
        np.testing.assert_allclose(W, W[::-1, ::-1], atol=1e-5)          # mirror symmetric (fp32 weights)
        taps = (W != 0).sum(axis=1).max()
        assert taps <= (int(np.ceil(4 * max(n_in / n_out, 1.0))) + 1)    # support 2 each side
    assert np.array_equal(restate.stb_axis_matrix(84, 84), np.eye(84, dtype=np.float32))
    for val in (0, 1, 200, 255):                                         # test/ai/vision-test.cc:5-30 analogue
        img = np.full((1, 210, 160), val, np.uint8)
        assert np.all(restate.resize_grayscale_stb(img, 84, 84)[0] == val)
    rng = np.random.default_rng(0)
    a, b = rng.integers(0, 256, (2, 2, 210, 160), dtype=np.uint8)
    assert np.array_equal(restate.preprocess_gray(a, b), restate.preprocess_gray(b, a))
