"""Seeded test inputs shared by tools/make_golden.py (which runs the REFERENCE's own code on
them to produce tests/golden/*.npz) and by the tests (which re-generate the same inputs and
compare the oracle restatement / the CUDA path against those goldens)."""
import numpy as np

import synth

GAMMA, LAMBDA = 0.99, 0.95
CLIP, VCOEF, ECOEF, MAX_NORM, LR = 0.1, 0.5, 0.01, 0.5, 2.5e-4

# ---- the six known-answer cases of /root/reference/test/ai/gae-test.cc:6-257 -----------------
def _kat(rewards, values, next_values, terminals=(), truncations=(), starts=()):
    r = np.asarray(rewards, np.float32)
    z = lambda idx: _flags(r.shape, idx)
    return dict(rewards=r, values=np.asarray(values, np.float32),
                next_values=np.asarray(next_values, np.float32), terminals=z(terminals),
                truncations=z(truncations), episode_starts=z(starts))


def _flags(shape, idx):
    f = np.zeros(shape, np.uint8)
    for e, t in idx:
        f[e, t] = 1
    return f


GAE_KATS = {
    "SimpleNoTerminalsNoTruncations": _kat([[1, 1, 1]], [[.5, .5, .5]], [.5]),
    "TerminalHandling": _kat([[1, 1, 1]], [[.5, .5, .5]], [0.], terminals=[(0, 2)]),
    "TruncationHandling": _kat([[1, 1, 1]], [[.5, .5, .5]], [0.], truncations=[(0, 2)]),
    "EpisodeStartHandling": _kat([[1, 1, 1]], [[.5, .5, .5]], [.5], starts=[(0, 1)]),
    "MultipleEnvironments": _kat([[1, 1, 1], [.5, .5, .5]], [[.5, .5, .5], [.3, .3, .3]],
                                 [.5, .3], terminals=[(1, 2)], starts=[(0, 1)]),
    "MixedStateFlags": _kat([[1] * 5], [[.5] * 5], [.5], terminals=[(0, 2)],
                            truncations=[(0, 4)], starts=[(0, 3)]),
}


def gae_kat_expected(c, gamma=np.float32(GAMMA), lam=np.float32(LAMBDA)):
    """The inline scalar re-derivation each reference test compares against (fp32 arithmetic,
    priority start > terminal > truncation > running; gae-test.cc:170-190)."""
    f = np.float32
    r, v = c["rewards"], c["values"]
    E, T = r.shape
    out = np.zeros((E, T), np.float32)
    for e in range(E):
        last, nv = f(0), f(c["next_values"][e])
        for i in range(T - 1, -1, -1):
            delta = f(f(r[e, i] + f(gamma * nv)) - v[e, i])
            if c["episode_starts"][e, i]:
                adv = f(0)
            elif c["terminals"][e, i]:
                adv = f(r[e, i] - v[e, i])
            elif c["truncations"][e, i]:
                adv = delta
            else:
                adv = f(delta + f(f(gamma * lam) * last))
            out[e, i] = adv
            last, nv = adv, v[e, i]
    return out


# ---- seeded cases ------------------------------------------------------------------------------
def gae_case(E=5, T=37, seed=7):
    return synth.rollout_scalars(E, T, seed=seed)


def frames_case():
    """5 frame pairs: 3 iid-uniform (worst case) + 2 ALE-shaped."""
    return np.concatenate([synth.frames(3, seed=11), synth.frames(2, seed=12, kind="ale")])


def loss_case(n=96, A=4, seed=5):
    rng = np.random.default_rng(seed)
    old = rng.standard_normal((n, A)).astype(np.float32)
    old = old - np.log(np.exp(old).sum(-1, keepdims=True)).astype(np.float32)
    logits = (old + 0.3 * rng.standard_normal((n, A))).astype(np.float32)
    logits[:8] = old[:8]  # ratio == 1 exactly inside the clip range (tie in torch.min)
    return dict(logits=logits, old_log_probs=old.astype(np.float32),
                actions=rng.integers(0, A, n).astype(np.int64),
                advantages=rng.standard_normal(n).astype(np.float32),
                values=rng.standard_normal(n).astype(np.float32),
                returns=rng.standard_normal(n).astype(np.float32),
                masks=(rng.random(n) > 0.1).astype(np.uint8))


PARAM_SHAPES = lambda H, A=4: [(32, 4, 8, 8), (32,), (64, 32, 4, 4), (64,), (64, 64, 3, 3), (64,),
                               (H, 3136), (H,), (A, H), (A,), (1, H), (1,)]


def net_params(H=32, A=4, seed=3):
    """He-scaled random weights (any weights are valid for forward/backward parity)."""
    rng = np.random.default_rng(seed)
    out = []
    for s in PARAM_SHAPES(H, A):
        if len(s) == 1:
            out.append(0.1 * rng.standard_normal(s))
        else:
            fan_in = int(np.prod(s[1:]))
            out.append(rng.standard_normal(s) * np.sqrt(2.0 / fan_in))
    return np.concatenate([x.ravel() for x in out]).astype(np.float32)


def param_offsets(H=32, A=4):
    sizes = [int(np.prod(s)) for s in PARAM_SHAPES(H, A)]
    return np.concatenate([[0], np.cumsum(sizes)])


def train_case(N=16, A=4, seed=109):
    """seed 109: no conv pre-activation of net_params(32) lies within 1.5e-5 of zero, so no
    ReLU mask can flip under fp32 summation-order noise (~2e-6) — a flipped tie changes the
    conv1/conv2 gradients by ~1e-3 relative and would swamp the 1e-4 tolerance."""
    lc = loss_case(N, A, seed + 1)
    return dict(obs=synth.observations(N, seed), actions=lc["actions"],
                log_probs=lc["old_log_probs"], advantages=lc["advantages"],
                returns=lc["returns"], masks=lc["masks"])


def sample_idx(P, k=4096, seed=1):
    return np.sort(np.random.default_rng(seed).choice(P, size=min(k, P), replace=False))


# ---- the benchmarked shapes (bench.py workloads; hyper-parameters of configs/v0.yaml / v1.yaml) ----
BENCH_SHAPES = {
    # whole train(): configs/v0.yaml — 8 envs x 128 steps, 4 epochs x 4 minibatches of 256
    "v0_train": dict(n=1024, epochs=4, nmb=4, clip=0.1, vcoef=0.5, ecoef=0.01, pseed=42, dseed=201),
    # one minibatch of configs/v1.yaml (4096 x 5 / 16 = 1280 rows)
    "v1_mb": dict(n=1280, clip=0.2, vcoef=0.4, ecoef=0.01, pseed=42, dseed=203),
    # one per-GPU minibatch of the 4096 x 128 sweep on 8 GPUs (16 384 / 8 = 2048 rows)
    "sweep_mb": dict(n=2048, clip=0.1, vcoef=0.5, ecoef=0.01, pseed=42, dseed=205),
}


def bench_shape_case(name):
    """(params, train_case) of a BENCH_SHAPES entry: the reference's orthogonal init (seed 42) and a
    seeded batch of the shape's row count, H = 512."""
    c = BENCH_SHAPES[name]
    return synth.orthogonal_params(512, 4, seed=c["pseed"]), train_case(c["n"], seed=c["dseed"])


def tensor_samples(P_offsets, k=2048, seed=7):
    """per-tensor sample indices (all of a tensor if it has <= k elements) into the flat vector"""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(len(P_offsets) - 1):
        lo, hi = int(P_offsets[i]), int(P_offsets[i + 1])
        n = hi - lo
        out.append(lo + (np.arange(n) if n <= k else np.sort(rng.choice(n, k, replace=False))))
    return out
