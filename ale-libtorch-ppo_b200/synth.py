"""Seeded synthetic inputs for the learner hot path: ALE-shaped frame streams
and reward/done streams obeying the rollout invariants of the reference
(src/ai/rollout.cc:239-265: episode_start[t+1] = terminal[t] | truncation[t]; the three
flags never co-occur, src/ai/gae.cc:49-53)."""
import numpy as np


def rollout_scalars(E, T, actions=4, seed=42, first_rollout=True):
    rng = np.random.default_rng(seed)
    rewards = np.where(rng.random((E, T)) < 0.95, 0.0,
                       rng.choice(np.array([1.0, 4.0, 7.0]), size=(E, T))).astype(np.float32)
    terminals = np.zeros((E, T), np.uint8)
    truncations = np.zeros((E, T), np.uint8)
    starts = np.zeros((E, T), np.uint8)
    pending = np.full(E, 1 if first_rollout else 0, np.uint8)
    t_draw, u_draw = rng.random((E, T)), rng.random((E, T))
    for t in range(T):
        starts[:, t] = pending
        live = pending == 0
        term = live & (t_draw[:, t] < 0.01)
        trunc = live & ~term & (u_draw[:, t] < 0.001)
        terminals[:, t], truncations[:, t] = term, trunc
        pending = (term | trunc).astype(np.uint8)
    rewards[starts == 1] = 0.0  # a reset slot carries reward 0 (rollout.cc:303-308)
    return dict(
        rewards=rewards, terminals=terminals, truncations=truncations, episode_starts=starts,
        logits=rng.standard_normal((E, T, actions)).astype(np.float32),
        values=rng.standard_normal((E, T)).astype(np.float32),
        next_values=rng.standard_normal(E).astype(np.float32),
        actions=rng.integers(0, actions, (E, T)).astype(np.int64))


def frames(n, seed=42, kind="uniform"):
    """[n,2,210,160,3] u8 HWC frame pairs. 'uniform' = iid bytes (worst case for parity);
    'ale' = piecewise-constant Breakout-like rectangles on a black field."""
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        return rng.integers(0, 256, (n, 2, 210, 160, 3), dtype=np.uint8)
    out = np.zeros((n, 2, 210, 160, 3), np.uint8)
    for i in range(n):
        for f in range(2):
            for _ in range(12):
                y, x = rng.integers(0, 200), rng.integers(0, 150)
                h, w = rng.integers(2, 12), rng.integers(4, 40)
                out[i, f, y:y + h, x:x + w] = rng.integers(0, 256, 3, dtype=np.uint8)
    return out


def observations(n, seed=42):
    return np.random.default_rng(seed).integers(0, 256, (n, 4, 84, 84), dtype=np.uint8)


def orthogonal_params(hidden=512, actions=4, seed=42):
    """Random-init weights of the reference architecture in `parameters()` order: orthogonal
    rows with gain sqrt(2) (policy head 0.01, value head 1) and zero biases — the scheme of
    layer_init, src/bin/train.cc:212-228 (torch::nn::init::orthogonal_ flattens to [out, rest]
    and orthogonalises via QR with sign correction)."""
    rng = np.random.default_rng(seed)

    def ortho(rows, cols, gain):
        a = rng.standard_normal((rows, cols))
        flip = rows < cols
        if flip:
            a = a.T
        q, r = np.linalg.qr(a)
        q = q * np.sign(np.diag(r))
        if flip:
            q = q.T
        return (gain * q).astype(np.float32).ravel()

    g = np.sqrt(2.0)
    parts = [ortho(32, 4 * 8 * 8, g), np.zeros(32, np.float32),
             ortho(64, 32 * 4 * 4, g), np.zeros(64, np.float32),
             ortho(64, 64 * 3 * 3, g), np.zeros(64, np.float32),
             ortho(hidden, 3136, g), np.zeros(hidden, np.float32),
             ortho(actions, hidden, 0.01), np.zeros(actions, np.float32),
             ortho(1, hidden, 1.0), np.zeros(1, np.float32)]
    return np.concatenate(parts)
