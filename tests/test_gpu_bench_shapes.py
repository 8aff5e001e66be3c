"""GPU parity at the BENCHMARKED shapes against the reference's own code.

tests/golden/bench_shapes.npz is produced by tools/make_golden.py from oracle/_ref (the reference's
src/ai/ppo/{losses,train}.cc + restated Network compiled against libtorch-CPU, fp32): a whole
configs/v0.yaml train() (H = 512, 4 epochs x 4 minibatches of 256), one configs/v1.yaml minibatch
(1 280 rows) and one per-GPU sweep minibatch (2 048 rows).  Every math mode the bench can run is held
to SURVEY.md §8c's bounds for it:
  fp32 (CUDA cores)      outputs rel-L2 1e-5..1e-4, per-tensor gradients rel-L2 <= 1e-3
  tf32 (tcgen05 kind::tf32) outputs <= 2e-3, gradients <= 1e-2 and cosine >= 0.9999, train() parameters <= 1e-3
  bf16 (tcgen05 kind::f16)  outputs <= 2e-2, gradients <= 3e-2 and cosine >= 0.999
"""
import os

import numpy as np
import pytest

import alegpu
import cases

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "bench_shapes.npz"))
OFFS = cases.param_offsets(512)
SAMP = cases.tensor_samples(OFFS)
BOUNDS = {  # math: (outputs rel-L2, grad rel-L2, grad cosine, train() update cosine, grad-norm rel)
    alegpu.MATH_FP32: (1e-4, 1e-3, 0.999999, 0.9999, 1e-3),
    alegpu.MATH_TF32: (2e-3, 1e-2, 0.9999, 0.999, 5e-3),
    alegpu.MATH_BF16: (2e-2, 3e-2, 0.999, 0.98, 3e-2),
}
NAMES = {alegpu.MATH_FP32: "fp32", alegpu.MATH_TF32: "tf32", alegpu.MATH_BF16: "bf16"}


def rel_l2(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def cosine(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-300))


def _batch(tc):
    return alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                        tc["masks"])


def _split(flat_samples):
    out, o = [], 0
    for idx in SAMP:
        out.append(flat_samples[o:o + len(idx)])
        o += len(idx)
    return out


@pytest.mark.parametrize("math", [alegpu.MATH_FP32, alegpu.MATH_TF32, alegpu.MATH_BF16])
@pytest.mark.parametrize("name", ["v1_mb", "sweep_mb"])
def test_minibatch_gradients_at_bench_shape(name, math):
    c = cases.BENCH_SHAPES[name]
    params, tc = cases.bench_shape_case(name)
    out_tol, g_tol, g_cos, _, n_tol = BOUNDS[math]
    L = alegpu.Learner(512, 4, c["n"], math, cases.LR)
    L.set_params(params)
    hp = alegpu.Hyperparameters(c["clip"], c["vcoef"], c["ecoef"], cases.MAX_NORM)
    grads, loss, logits, values = L.grads(_batch(tc), hp)
    assert rel_l2(logits, G[name + ".logits"]) <= out_tol, rel_l2(logits, G[name + ".logits"])
    assert rel_l2(values, G[name + ".values"]) <= out_tol, rel_l2(values, G[name + ".values"])
    assert abs(loss - float(G[name + ".loss"])) <= out_tol * abs(float(G[name + ".loss"])) + 1e-5
    want = _split(G[name + ".grads"])
    report = []
    for i, idx in enumerate(SAMP):
        norm = np.linalg.norm(grads[OFFS[i]:OFFS[i + 1]].astype(np.float64))
        report.append((i, rel_l2(grads[idx], want[i]), cosine(grads[idx], want[i]),
                       abs(norm - G[name + ".grad_norms"][i]) / G[name + ".grad_norms"][i]))
    print(f"{name} {NAMES[math]} grad report (tensor, rel-L2, cosine, norm rel):",
          [(i, float(f"{r:.2e}"), round(co, 7), float(f"{nr:.1e}")) for i, r, co, nr in report])
    for i, r, co, nr in report:
        assert r <= g_tol and co >= g_cos and nr <= n_tol, report


@pytest.mark.parametrize("math", [alegpu.MATH_FP32, alegpu.MATH_TF32, alegpu.MATH_BF16])
def test_v0_train_against_the_reference(math):
    """16 optimiser steps of configs/v0.yaml (4 epochs x 4 contiguous minibatches of 256 rows, CUDA
    graph, H = 512) vs the reference's train(): parameters, the update direction per tensor, the
    per-minibatch loss and pre-clip gradient norms."""
    name = "v0_train"
    c = cases.BENCH_SHAPES[name]
    params, tc = cases.bench_shape_case(name)
    _, _, _, u_cos, n_tol = BOUNDS[math]
    mb = c["n"] // c["nmb"]
    L = alegpu.Learner(512, 4, mb, math, cases.LR)
    L.set_params(params)
    met = alegpu.Metrics(c["epochs"], c["nmb"], mb)
    hp = alegpu.Hyperparameters(c["clip"], c["vcoef"], c["ecoef"], cases.MAX_NORM)
    L.train(_batch(tc), met, c["epochs"], c["nmb"], hp, cuda_graph=True)
    got = L.get_params()
    assert L.get_adam()[2] == c["epochs"] * c["nmb"]
    flat = np.concatenate(SAMP)
    assert rel_l2(got[flat], G[name + ".params"]) <= 1e-3, rel_l2(got[flat], G[name + ".params"])
    upd, want = _split((got - params)[flat]), _split(G[name + ".update"])
    report = [(i, cosine(upd[i], want[i]), rel_l2(upd[i], want[i])) for i in range(12)]
    print(f"v0 train() {NAMES[math]} update report (tensor, cosine, rel-L2):",
          [(i, round(co, 6), float(f"{r:.2e}")) for i, co, r in report])
    m = met.numpy()
    # Adam normalises every coordinate's step to ~lr, so a coordinate whose gradient is ~0 moves in a
    # noise-determined direction: compare the update where it carries signal (weights, not the
    # biases that start at 0), and the metrics everywhere
    for i, co, r in report:
        if i in (0, 2, 4, 6):
            assert co >= u_cos, report
    np.testing.assert_allclose(m["loss"][:, :, 0], G[name + ".loss"], rtol=10 * n_tol, atol=1e-4)
    np.testing.assert_allclose(m["clipped_gradients"], G[name + ".clipped_gradients"], rtol=10 * n_tol)
    for k in ("clipped_losses", "value_losses", "entropies", "ratio"):
        np.testing.assert_allclose(m[k][:, :, :32], G[name + "." + k], rtol=20 * n_tol, atol=1e-3)
