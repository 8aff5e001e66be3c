"""Multi-GPU parity of the data-parallel path: G ranks, each training its interleaved
share of one batch through alegpu_learner_train with world_size = G (NCCL all-reduce of the
un-normalised gradient + mask counts, or the fused NVLink exchange), must end where ONE GPU training
the whole batch ends, and where the oracle ends.  Needs >= 2 GPUs (skipped otherwise; the log of a
`gpurun --gpus 2` run is kept under profiles/)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import alegpu
import cases
from oracle import restate

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = {alegpu.MATH_FP32: "fp32", alegpu.MATH_TF32: "tf32", alegpu.MATH_BF16: "bf16"}


def rel_l2(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def run_ranks(tmp_path, world, math, graph, H, N, nmb, epochs, extras="", env=None):
    work = str(tmp_path)
    for f in os.listdir(work):
        os.remove(os.path.join(work, f))
    e = dict(os.environ)
    e.update(env or {})
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "multi_worker.py"), str(r), str(world),
                               work, str(math), str(int(graph)), str(H), str(N), str(nmb), str(epochs),
                               extras], env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = []
    try:
        for p in procs:
            outs.append(p.communicate(timeout=240)[0])
    finally:
        for p in procs:  # exactly the processes started here
            if p.poll() is None:
                p.kill()
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"rank {r} done" in o, (r, p.returncode, o[-2000:])
    return [np.load(os.path.join(work, f"rank{r}.npz")) for r in range(world)]


def single_gpu(math, graph, H, N, nmb, epochs, extras=""):
    params = cases.net_params(H, seed=21)
    tc = cases.train_case(N, seed=77)
    old_v = (tc["returns"] + 0.3 * np.random.default_rng(2).standard_normal(N)).astype(np.float32)
    batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                         tc["masks"], values=old_v)
    L = alegpu.Learner(H, 4, N // nmb, math, cases.LR)
    L.set_params(params)
    met = alegpu.Metrics(epochs, nmb, N // nmb)
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    for _ in range(2 if graph else 1):
        L.train(batch, met, epochs, nmb, hp, cuda_graph=graph, normalize_advantages="norm" in extras,
                clip_value_loss="clipv" in extras)
    return L.get_params(), met.numpy(), (params, tc, old_v)


def world_size():
    n = alegpu.device_count()
    return 4 if n >= 4 else 2 if n >= 2 else 0


@pytest.mark.parametrize("env", [{}, {"ALEGPU_NO_COMM_OVERLAP": "1"}, {"ALEGPU_NO_FUSED_AR": "1"}],
                         ids=["default", "no-overlap", "nccl-only"])
@pytest.mark.parametrize("graph", [False, True], ids=["eager", "graph"])
@pytest.mark.parametrize("math", [alegpu.MATH_FP32, alegpu.MATH_TF32, alegpu.MATH_BF16],
                         ids=["fp32", "tf32", "bf16"])
def test_multi_gpu_train_equals_single_gpu(tmp_path, math, graph, env):
    G = world_size()
    if G < 2:
        pytest.skip("needs >= 2 GPUs")
    H, N, nmb, epochs = 64, 256, 2, 2  # minibatch 128 rows = 128/G per rank
    ranks = run_ranks(tmp_path, G, math, graph, H, N, nmb, epochs, env=env)
    # the default exchanges gradients inside the optimiser kernel over NVLink peer mappings; the other
    # two environments are the ncclAllReduce path (with / without the overlapped FC bucket)
    want_mode = "nccl" if "ALEGPU_NO_FUSED_AR" in env else "fused-nvlink"
    if "ALEGPU_NO_COMM_OVERLAP" in env:
        want_mode = str(ranks[0]["comm_mode"])  # overlap only matters on the NCCL path
    assert all(str(r["comm_mode"]) == want_mode for r in ranks), [str(r["comm_mode"]) for r in ranks]
    for r in ranks[1:]:  # replicated parameters stay bit-identical on every rank
        assert np.array_equal(r["params"], ranks[0]["params"])
        assert np.array_equal(r["loss"], ranks[0]["loss"])  # the GLOBAL minibatch loss on every rank
        assert np.array_equal(r["clipped_gradients"], ranks[0]["clipped_gradients"])
    steps = epochs * nmb * (2 if graph else 1)
    assert int(ranks[0]["step"]) == steps
    p1, m1, (params, tc, _) = single_gpu(math, graph, H, N, nmb, epochs)
    tol = {alegpu.MATH_FP32: 1e-5, alegpu.MATH_TF32: 1e-4, alegpu.MATH_BF16: 1e-3}[math]
    err = rel_l2(ranks[0]["params"], p1)
    print(f"G={G} {NAMES[math]} graph={graph} {env} [{ranks[0]['comm_mode']}]: params rel-L2 vs 1 GPU {err:.2e}")
    assert err <= tol, err
    mt = 1e-4 if math == alegpu.MATH_FP32 else 2e-2
    np.testing.assert_allclose(ranks[0]["loss"][:, :, 0], m1["loss"][:, :, 0], rtol=mt, atol=1e-5)
    np.testing.assert_allclose(ranks[0]["clipped_gradients"], m1["clipped_gradients"], rtol=mt)
    # per-sample metrics stay sharded: rank g's [ep,k,:] segment is its share of the minibatch
    mb, per = N // nmb, N // nmb // G
    for g, r in enumerate(ranks):
        for k in range(nmb):
            np.testing.assert_allclose(r["total_losses"][0, k], m1["total_losses"][0, k, g * per:(g + 1) * per],
                                       rtol=mt, atol=1e-4)
    if not graph:  # and the oracle on the whole batch (fp32 CPU restatement of the reference's train())
        z = np.zeros_like(params)
        want = restate.train(params, z, z, 0, H, tc["obs"], tc["actions"], tc["log_probs"],
                             tc["advantages"], tc["returns"], tc["masks"], epochs, nmb, cases.CLIP,
                             cases.VCOEF, cases.ECOEF, cases.MAX_NORM, cases.LR)[0]
        assert rel_l2(ranks[0]["params"], want) <= max(tol, 1e-4) * 10


def test_multi_gpu_extras_use_global_statistics(tmp_path):
    """advantage normalisation (global count / mean / std over the ranks) and the clipped value loss
    on 2+ ranks == 1 GPU on the whole batch."""
    G = world_size()
    if G < 2:
        pytest.skip("needs >= 2 GPUs")
    H, N, nmb, epochs = 64, 256, 2, 1
    ranks = run_ranks(tmp_path, G, alegpu.MATH_FP32, False, H, N, nmb, epochs, extras="norm,clipv")
    p1, m1, _ = single_gpu(alegpu.MATH_FP32, False, H, N, nmb, epochs, extras="norm,clipv")
    assert rel_l2(ranks[0]["params"], p1) <= 1e-5
    np.testing.assert_allclose(ranks[0]["loss"][:, :, 0], m1["loss"][:, :, 0], rtol=1e-4, atol=1e-6)
