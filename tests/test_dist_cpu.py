"""world_size-2 gloo test (CPU) of the multi-GPU algebra the learner uses.

Each rank owns half of the envs of every minibatch, back-propagates the UN-normalised masked loss
sum of its rows, and the ranks all-reduce (sum) the flat gradient and the unmasked-row count; the
gradient divided by the GLOBAL count must equal the single-process gradient of the whole minibatch
(src/ai/ppo/losses.cc:19 normalises by the mask count of the full minibatch), up to fp32 summation
order.  The per-rank maths is the oracle's (restate.c); the exchange is torch.distributed/gloo, the
same plumbing bench.py uses to hand out the NCCL unique id."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import cases
    from oracle import restate
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, n = 32, 16
    params = cases.net_params(H)
    tc = cases.train_case(n)
    rows = np.arange(n).reshape(world, n // world)[rank]  # this rank's env slice of the minibatch
    lg, v, acts = restate.net_forward(params, H, tc["obs"][rows], keep_acts=True)
    lo = restate.loss(lg, tc["log_probs"][rows], tc["actions"][rows], tc["advantages"][rows], v,
                      tc["returns"][rows], tc["masks"][rows], cases.CLIP, cases.VCOEF, cases.ECOEF)
    count = float(tc["masks"][rows].sum())
    # restate.loss normalises by the LOCAL count: undo it to get the un-normalised sums
    g_local = restate.net_backward(params, H, acts, lo["dlogits"] * np.float32(count),
                                   lo["dvalues"] * np.float32(count))
    buf = torch.from_numpy(np.concatenate([g_local, [np.float32(count)],
                                           [np.float32(lo["loss"] * count)]]).astype(np.float32))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    total = buf.numpy()
    g = total[:-2] / total[-2]
    np.save(os.path.join(out_dir, f"grad_{rank}.npy"), g)
    np.save(os.path.join(out_dir, f"loss_{rank}.npy"), total[-1] / total[-2])
    dist.destroy_process_group()


def test_sharded_gradient_equals_single_process(tmp_path):
    import torch.multiprocessing as mp
    sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
    import cases
    from oracle import restate
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    H, n = 32, 16
    params, tc = cases.net_params(H), cases.train_case(n)
    lg, v, acts = restate.net_forward(params, H, tc["obs"], keep_acts=True)
    lo = restate.loss(lg, tc["log_probs"], tc["actions"], tc["advantages"], v, tc["returns"],
                      tc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    want = restate.net_backward(params, H, acts, lo["dlogits"], lo["dvalues"])
    g0, g1 = (np.load(tmp_path / f"grad_{r}.npy") for r in range(world))
    assert np.array_equal(g0, g1)  # all ranks hold the same reduced gradient
    err = np.linalg.norm((g0 - want).astype(np.float64)) / np.linalg.norm(want.astype(np.float64))
    assert err <= 1e-5, err
    assert abs(float(np.load(tmp_path / "loss_0.npy")) - lo["loss"]) <= 1e-5 * abs(lo["loss"]) + 1e-6


def _adv_worker(rank, world, port, out_dir):
    import numpy as np
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(77)
    adv = (rng.standard_normal(64) * 2 + 0.5).astype(np.float32)
    masks = rng.random(64) > 0.15
    lo, hi = rank * 32, (rank + 1) * 32  # each rank holds half of the global minibatch
    a, m = adv[lo:hi].astype(np.float64), masks[lo:hi]
    # what adv_stats_kernel leaves per rank: (count, sum, sum of squares) in double; one
    # all-reduce(sum) of 3 doubles gives the GLOBAL statistics (learner.cu, ncclDouble)
    st = torch.tensor([m.sum(), a[m].sum(), (a[m] ** 2).sum()], dtype=torch.float64)
    dist.all_reduce(st)
    c, mean = st[0].item(), st[1].item() / st[0].item()
    var = (st[2].item() - c * mean * mean) / (c - 1.0)
    out = ((a - mean) / (np.sqrt(max(var, 0.0)) + 1e-8)).astype(np.float32)
    np.save(os.path.join(out_dir, f"adv_{rank}.npy"), out)
    dist.destroy_process_group()


def test_global_advantage_statistics_world2(tmp_path):
    """normalize_advantages with world_size 2: local (count, sum, sumsq) all-reduced == the
    single-process normalisation of the whole minibatch (oracle/restate.normalize_advantages)."""
    import torch.multiprocessing as mp
    sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
    from oracle import restate
    world, port = 2, _free_port()
    mp.spawn(_adv_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(77)
    adv = (rng.standard_normal(64) * 2 + 0.5).astype(np.float32)
    masks = rng.random(64) > 0.15
    want = restate.normalize_advantages(adv, masks)
    got = np.concatenate([np.load(tmp_path / f"adv_{r}.npy") for r in range(world)])
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
