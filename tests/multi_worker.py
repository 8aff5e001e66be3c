"""One rank of the multi-GPU parity test (tests/test_gpu_multi.py): trains its interleaved share of a
seeded batch through alegpu_learner_train with world_size > 1 and writes what it ended up with.

    python multi_worker.py <rank> <world> <workdir> <math> <graph 0|1> <H> <N> <nmb> <epochs> [extras]

The ncclUniqueId travels through <workdir>/uid.bin (rank 0 writes it) — no torch.distributed here, so
the only NCCL user in the process is libalegpu.so."""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path[:0] = [ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200"), HERE]
import ctypes as C  # noqa: E402

import alegpu  # noqa: E402
import cases  # noqa: E402


def main():
    rank, world, work = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    math, graph = int(sys.argv[4]), bool(int(sys.argv[5]))
    H, N, nmb, epochs = (int(x) for x in sys.argv[6:10])
    extras = sys.argv[10] if len(sys.argv) > 10 else ""
    alegpu._check(alegpu.lib().alegpu_set_device(rank))
    uid_path = os.path.join(work, "uid.bin")
    if rank == 0:
        raw = C.create_string_buffer(128)
        alegpu._check(alegpu.lib().alegpu_nccl_unique_id(raw))
        with open(uid_path + ".tmp", "wb") as f:
            f.write(raw.raw)
        os.replace(uid_path + ".tmp", uid_path)
    t0 = time.time()
    while not os.path.exists(uid_path):
        assert time.time() - t0 < 60, "no uid"
        time.sleep(0.01)
    uid = open(uid_path, "rb").read()
    params = cases.net_params(H, seed=21)
    tc = cases.train_case(N, seed=77)
    old_v = (tc["returns"] + 0.3 * np.random.default_rng(2).standard_normal(N)).astype(np.float32)
    rows = alegpu.shard_rows(N, nmb, world, rank)  # this rank's share of every minibatch
    mb = N // nmb // world
    batch = alegpu.Batch(tc["obs"][rows], tc["actions"][rows], tc["log_probs"][rows],
                         tc["advantages"][rows], tc["returns"][rows], tc["masks"][rows],
                         values=old_v[rows])
    L = alegpu.Learner(H, 4, mb, math, cases.LR, device=rank, world_size=world, rank=rank,
                       nccl_unique_id=uid)
    L.set_params(params)
    met = alegpu.Metrics(epochs, nmb, mb)
    hp = alegpu.Hyperparameters(cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    for _ in range(2 if graph else 1):  # graph mode: second call replays the cached graph
        L.train(batch, met, epochs, nmb, hp, cuda_graph=graph,
                normalize_advantages="norm" in extras, clip_value_loss="clipv" in extras)
    m = met.numpy()
    np.savez(os.path.join(work, f"rank{rank}.npz"), params=L.get_params(), loss=m["loss"],
             clipped_gradients=m["clipped_gradients"], total_losses=m["total_losses"],
             step=L.get_adam()[2], comm_mode=L.comm_mode())
    L.close()  # graceful NCCL teardown on every rank
    print(f"rank {rank} done", flush=True)


if __name__ == "__main__":
    main()
