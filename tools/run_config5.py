#!/usr/bin/env python
"""BASELINE.json config 5: multi-threaded CPU frame source feeding the pinned ring into the GPU
learner, end to end: collect a rollout (policy call per step, CPU emulator threads, per-step
hand-off) then train() on it, for K iterations.  ALE and its ROMs are not in this image, so the
frame source is the library's Breakout-shaped emulator (collector.cu).  Prints one JSON line.
usage: run_config5.py [E] [T] [workers] [iterations]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200"), os.path.join(ROOT, "tests")]
import alegpu  # noqa: E402
import synth  # noqa: E402
from bench import load_workload  # noqa: E402
HP = load_workload("v0", 1)  # configs/v0.yaml hyper-parameters

E = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 128
workers = int(sys.argv[3]) if len(sys.argv) > 3 else (os.cpu_count() or 1)
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
H, A, epochs, nmb = 512, 4, 4, 4
mb = E * T // nmb
L = alegpu.Learner(H, A, max(mb, E), alegpu.MATH_BF16, HP["lr"])
L.set_params(synth.orthogonal_params(H, A, seed=42))
buf = alegpu.Buffer(E, T, A)
col = alegpu.Collector(E, T, A, num_workers=workers, seed=42)
met = alegpu.Metrics(epochs, nmb, mb)
hp = alegpu.Hyperparameters(HP["clip"], HP["vcoef"], HP["ecoef"], HP["max_norm"])
t_roll = t_train = 0.0
log = {}
for i in range(iters + 1):
    t0 = time.perf_counter()
    batch, log = col.rollout(buf, L, HP["gamma"], HP["lam"])
    t1 = time.perf_counter()
    L.train(batch, met, epochs, nmb, hp, cuda_graph=True)
    summary = met.summary(L.stream)
    t2 = time.perf_counter()
    if i:  # iteration 0 warms up (graph capture, allocations)
        t_roll += t1 - t0
        t_train += t2 - t1
total = t_roll + t_train
print(json.dumps({
    "metric": "agent-steps/sec end to end (CPU frame threads -> pinned ring -> GPU learner)",
    "value": iters * E * T / total, "unit": "agent-steps/s", "n_gpus": 1,
    "config": {"workload": "config5", "envs": E, "horizon": T, "workers": workers, "hidden_size": H,
               "num_epochs": epochs, "num_mini_batches": nmb,
               "frame_source": "built-in Breakout-shaped emulator (ALE absent from the image)"},
    "rollout_s_per_iteration": t_roll / iters, "train_s_per_iteration": t_train / iters,
    "rollout_us_per_env_step": 1e6 * t_roll / (iters * E * T),
    "log": log, "mean_loss": summary["mean_loss"], "mean_entropy": summary["mean_entropy"],
    "host_cores": os.cpu_count()}))
