#!/usr/bin/env python
"""train() time per minibatch step against the minibatch size (H=512, graph replay): where the
fixed per-step cost (launch chain, optimiser over 1.69 M parameters) ends and the GEMMs begin."""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import alegpu, cases, synth
from bench import load_workload
HP = load_workload("v0", 1)  # configs/v0.yaml hyper-parameters
H, nmb, epochs = 512, 4, 4
hp = alegpu.Hyperparameters(HP["clip"], HP["vcoef"], HP["ecoef"], HP["max_norm"])
for mb in (8, 32, 64, 128, 256, 512, 1024):
    N = mb * nmb
    tc = cases.train_case(N, seed=5)
    L = alegpu.Learner(H, 4, mb, alegpu.MATH_BF16, HP["lr"])
    L.set_params(synth.orthogonal_params(H, 4, seed=42))
    batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"], tc["masks"])
    met = alegpu.Metrics(epochs, nmb, mb)
    for _ in range(3):
        L.train(batch, met, epochs, nmb, hp, cuda_graph=True)
    t0 = time.perf_counter()
    K = 20
    for _ in range(K):
        L.train(batch, met, epochs, nmb, hp, cuda_graph=True, sync=False)
    alegpu.synchronize(L.stream)
    dt = (time.perf_counter() - t0) / K
    print(f"mb={mb:5d}  train() {dt*1e3:7.3f} ms  -> {dt*1e6/(epochs*nmb):7.1f} us per minibatch step")
    L.close()
