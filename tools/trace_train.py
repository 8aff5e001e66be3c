#!/usr/bin/env python
"""Per-launch-site device time of one eager train() call (alegpu_trace_begin/_end).
usage: python tools/trace_train.py [workload] [math]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200"), os.path.join(ROOT, "tests")]
import alegpu  # noqa: E402
import cases  # noqa: E402
import synth  # noqa: E402
from bench import load_workload  # noqa: E402

w = load_workload(sys.argv[1] if len(sys.argv) > 1 else "v0", 1)
HP = w
math = {"bf16": alegpu.MATH_BF16, "tf32": alegpu.MATH_TF32, "fp32": alegpu.MATH_FP32}[sys.argv[2] if len(sys.argv) > 2 else "bf16"]
E, T, H, epochs, nmb = w["E"], w["T"], w["H"], w["epochs"], w["nmb"]
N, mb = E * T, E * T // nmb
tc = cases.train_case(N, seed=5)
L = alegpu.Learner(H, 4, mb, math, HP["lr"])
L.set_params(synth.orthogonal_params(H, 4, seed=42))
batch = alegpu.Batch(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"], tc["masks"])
hp = alegpu.Hyperparameters(HP["clip"], HP["vcoef"], HP["ecoef"], HP["max_norm"])
met = alegpu.Metrics(epochs, nmb, mb)
for _ in range(2):
    L.train(batch, met, epochs, nmb, hp, cuda_graph=False)
alegpu.synchronize(L.stream)
lib = alegpu.lib()
alegpu._check(lib.alegpu_trace_begin(C.c_void_p(L.stream)))
L.train(batch, met, epochs, nmb, hp, cuda_graph=False)
buf = C.create_string_buffer(1 << 16)
alegpu._check(lib.alegpu_trace_end(buf, C.c_size_t(len(buf))))
tot = 0.0
rows = []
for line in buf.value.decode().strip().split("\n"):
    k, n, us = line.rsplit(" ", 2)
    if "AOp = " in k:
        import re
        m = re.search(r"BN = (\d+).*AOp = ([^;\]]+); BOp = ([^;\]]+); Epi = ([^;\]]+)", k)
        k = "gemm_tc<%s> %s x %s -> %s" % tuple(x.replace("alegpu::", "").replace("{anonymous}::", "") for x in m.groups())
    rows.append((k, int(n), float(us)))
    tot += float(us)
print(f"train() eager trace: {tot:.1f} us total, {epochs * nmb} minibatch steps -> {tot / (epochs * nmb):.1f} us/step")
for k, n, us in rows:
    print(f"{us:10.1f} us {n:4d} x {us / n:8.2f} us {100 * us / tot:5.1f}%  {k}")
