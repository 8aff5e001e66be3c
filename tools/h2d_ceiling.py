#!/usr/bin/env python
"""Host -> device copy ceiling of the box, one rank per GPU (torchrun): every rank copies a pinned
buffer to its own GPU, first ALONE (ranks take turns), then ALL AT ONCE.  The e2e bench number is this
ceiling divided by 201 600 B per agent step (two raw 210x160x3 frames by definition of the metric).

    python -m torch.distributed.run --nproc-per-node N tools/h2d_ceiling.py [MB per copy]
"""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200")]
import torch
import torch.distributed as dist

import alegpu

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = alegpu.lib()
alegpu._check(lib.alegpu_set_device(local))
n = mb << 20
h = C.c_void_p()
alegpu._check(lib.alegpu_host_alloc(C.byref(h), C.c_size_t(n)))
C.memset(h, 1, n)
d = alegpu.DeviceArray((n,), "uint8")
st = C.c_void_p()
alegpu._check(lib.alegpu_stream_create(C.byref(st)))


def copy_gbs(reps=3):
    alegpu._check(lib.alegpu_memcpy_h2d(C.c_void_p(d.ptr), h, C.c_size_t(n), st))
    alegpu.synchronize(st.value)
    if world > 1:
        dist.barrier()
    t0 = time.time()
    for _ in range(reps):
        alegpu._check(lib.alegpu_memcpy_h2d(C.c_void_p(d.ptr), h, C.c_size_t(n), st))
    alegpu.synchronize(st.value)
    return reps * n / (time.time() - t0) / 1e9


alone = 0.0
for r in range(world):  # ranks take turns; the others wait at the barriers inside copy_gbs
    if r == rank:
        alone = copy_gbs()
    elif world > 1:
        dist.barrier()
    if world > 1:
        dist.barrier()
together = copy_gbs()
vals = torch.tensor([alone, together], device="cuda", dtype=torch.float64)
if world > 1:
    allv = [torch.zeros_like(vals) for _ in range(world)]
    dist.all_gather(allv, vals)
else:
    allv = [vals]
if rank == 0:
    a = [float(v[0]) for v in allv]
    t = [float(v[1]) for v in allv]
    numa = {}
    try:
        for i in range(world):
            bus = torch.cuda.get_device_properties(i).pci_bus_id
            numa[i] = bus
    except Exception:
        pass
    print(json.dumps({"gpus": world, "mb_per_copy": mb, "alone_gbs_per_gpu": [round(x, 1) for x in a],
                      "together_gbs_per_gpu": [round(x, 1) for x in t],
                      "together_aggregate_gbs": round(sum(t), 1),
                      "e2e_ceiling_agent_steps_per_s": round(sum(t) * 1e9 / 201600),
                      "host_cpus": os.cpu_count()}))
os._exit(0)
