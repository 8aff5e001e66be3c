#!/usr/bin/env python
"""bench.py — agent-steps/sec through preprocess + GAE + PPO update (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload v0|v1|debug|sweep]
                    [--math fp32|bf16] [--impl ours|reference]

A "step" is one learner iteration over one rollout batch of synthetic input:
  preprocess (2-frame max + area resize + gray + 4-frame stack) of E*T frame pairs
  -> Buffer::get (reward clamp + GAE + returns + masks) + prepare_batch (log-softmax)
  -> train(): num_epochs x num_mini_batches of {Nature-CNN fwd/bwd, PPO loss, clip, Adam}.
`value`  : inputs (raw frames [T,E,2,210,160,3] u8, reward/flag/logit/value streams) already
           resident in HBM when the timed region starts; CUDA events on the launching stream.
`e2e`    : the same iteration through the reference-facing API with HOST buffers: the pinned
           frame ring is cudaMemcpyAsync'ed slab by slab on a side stream (alegpu_rollout_ingest),
           scalar streams are H2D-copied, and the per-minibatch loss/grad-norm metrics are read
           back D2H, all inside the timed region.
N > 1 (torchrun): one process per GPU, E envs per GPU (weak scaling: the global batch is N
times larger), the flat gradient is NCCL-allreduced every minibatch.
`--impl reference` times the reference's own CPU implementation (oracle/_ref: the reference
sources compiled unmodified against libtorch-CPU) with all host threads.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "ale-libtorch-ppo_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "agent-steps/sec via preprocess+GAE+PPO update"
WORKLOADS = {  # configs/*.yaml shapes; E is per GPU
    "debug": dict(E=8, T=32, H=32, epochs=2, nmb=4),
    "v0": dict(E=8, T=128, H=512, epochs=4, nmb=4),
    "v1": dict(E=4096, T=5, H=512, epochs=1, nmb=16),
    # BASELINE.json config 4: 4096 envs x 128 steps env-sharded over 8 GPUs = 512 envs per GPU
    # (weak-scaling form: the per-GPU slice is fixed, N GPUs hold N*512 envs); minibatch 2048/GPU
    "sweep": dict(E=512, T=128, H=512, epochs=4, nmb=32),
}
HP = dict(clip=0.1, vcoef=0.5, ecoef=0.01, max_norm=0.5, lr=2.5e-4, gamma=0.99, lam=0.95)
PREPROCESS_BYTES_PER_STEP = 250_992  # 2*100800 read + 28224 write + 21168 carried


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled DURING the timed region: NVML every 2 ms (a v0 timed
    region is ~50 ms, shorter than one nvidia-smi call), nvidia-smi when NVML cannot be opened."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    # nvmlClocksEventReason bits, in the order of Q's reason columns
    BITS = (0x8, 0x40, 0x20, 0x4)

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.source = index, [], False, "nvidia-smi"
        self.nvml_error = None

    def _nvml_loop(self):
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
        mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
        while not self.stop_flag:
            sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            self.samples.append([str(sm), str(mx)] +
                                ["Active" if r & b else "Not Active" for b in self.BITS])
            time.sleep(0.002)

    def _smi_once(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                              "--format=csv,noheader,nounits"], capture_output=True,
                             text=True, timeout=5).stdout.strip()
        if out:
            self.samples.append([x.strip() for x in out.split(",")])

    def run(self):
        try:
            self.source = "nvml"
            self._nvml_loop()
            if self.samples:
                return
        except Exception as e:  # reported in the bench line so the cause is visible next run
            self.nvml_error = repr(e)[:120]
        self.source = "nvidia-smi"
        self.samples = []
        while True:  # at least one sample even if the region already ended
            try:
                self._smi_once()
            except Exception:
                pass
            if self.stop_flag:
                break
            time.sleep(0.2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names)
                   if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]),
                "reasons": reasons, "samples": len(self.samples), "source": self.source,
                **({"nvml_error": self.nvml_error} if self.nvml_error else {})}


def synth_inputs(w, seed):
    import synth
    E, T = w["E"], w["T"]
    s = synth.rollout_scalars(E, T, seed=seed)
    # a pool of distinct ALE-shaped + uniform frame pairs, tiled to [T,E,2,210,160,3]
    pool = np.concatenate([synth.frames(48, seed=seed, kind="uniform"),
                           synth.frames(16, seed=seed + 1, kind="ale")])
    return s, pool


def cpu_reference_iteration(w, threads, sample_frac=1.0, iters=1, warm=1):
    """The reference's own CPU code (oracle/_ref) on the same workload shape.  Returns
    (steps/s, description).  For big shapes a bounded sample: E_s = E*sample_frac envs."""
    from oracle import ref
    if not ref.available():
        return None, "oracle/_ref missing"
    import synth
    ref.set_num_threads(threads)
    E = max(w["nmb"], int(round(w["E"] * sample_frac)))
    E -= E % w["nmb"] if (E * w["T"]) % w["nmb"] else 0
    T = w["T"]
    s = synth.rollout_scalars(E, T, seed=42)
    pool = synth.frames(64, seed=42, kind="uniform")
    L = ref.Learner(w["H"], 4, 42, HP["lr"])
    times = []
    for i in range(warm + iters):
        secs, _ = L.iteration(pool, s["actions"], s["rewards"], s["terminals"], s["truncations"],
                              s["episode_starts"], s["logits"], s["values"], s["next_values"],
                              w["epochs"], w["nmb"], HP["gamma"], HP["lam"], HP["clip"],
                              HP["vcoef"], HP["ecoef"], HP["max_norm"])
        if i >= warm:
            times.append(sum(secs))
    t = float(np.median(times))
    return E * T / t, (f"{iters} full iteration(s) of {E} envs x {T} steps "
                       f"({w['epochs']} epochs x {w['nmb']} minibatches), median; "
                       f"stages preprocess/get/train = {secs[0]:.3f}/{secs[1]:.3f}/{secs[2]:.3f} s")


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    frac = 1.0 if w["E"] * w["T"] <= 4096 else 512.0 / (w["E"] * w["T"]) * 8
    t0 = time.time()
    vals = []
    desc = ""
    for _ in range(max(1, args.warmup if args.warmup < 2 else 1)):
        cpu_reference_iteration(w, threads, frac, iters=1, warm=0)
    for _ in range(args.steps):
        v, desc = cpu_reference_iteration(w, threads, frac, iters=1, warm=0)
        if v is None:
            print(json.dumps({"impl": "reference", "unavailable": desc}))
            return
        vals.append(v)
        if time.time() - t0 > 240:
            break
    v = float(np.median(vals))
    n_steps = len(vals)
    E = max(w["nmb"], int(round(w["E"] * frac)))
    line = {"metric": METRIC, "value": v, "unit": "agent-steps/s", "n_gpus": args.gpus,
            "steps": n_steps, "warmup": 1, "ms_per_step": 1e3 * E * w["T"] / v,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference",
            # same keys as our arm's config; sample_envs = the bounded sample one step runs
            "config": {"workload": args.workload, "envs_per_gpu": w["E"], "horizon": w["T"],
                       "hidden_size": w["H"], "num_epochs": w["epochs"], "num_mini_batches": w["nmb"],
                       "mini_batch_per_gpu": w["E"] * w["T"] // w["nmb"], "math": "fp32 (libtorch CPU)",
                       "sample_envs": E},
            "cpu_baseline": {"value": v, "unit": "agent-steps/s", "cores": threads,
                             "kind": "reference", "sample": desc},
            "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="v0", choices=list(WORKLOADS))
    ap.add_argument("--math", default=os.environ.get("ALEGPU_MATH", "bf16"), choices=["fp32", "bf16", "tf32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.impl == "reference":
        return run_reference(args, w)

    import alegpu
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    uid = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            raw = C.create_string_buffer(128)
            alegpu._check(alegpu.lib().alegpu_nccl_unique_id(raw))
            buf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        uid = bytes(buf.cpu().numpy().tobytes())
    alegpu._check(alegpu.lib().alegpu_set_device(local))
    DA = alegpu.DeviceArray
    E, T, H, epochs, nmb, A = w["E"], w["T"], w["H"], w["epochs"], w["nmb"], 4
    N = E * T
    mb = N // nmb
    math = {"bf16": alegpu.MATH_BF16, "tf32": alegpu.MATH_TF32, "fp32": alegpu.MATH_FP32}[args.math]

    # ---- synthetic inputs (seeded; different env slice per rank) ------------------------------
    s, pool = synth_inputs(w, seed=42 + rank)
    npool = pool.shape[0]
    idx = (np.arange(T * E) * 7 + 3) % npool
    slab = E * 2 * alegpu.FRAME_BYTES
    hptr = C.c_void_p()
    alegpu._check(alegpu.lib().alegpu_host_alloc(C.byref(hptr), C.c_size_t(T * slab)))
    frames_host = np.ctypeslib.as_array(C.cast(hptr, C.POINTER(C.c_uint8)), shape=(T * slab,))
    fh = frames_host.reshape(T * E, 2, 210, 160, 3)
    for lo in range(0, T * E, 4096):
        hi = min(T * E, lo + 4096)
        fh[lo:hi] = pool[idx[lo:hi]]
    frames_dev = DA((T * slab,), np.uint8)
    alegpu._check(alegpu.lib().alegpu_memcpy_h2d(C.c_void_p(frames_dev.ptr), hptr,
                                                 C.c_size_t(T * slab), None))
    alegpu.synchronize()

    import synth
    params = synth.orthogonal_params(H, A, seed=42)
    L = alegpu.Learner(H, A, mb, math, HP["lr"], device=local, world_size=world, rank=rank,
                       nccl_unique_id=uid)
    L.set_params(params)
    stream = L.stream
    buf = alegpu.Buffer(E, T, A, device=local)
    pristine = {k: DA.from_numpy(s[k]) for k in
                ("actions", "rewards", "values", "logits", "terminals", "truncations",
                 "episode_starts")}
    nv = DA.from_numpy(s["next_values"])
    met = alegpu.Metrics(epochs, nmb, mb)
    hp = alegpu.Hyperparameters(HP["clip"], HP["vcoef"], HP["ecoef"], HP["max_norm"])
    lib = alegpu.lib()
    carry0 = buf.carry.view(0, (E, 4, 84, 84))
    carry1 = buf.carry.view(E * alegpu.STACK_BYTES, (E, 4, 84, 84))
    use_graph = not args.no_graph

    def restore_scalars_d2d():
        for k, d in pristine.items():
            alegpu._check(lib.alegpu_memcpy_d2d(C.c_void_p(getattr(buf, k).ptr), C.c_void_p(d.ptr),
                                                C.c_size_t(d.nbytes), C.c_void_p(stream)))

    def evt():
        e = C.c_void_p()
        alegpu._check(lib.alegpu_event_create(C.byref(e)))
        return e

    ev = [evt() for _ in range(6)]
    pre_ms, train_ms = [], []

    def device_step(i, timed):
        restore_scalars_d2d()
        cin, cout = (carry0, carry1) if i % 2 == 0 else (carry1, carry0)
        if timed:
            alegpu._check(lib.alegpu_event_record(ev[2], C.c_void_p(stream)))
        alegpu.preprocess_rollout(frames_dev, buf.episode_starts, cin, cout, buf.observations, E, T,
                                  stream=stream)
        if timed:
            alegpu._check(lib.alegpu_event_record(ev[3], C.c_void_p(stream)))
        b = alegpu._Batch()
        alegpu._check(lib.alegpu_rollout_get(buf.h, C.c_void_p(nv.ptr), C.c_float(HP["gamma"]),
                                             C.c_float(HP["lam"]), C.byref(b), 1,
                                             C.c_void_p(stream)))
        cm = met._c()
        if timed:
            alegpu._check(lib.alegpu_event_record(ev[4], C.c_void_p(stream)))
        alegpu._check(lib.alegpu_learner_train(L.h, C.byref(b), C.byref(cm), C.c_int64(epochs),
                                               C.c_int64(nmb), C.byref(hp), 1 if use_graph else 0))
        if timed:
            alegpu._check(lib.alegpu_event_record(ev[5], C.c_void_p(stream)))
            alegpu.synchronize(stream)
            ms = C.c_float(0)
            alegpu._check(lib.alegpu_event_elapsed_ms(ev[2], ev[3], C.byref(ms)))
            pre_ms.append(ms.value)
            alegpu._check(lib.alegpu_event_elapsed_ms(ev[4], ev[5], C.byref(ms)))
            train_ms.append(ms.value)

    def barrier():
        alegpu.synchronize(stream)
        if dist is not None:
            dist.barrier()
            import torch
            torch.cuda.synchronize()

    def timed_loop(fn, steps):
        barrier()
        alegpu._check(lib.alegpu_event_record(ev[0], C.c_void_p(stream)))
        for i in range(steps):
            fn(i)
        alegpu._check(lib.alegpu_event_record(ev[1], C.c_void_p(stream)))
        alegpu.synchronize(stream)
        ms = C.c_float(0)
        alegpu._check(lib.alegpu_event_elapsed_ms(ev[0], ev[1], C.byref(ms)))
        t = ms.value
        if dist is not None:
            import torch
            tt = torch.tensor([t], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        barrier()
        return t

    # ---- value: device-resident inputs ---------------------------------------------------------
    for i in range(args.warmup):
        device_step(i, False)
    alegpu.synchronize(stream)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = alegpu.launch_count()
    total_ms = timed_loop(lambda i: device_step(i, False), args.steps)
    launches = alegpu.launch_count() - l0
    # the dominant memory-bound kernel, timed live with events on the launching stream
    for i in range(min(args.steps, 10)):
        device_step(i, True)
    sampler.stop_flag = True
    ms_per_step = total_ms / args.steps
    value = world * N / (ms_per_step * 1e-3)

    # ---- e2e: host buffers through the reference-facing API ---------------------------------------
    # Every step copies its inputs from PINNED host memory (raw frames through the ring of
    # alegpu_rollout_ingest, scalar streams, next_values) and reads its results (loss and pre-clip
    # gradient norms of every minibatch) back into pinned host memory.  The host runs one step
    # ahead of the device: step i is enqueued before the results of step i-1 are awaited, so the
    # frame copies of step i (copy stream) overlap train() of step i-1; every step's results are
    # read inside the timed region (the loop ends on a full synchronize).
    def pinned_like(a):
        p = C.c_void_p()
        alegpu._check(lib.alegpu_host_alloc(C.byref(p), C.c_size_t(max(a.nbytes, 16))))
        v = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(a.nbytes,))
        v = v.view(a.dtype).reshape(a.shape)
        v[...] = a
        return v

    host_scalars = {k: pinned_like(s[k]) for k in pristine}
    nv_host = pinned_like(s["next_values"])
    loss_host = [pinned_like(np.zeros((epochs, nmb, mb), np.float32)) for _ in range(2)]
    gn_host = [pinned_like(np.zeros((epochs, nmb), np.float32)) for _ in range(2)]
    done = [evt(), evt()]
    pending = [False, False]

    def e2e_step(i):
        for k, a in host_scalars.items():
            d = getattr(buf, k)
            alegpu._check(lib.alegpu_memcpy_h2d(C.c_void_p(d.ptr), a.ctypes.data_as(C.c_void_p),
                                                C.c_size_t(a.nbytes), C.c_void_p(stream)))
        alegpu._check(lib.alegpu_memcpy_h2d(C.c_void_p(nv.ptr), nv_host.ctypes.data_as(C.c_void_p),
                                            C.c_size_t(nv_host.nbytes), C.c_void_p(stream)))
        alegpu._check(lib.alegpu_rollout_ingest(buf.h, hptr, C.c_int(i % 2), C.c_void_p(stream)))
        b = alegpu._Batch()
        alegpu._check(lib.alegpu_rollout_get(buf.h, C.c_void_p(nv.ptr), C.c_float(HP["gamma"]),
                                             C.c_float(HP["lam"]), C.byref(b), 1,
                                             C.c_void_p(stream)))
        cm = met._c()
        alegpu._check(lib.alegpu_learner_train(L.h, C.byref(b), C.byref(cm), C.c_int64(epochs),
                                               C.c_int64(nmb), C.byref(hp), 1 if use_graph else 0))
        j = i % 2
        if pending[j]:  # results of step i-2 were consumed below before this buffer is reused
            pending[j] = False
        alegpu._check(lib.alegpu_memcpy_d2h(loss_host[j].ctypes.data_as(C.c_void_p),
                                            C.c_void_p(met.loss.ptr),
                                            C.c_size_t(loss_host[j].nbytes), C.c_void_p(stream)))
        alegpu._check(lib.alegpu_memcpy_d2h(gn_host[j].ctypes.data_as(C.c_void_p),
                                            C.c_void_p(met.clipped_gradients.ptr),
                                            C.c_size_t(gn_host[j].nbytes), C.c_void_p(stream)))
        alegpu._check(lib.alegpu_event_record(done[j], C.c_void_p(stream)))
        pending[j] = True
        if pending[1 - j]:  # await and read the PREVIOUS step's results
            alegpu._check(lib.alegpu_event_synchronize(done[1 - j]))
            assert np.isfinite(loss_host[1 - j][0, 0, 0]) and np.isfinite(gn_host[1 - j]).all()
            pending[1 - j] = False

    e2e_steps = max(3, min(args.steps, 10))
    for i in range(2):
        e2e_step(i)
    e2e_ms = timed_loop(e2e_step, e2e_steps) / e2e_steps
    e2e_value = world * N / (e2e_ms * 1e-3)
    h2d = T * slab + sum(a.nbytes for a in host_scalars.values()) + s["next_values"].nbytes
    d2h = loss_host[0].nbytes + gn_host[0].nbytes
    for j in range(2):
        assert np.isfinite(loss_host[j]).all() and np.isfinite(gn_host[j]).all(), "non-finite metrics"

    # Multi-rank exit: every collective of this run has completed (timed_loop ends on a barrier).
    # Graceful NCCL teardown (ncclCommDestroy / destroy_process_group) of two communicators in one
    # process hung on this image, so ranks leave through os._exit once rank 0 has printed its line.
    if rank != 0:
        sys.stdout.flush()
        os._exit(0)
    hbm, tflops, which = peaks()
    pre = float(np.median(pre_ms)) if pre_ms else None
    roof = None
    traffic = None
    try:  # measured DRAM bytes per launch of this kernel (one ncu --set full capture, see profiles/)
        with open(os.path.join(ROOT, "profiles", "r01_roofline_traffic.json")) as f:
            traffic = json.load(f).get(args.workload)
    except Exception:
        pass
    if pre:
        ach = N * PREPROCESS_BYTES_PER_STEP / (pre * 1e-3) / 1e9
        roof = {"kernel": "preprocess_rollout_pipe_kernel", "bound": "hbm", "achieved": ach, "peak": hbm,
                "unit": "GB/s", "frac": ach / hbm, "traffic": traffic, "peak_source": which,
                "ms_per_launch": pre, "algorithmic_bytes_per_launch": N * PREPROCESS_BYTES_PER_STEP}
    # the tensor-core side of the step, for the same honesty: train() = epochs x N samples x
    # 49.52 MFLOP (18.69 fwd + 30.83 bwd at H=512; 40.47 at H=32) against the measured
    # dense bf16 peak.  At these minibatch sizes the chain of small GEMMs is latency / L2-bound, far
    # from the tensor roof; the number is reported so that nobody has to guess it.
    roof_tc = None
    if train_ms:
        tr = float(np.median(train_ms))
        flop = epochs * N * (49.52e6 if H == 512 else 40.47e6 if H == 32 else 0.0)
        if flop:
            ach = flop / (tr * 1e-3) / 1e12
            roof_tc = {"kernel": "train(): 11 gemm_tc_kernel launches per minibatch step", "bound": "tensor",
                       "achieved": ach, "peak": tflops, "unit": "TFLOP/s", "frac": ach / tflops,
                       "ms_per_train": tr, "peak_source": which}
    line = {"metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"bf16": "bf16", "tf32": "tf32", "fp32": "f32"}[args.math], "data": "synthetic",
            "config": {"workload": args.workload, "envs_per_gpu": E, "horizon": T, "hidden_size": H,
                       "num_epochs": epochs, "num_mini_batches": nmb, "mini_batch_per_gpu": mb,
                       "cuda_graph": use_graph, "math": args.math,
                       "l2": "inputs larger than L2 (raw frames %.0f MB per step)" % (T * slab / 1e6)},
            "e2e": {"value": e2e_value, "unit": "agent-steps/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms},
            "gpu_launches": int(launches), "clocks": sampler.summary(), "roofline": roof,
            "roofline_tensor": roof_tc}
    if not args.no_cpu_baseline and world == 1:  # reported on rank 0 at N=1 only
        threads = os.cpu_count() or 1
        frac = 1.0 if N <= 4096 else 4096.0 / N
        v, desc = cpu_reference_iteration(w, threads, frac, iters=2 if N <= 1024 else 1, warm=1)
        if v is not None:
            line["cpu_baseline"] = {"value": v, "unit": "agent-steps/s", "cores": threads,
                                    "kind": "reference", "sample": desc}
        else:
            line["cpu_baseline"] = {"value": None, "unit": "agent-steps/s", "cores": threads,
                                    "kind": "reference", "sample": desc}
    print(json.dumps(line))
    sys.stdout.flush()
    if world > 1:
        os._exit(0)  # never linger in interpreter teardown with NCCL state around


if __name__ == "__main__":
    try:
        main()
    except BaseException:
        import traceback
        traceback.print_exc()
        sys.stdout.flush()
        sys.stderr.flush()
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            os._exit(1)  # do not hang in teardown with NCCL communicators alive
        raise
