#!/usr/bin/env python
"""bench.py — agent-steps/sec through preprocess + GAE + PPO update (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload v0|v1|debug|sweep]
                    [--math tf32|bf16|fp32] [--impl ours|reference]

A "step" is one learner iteration over one rollout batch of synthetic input:
  preprocess (2-frame max + area resize + gray + 4-frame stack) of E*T frame pairs
  -> Buffer::get (reward clamp + GAE + returns + masks) + prepare_batch (log-softmax)
  -> train(): num_epochs x num_mini_batches of {Nature-CNN fwd/bwd, PPO loss, clip, Adam}.
Workloads are the reference's configs/*.yaml (shape AND hyper-parameters), read through
alegpu_config_load.  Default: N = 1 -> configs/v1.yaml (the largest single-GPU configuration of
BASELINE.json), N > 1 -> the 4096 x 128 sweep of BASELINE config 4 in weak-scaling form (512 envs
per GPU; N = 8 is the whole config).  Default math: tf32 (tcgen05 kind::tf32 — the precision cuDNN
gives the reference's fp32 network on a GPU); the bf16 line rides along as an extra key.
`value`  : inputs (raw frames [T,E,2,210,160,3] u8, reward/flag/logit/value streams) already
           resident in HBM when the timed region starts; CUDA events on the launching stream; the K
           steps are repeated until the timed region is >= 2 s.
`e2e`    : the same iteration through the reference-facing API with HOST buffers: the pinned
           frame ring is cudaMemcpyAsync'ed on a side stream (alegpu_rollout_ingest), scalar streams
           are H2D-copied, and the per-minibatch loss / grad-norm metrics are read back D2H, all
           inside the timed region.
`--impl reference` times the reference's own CPU implementation (oracle/_ref: the reference
sources compiled unmodified against libtorch-CPU) with all host threads on the same workload.
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
WORKLOAD_NAMES = ("debug", "v0", "v1", "sweep")
SWEEP_ENVS_PER_GPU = 512  # BASELINE config 4: 4096 envs env-sharded over 8 GPUs
# 2*100800 read + 28224 write + 21168 carried planes; the rollout kernel re-derives the
# carried planes from the frames instead of reading them, so it really moves 229 824 B per agent step
PREPROCESS_ALGO_BYTES = 250_992
PREPROCESS_MOVED_BYTES = 229_824
TRAIN_MFLOP_PER_SAMPLE = {512: 49.52, 32: 40.47}  # fwd 18.69 + bwd 30.83 at H=512
MIN_TIMED_S = 2.0


def parse_yaml_cpu(path):
    """the flat `key: value` schema without the CUDA library (the reference arm never loads it)"""
    out = {}
    for line in open(path):
        line = line.split(" #")[0].strip()
        if not line or line.startswith("#") or ":" not in line:
            continue
        k, v = (x.strip() for x in line.split(":", 1))
        try:
            out[k] = int(v)
        except ValueError:
            try:
                out[k] = float(v)
            except ValueError:
                out[k] = v.lower() in ("true", "yes", "on")
    return out


def load_workload(name, gpus, use_lib=True):
    """configs/<name>.yaml -> the shape / hyper-parameter dict one rank runs"""
    path = os.path.join(PKG, "configs", name + ".yaml")
    if use_lib:
        import alegpu
        c = alegpu.Config.load(path).as_dict()
    else:
        c = parse_yaml_cpu(path)
    E = c["total_environments"]
    if name == "sweep":  # weak-scaling form: every GPU holds the slice it holds in the 8-GPU config
        E = SWEEP_ENVS_PER_GPU
    return dict(name=name, E=E, T=c["horizon"], H=c["hidden_size"], epochs=c["num_epochs"],
                nmb=c["num_mini_batches"], clip=c["clip_param"], vcoef=c["value_loss_coef"],
                ecoef=c["entropy_coef"], max_norm=c["max_gradient_norm"], lr=c["learning_rate"],
                gamma=c["gae_discount"], lam=c["gae_lambda"], config_file=f"configs/{name}.yaml",
                global_envs=E * gpus)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


# ---- clocks: sampled by a SEPARATE process ------------------------------------------------------------
# Round 1 polled NVML from a thread of this process every 2 ms; that thread fought the launch thread for
# the GIL / a core inside a timed region that enqueues ~10 host calls per step (BENCH_r01 3.79 ms vs
# SCALE_r01 2.32 ms per v0 step on the same command; A/B in profiles/r02_sampler_ab.txt).  The sampler
# is now its own process (`bench.py --_sampler`): it appends "t sm max hw_slowdown hw_thermal
# sw_thermal sw_power_cap" lines until it is terminated, and the bench keeps the samples whose
# timestamps fall inside its timed regions.
SAMPLER_BITS = (0x8, 0x40, 0x20, 0x4)  # nvmlClocksEventReason: hw_slowdown, hw_thermal, sw_thermal, sw_power_cap
SAMPLER_Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")


def sampler_main(index, path, period=0.02):
    with open(path, "w", buffering=1) as out:
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            out.write("# nvml\n")
            while True:
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                out.write(f"{time.time():.4f} {sm} {mx} " +
                          " ".join("1" if r & b else "0" for b in SAMPLER_BITS) + "\n")
                time.sleep(period)
        except Exception as e:  # NVML not usable: nvidia-smi, one sample per call
            out.write(f"# nvidia-smi ({repr(e)[:80]})\n")
            while True:
                try:
                    o = subprocess.run(["nvidia-smi", f"--id={index}", f"--query-gpu={SAMPLER_Q}",
                                        "--format=csv,noheader,nounits"], capture_output=True,
                                       text=True, timeout=5).stdout.strip().split(",")
                    out.write(f"{time.time():.4f} {float(o[0]):.0f} {float(o[1]):.0f} " + " ".join(
                        "1" if x.strip().lower().startswith("active") else "0" for x in o[2:6]) + "\n")
                except Exception:
                    pass
                time.sleep(0.1)


class ClockSampler:
    def __init__(self, index, mode="process"):
        self.mode, self.index, self.windows, self.proc = mode, index, [], None
        self.path = f"/tmp/alegpu_clocks_{os.getpid()}_{index}.txt"
        if mode == "process":
            self.proc = subprocess.Popen([sys.executable, os.path.abspath(__file__), "--_sampler",
                                          str(index), self.path], stdout=subprocess.DEVNULL,
                                         stderr=subprocess.DEVNULL)
        elif mode == "thread":  # round 1's in-process sampler, kept for the A/B only
            threading.Thread(target=sampler_main, args=(index, self.path, 0.002), daemon=True).start()

    def window(self, t0, t1):
        self.windows.append((t0, t1))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()  # the exact process started above
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        rows, source = [], "none"
        try:
            for line in open(self.path):
                if line.startswith("#"):
                    source = line[1:].strip()
                    continue
                f = line.split()
                if len(f) == 7:
                    rows.append([float(f[0]), float(f[1]), float(f[2])] + [int(x) for x in f[3:]])
        except Exception:
            pass
        load = [r for r in rows if any(a <= r[0] <= b for a, b in self.windows)]
        if not load:
            return {"sm_mhz": None, "sm_max_mhz": rows[0][2] if rows else None,
                    "reasons": ["unsampled"], "samples": 0, "source": source, "sampler": self.mode}
        sm = sorted(r[1] for r in load)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": load[0][2],
                "reasons": [n for i, n in enumerate(names) if any(r[3 + i] for r in load)],
                "samples": len(load), "source": source,
                "sampler": "separate process" if self.mode == "process" else self.mode}


# ---- the reference's own CPU code (oracle/_ref) on the same workload ----------------------------------
def synth_inputs(w, seed):
    import synth
    s = synth.rollout_scalars(w["E"], w["T"], seed=seed)
    # a pool of distinct ALE-shaped + uniform frame pairs, tiled to [T,E,2,210,160,3]
    pool = np.concatenate([synth.frames(48, seed=seed, kind="uniform"),
                           synth.frames(16, seed=seed + 1, kind="ale")])
    return s, pool


def cpu_sample_envs(w, budget_steps):
    """a bounded sample of the workload: as many whole envs as keep one iteration <= budget_steps agent
    steps, at least one env per minibatch so the epochs x minibatches structure is the workload's"""
    E = min(w["global_envs"], max(1, budget_steps // w["T"]))
    while (E * w["T"]) % w["nmb"]:
        E += 1
    return E


def cpu_reference_iteration(w, threads, sample_envs, iters=1, warm=1):
    """The reference's own CPU code (oracle/_ref) on `sample_envs` envs of the workload.  Returns
    (agent-steps/s, description)."""
    from oracle import ref
    if not ref.available():
        return None, "oracle/_ref missing"
    import synth
    ref.set_num_threads(threads)
    E, T = sample_envs, w["T"]
    s = synth.rollout_scalars(E, T, seed=42)
    pool = synth.frames(64, seed=42, kind="uniform")
    L = ref.Learner(w["H"], 4, 42, w["lr"])
    times, secs = [], (0, 0, 0)
    for i in range(warm + iters):
        secs, _ = L.iteration(pool, s["actions"], s["rewards"], s["terminals"], s["truncations"],
                              s["episode_starts"], s["logits"], s["values"], s["next_values"],
                              w["epochs"], w["nmb"], w["gamma"], w["lam"], w["clip"], w["vcoef"],
                              w["ecoef"], w["max_norm"])
        if i >= warm:
            times.append(sum(secs))
    t = float(np.median(times))
    return E * T / t, (f"{iters} full iteration(s) of {E} of the workload's {w['global_envs']} envs x {T} "
                       f"steps ({w['epochs']} epochs x {w['nmb']} minibatches of {E * T // w['nmb']} rows), "
                       f"median; stages preprocess/get/train = {secs[0]:.3f}/{secs[1]:.3f}/{secs[2]:.3f} s")


def shape_config(w, world):
    """the config keys both arms print (same_config is judged on these)"""
    return {"workload": w["name"], "config_file": w["config_file"], "envs_per_gpu": w["E"],
            "global_envs": w["global_envs"], "horizon": w["T"], "hidden_size": w["H"],
            "num_epochs": w["epochs"], "num_mini_batches": w["nmb"],
            "mini_batch_per_gpu": w["E"] * w["T"] // w["nmb"],
            "global_mini_batch": w["global_envs"] * w["T"] // w["nmb"],
            "clip_param": round(w["clip"], 6), "value_loss_coef": round(w["vcoef"], 6)}


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    w = load_workload(args.workload, args.gpus, use_lib=False)
    threads = os.cpu_count() or 1
    E = cpu_sample_envs(w, 4096)  # ~1-2 s of CPU work per step
    t0, vals, desc = time.time(), [], ""
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_reference_iteration(w, threads, E, iters=1, warm=0)
    for _ in range(args.steps):
        v, desc = cpu_reference_iteration(w, threads, E, iters=1, warm=0)
        if v is None:
            print(json.dumps({"impl": "reference", "unavailable": desc}))
            return
        vals.append(v)
        if time.time() - t0 > 200:
            break
    v = float(np.median(vals))
    cfg = shape_config(w, args.gpus)
    cfg.update({"math": "fp32 (libtorch CPU)", "sample_envs": E})
    line = {"metric": METRIC, "value": v, "unit": "agent-steps/s", "n_gpus": args.gpus,
            "steps": len(vals), "warmup": 1, "ms_per_step": 1e3 * E * w["T"] / v,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference", "config": cfg,
            "cpu_baseline": {"value": v, "unit": "agent-steps/s", "cores": threads,
                             "kind": "reference", "sample": desc},
            "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ---- the product arm ------------------------------------------------------------------------------------
class Run:
    """One workload x math mode on this rank's GPU: buffers, learner, the device-resident step and the
    host-buffer (e2e) step."""

    def __init__(self, w, math_name, local, world, rank, uid, use_graph, seed_rank):
        import alegpu
        import synth
        self.alegpu, self.w, self.math_name, self.world = alegpu, w, math_name, world
        self.lib = lib = alegpu.lib()
        DA = alegpu.DeviceArray
        E, T, H, A = w["E"], w["T"], w["H"], 4
        self.N, self.mb = E * T, E * T // w["nmb"]
        self.use_graph = use_graph
        math = {"bf16": alegpu.MATH_BF16, "tf32": alegpu.MATH_TF32, "fp32": alegpu.MATH_FP32}[math_name]
        # synthetic inputs (seeded; a different env slice per rank)
        s, pool = synth_inputs(w, seed=42 + seed_rank)
        self.s = s
        idx = (np.arange(T * E) * 7 + 3) % pool.shape[0]
        self.slab = slab = E * 2 * alegpu.FRAME_BYTES
        self.hptr = C.c_void_p()
        alegpu._check(lib.alegpu_host_alloc(C.byref(self.hptr), C.c_size_t(T * slab)))
        fh = np.ctypeslib.as_array(C.cast(self.hptr, C.POINTER(C.c_uint8)), shape=(T * slab,))
        fh = fh.reshape(T * E, 2, 210, 160, 3)
        for lo in range(0, T * E, 4096):
            hi = min(T * E, lo + 4096)
            fh[lo:hi] = pool[idx[lo:hi]]
        self.frames_dev = DA((T * slab,), np.uint8)
        alegpu._check(lib.alegpu_memcpy_h2d(C.c_void_p(self.frames_dev.ptr), self.hptr,
                                            C.c_size_t(T * slab), None))
        alegpu.synchronize()
        self.L = alegpu.Learner(H, A, self.mb, math, w["lr"], device=local, world_size=world, rank=rank,
                                nccl_unique_id=uid)
        self.L.set_params(synth.orthogonal_params(H, A, seed=42))
        self.stream = self.L.stream
        self.buf = alegpu.Buffer(E, T, A, device=local)
        self.pristine = {k: DA.from_numpy(s[k]) for k in
                         ("actions", "rewards", "values", "logits", "terminals", "truncations",
                          "episode_starts")}
        self.nv = DA.from_numpy(s["next_values"])
        self.met = alegpu.Metrics(w["epochs"], w["nmb"], self.mb)
        self.hp = alegpu.Hyperparameters(w["clip"], w["vcoef"], w["ecoef"], w["max_norm"])
        self.carry = (self.buf.carry.view(0, (E, 4, 84, 84)),
                      self.buf.carry.view(E * alegpu.STACK_BYTES, (E, 4, 84, 84)))
        self.ev = [self.evt() for _ in range(6)]
        self.pre_ms, self.train_ms = [], []
        self.e2e_ready = False

    def evt(self):
        e = C.c_void_p()
        self.alegpu._check(self.lib.alegpu_event_create(C.byref(e)))
        return e

    def close(self):
        self.L.close()
        self.alegpu._check(self.lib.alegpu_host_free(self.hptr))
        for a in getattr(self, "pinned", []):
            self.lib.alegpu_host_free(a)
        self.buf = self.frames_dev = self.pristine = self.met = None

    # one learner iteration with device-resident inputs
    def device_step(self, i, timed=False):
        a, lib, w, st = self.alegpu, self.lib, self.w, C.c_void_p(self.stream)
        for k, d in self.pristine.items():
            a._check(lib.alegpu_memcpy_d2d(C.c_void_p(getattr(self.buf, k).ptr), C.c_void_p(d.ptr),
                                           C.c_size_t(d.nbytes), st))
        cin, cout = (self.carry[0], self.carry[1]) if i % 2 == 0 else (self.carry[1], self.carry[0])
        if timed:
            a._check(lib.alegpu_event_record(self.ev[2], st))
        a.preprocess_rollout(self.frames_dev, self.buf.episode_starts, cin, cout, self.buf.observations,
                             w["E"], w["T"], stream=self.stream)
        if timed:
            a._check(lib.alegpu_event_record(self.ev[3], st))
        b = a._Batch()
        a._check(lib.alegpu_rollout_get(self.buf.h, C.c_void_p(self.nv.ptr), C.c_float(w["gamma"]),
                                        C.c_float(w["lam"]), C.byref(b), 1, st))
        cm = self.met._c()
        if timed:
            a._check(lib.alegpu_event_record(self.ev[4], st))
        a._check(lib.alegpu_learner_train(self.L.h, C.byref(b), C.byref(cm), C.c_int64(w["epochs"]),
                                          C.c_int64(w["nmb"]), C.byref(self.hp),
                                          1 if self.use_graph else 0))
        if timed:
            a._check(lib.alegpu_event_record(self.ev[5], st))
            a.synchronize(self.stream)
            ms = C.c_float(0)
            a._check(lib.alegpu_event_elapsed_ms(self.ev[2], self.ev[3], C.byref(ms)))
            self.pre_ms.append(ms.value)
            a._check(lib.alegpu_event_elapsed_ms(self.ev[4], self.ev[5], C.byref(ms)))
            self.train_ms.append(ms.value)

    # ---- e2e: host buffers through the reference-facing API -------------------------------------------
    # Every step copies its inputs from PINNED host memory (raw frames through the ring of
    # alegpu_rollout_ingest, scalar streams, next_values) and reads its results (loss and pre-clip
    # gradient norm of every minibatch) back into pinned host memory.  The host runs one step ahead of
    # the device: step i is enqueued before the results of step i-1 are awaited, so the frame copies of
    # step i (copy stream) overlap train() of step i-1; every step's results are read inside the timed
    # region (the loop ends on a full synchronize).
    def pinned_like(self, arr):
        p = C.c_void_p()
        self.alegpu._check(self.lib.alegpu_host_alloc(C.byref(p), C.c_size_t(max(arr.nbytes, 16))))
        self.pinned.append(p)
        v = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(arr.nbytes,))
        v = v.view(arr.dtype).reshape(arr.shape)
        v[...] = arr
        return v

    def e2e_setup(self):
        w = self.w
        self.pinned = []
        self.host_scalars = {k: self.pinned_like(self.s[k]) for k in self.pristine}
        self.nv_host = self.pinned_like(self.s["next_values"])
        self.loss_host = [self.pinned_like(np.zeros((w["epochs"], w["nmb"], self.mb), np.float32))
                          for _ in range(2)]
        self.gn_host = [self.pinned_like(np.zeros((w["epochs"], w["nmb"]), np.float32)) for _ in range(2)]
        self.done = [self.evt(), self.evt()]
        self.pending = [False, False]
        self.h2d = (w["T"] * self.slab + sum(a.nbytes for a in self.host_scalars.values()) +
                    self.s["next_values"].nbytes)
        self.d2h = self.loss_host[0].nbytes + self.gn_host[0].nbytes
        self.e2e_ready = True

    def e2e_step(self, i):
        a, lib, w, st = self.alegpu, self.lib, self.w, C.c_void_p(self.stream)
        for k, h in self.host_scalars.items():
            d = getattr(self.buf, k)
            a._check(lib.alegpu_memcpy_h2d(C.c_void_p(d.ptr), h.ctypes.data_as(C.c_void_p),
                                           C.c_size_t(h.nbytes), st))
        a._check(lib.alegpu_memcpy_h2d(C.c_void_p(self.nv.ptr), self.nv_host.ctypes.data_as(C.c_void_p),
                                       C.c_size_t(self.nv_host.nbytes), st))
        a._check(lib.alegpu_rollout_ingest(self.buf.h, self.hptr, C.c_int(i % 2), st))
        b = a._Batch()
        a._check(lib.alegpu_rollout_get(self.buf.h, C.c_void_p(self.nv.ptr), C.c_float(w["gamma"]),
                                        C.c_float(w["lam"]), C.byref(b), 1, st))
        cm = self.met._c()
        a._check(lib.alegpu_learner_train(self.L.h, C.byref(b), C.byref(cm), C.c_int64(w["epochs"]),
                                          C.c_int64(w["nmb"]), C.byref(self.hp),
                                          1 if self.use_graph else 0))
        j = i % 2
        a._check(lib.alegpu_memcpy_d2h(self.loss_host[j].ctypes.data_as(C.c_void_p),
                                       C.c_void_p(self.met.loss.ptr),
                                       C.c_size_t(self.loss_host[j].nbytes), st))
        a._check(lib.alegpu_memcpy_d2h(self.gn_host[j].ctypes.data_as(C.c_void_p),
                                       C.c_void_p(self.met.clipped_gradients.ptr),
                                       C.c_size_t(self.gn_host[j].nbytes), st))
        a._check(lib.alegpu_event_record(self.done[j], st))
        self.pending[j] = True
        if self.pending[1 - j]:  # await and read the PREVIOUS step's results
            a._check(lib.alegpu_event_synchronize(self.done[1 - j]))
            assert np.isfinite(self.loss_host[1 - j][0, 0, 0]) and np.isfinite(self.gn_host[1 - j]).all()
            self.pending[1 - j] = False


class Harness:
    """barrier + CUDA-event timing on the learner stream, max over ranks"""

    def __init__(self, dist, sampler):
        self.dist, self.sampler = dist, sampler

    def barrier(self, run):
        run.alegpu.synchronize(run.stream)
        if self.dist is not None:
            import torch
            self.dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(self, t):
        if self.dist is None:
            return t
        import torch
        tt = torch.tensor([t], device="cuda", dtype=torch.float64)
        self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
        return float(tt.item())

    def timed_loop(self, run, fn, steps):
        """ms for `steps` calls of fn(i): events on the launching stream, barrier + synchronize on both
        sides, max over ranks"""
        a, lib = run.alegpu, run.lib
        self.barrier(run)
        t0 = time.time()
        a._check(lib.alegpu_event_record(run.ev[0], C.c_void_p(run.stream)))
        for i in range(steps):
            fn(i)
        a._check(lib.alegpu_event_record(run.ev[1], C.c_void_p(run.stream)))
        a.synchronize(run.stream)
        t1 = time.time()
        ms = C.c_float(0)
        a._check(lib.alegpu_event_elapsed_ms(run.ev[0], run.ev[1], C.byref(ms)))
        t = self.max_over_ranks(ms.value)
        self.barrier(run)
        if self.sampler is not None:
            self.sampler.window(t0, t1)
        return t

    def measure(self, run, steps, warmup, min_s=MIN_TIMED_S, e2e=True, stage_steps=10):
        """value (device-resident inputs), per-stage event times, e2e (host buffers).  The K steps are
        repeated R times inside ONE timed region so that it lasts >= min_s."""
        a = run.alegpu
        for i in range(warmup):
            run.device_step(i)
        pilot = self.timed_loop(run, run.device_step, steps)  # also warm-up; sizes the repeats
        reps = int(min(500, max(1, np.ceil(min_s * 1e3 / max(pilot, 1e-3)))))
        l0 = a.launch_count()
        total = self.timed_loop(run, run.device_step, steps * reps)
        launches = a.launch_count() - l0
        out = {"ms_per_step": total / (steps * reps), "repeats": reps, "timed_steps": steps * reps,
               "timed_region_s": total * 1e-3, "launches": int(launches)}
        out["value"] = run.world * run.N / (out["ms_per_step"] * 1e-3)
        for i in range(min(steps, stage_steps)):  # per-stage times, each step synchronised
            run.device_step(i, timed=True)
        out["pre_ms"] = float(np.median(run.pre_ms)) if run.pre_ms else None
        out["train_ms"] = float(np.median(run.train_ms)) if run.train_ms else None
        if e2e:
            run.e2e_setup()
            for i in range(2):
                run.e2e_step(i)
            pilot = self.timed_loop(run, run.e2e_step, max(2, min(steps, 4)))
            per = pilot / max(2, min(steps, 4))
            n = int(min(2000, max(steps, np.ceil(min_s * 1e3 / max(per, 1e-3)))))
            ms = self.timed_loop(run, run.e2e_step, n) / n
            for j in range(2):
                assert np.isfinite(run.loss_host[j]).all() and np.isfinite(run.gn_host[j]).all(), \
                    "non-finite metrics"
            out["e2e"] = {"value": run.world * run.N / (ms * 1e-3), "unit": "agent-steps/s",
                          "h2d_bytes_per_step": int(run.h2d), "d2h_bytes_per_step": int(run.d2h),
                          "ms_per_step": ms, "timed_steps": n,
                          "preprocessing_path": "RGB 2-frame max -> area resize -> gray (the reference's "
                                                "libtorch vision path, vision.cc:38-84): 201 600 B H2D per agent step"}
        return out


def rooflines(w, math_name, res):
    hbm, bf16_tf, which = peaks()
    N, H = w["E"] * w["T"], w["H"]
    roof = roof_hbm = None
    if res.get("train_ms") and H in TRAIN_MFLOP_PER_SAMPLE and math_name in ("tf32", "bf16"):
        flop = w["epochs"] * N * TRAIN_MFLOP_PER_SAMPLE[H] * 1e6
        ach = flop / (res["train_ms"] * 1e-3) / 1e12
        # tf32 dense tensor throughput is half the bf16 rate (B200_PROFILING.md: 1.1 vs 2.25 PFLOP/s);
        # MEASURED_PEAKS.json holds the measured bf16 figure
        peak = bf16_tf if math_name == "bf16" else bf16_tf / 2
        roof = {"kernel": "gemm_tc_kernel: the 11 implicit-GEMM launches of every minibatch step, summed "
                          "over one train() (the unit the CUDA events bracket)",
                "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                "traffic": None, "peak_source": which + (" bf16" if math_name == "bf16" else " bf16 / 2 (tf32)"),
                "ms_per_train": res["train_ms"], "algorithmic_flop_per_train": flop,
                "share_of_step": res["train_ms"] / res["ms_per_step"]}
    if res.get("pre_ms"):
        traffic = None
        for f in ("r02_roofline_traffic.json", "r01_roofline_traffic.json"):
            try:  # measured DRAM bytes per launch of this kernel (one ncu --set full capture, profiles/)
                with open(os.path.join(ROOT, "profiles", f)) as fh:
                    traffic = json.load(fh).get(w["name"])
                if traffic is not None:
                    break
            except Exception:
                pass
        ach = N * PREPROCESS_MOVED_BYTES / (res["pre_ms"] * 1e-3) / 1e9
        roof_hbm = {"kernel": "preprocess_rollout_pipe_kernel", "bound": "hbm", "achieved": ach, "peak": hbm,
                    "unit": "GB/s", "frac": ach / hbm, "traffic": traffic, "peak_source": which,
                    "ms_per_launch": res["pre_ms"], "bytes_per_launch": N * PREPROCESS_MOVED_BYTES,
                    "note": "229 824 B per agent step really moved (2 x 100 800 read + 28 224 written); "
                            "SURVEY's 250 992 B also counts the 21 168 B of carried planes the kernel re-derives",
                    "achieved_on_250992": N * PREPROCESS_ALGO_BYTES / (res["pre_ms"] * 1e-3) / 1e9,
                    "share_of_step": res["pre_ms"] / res["ms_per_step"]}
    return roof, roof_hbm


def brief(w, math_name, res):
    """an extra workload / math mode carried next to the headline"""
    roof, roof_hbm = rooflines(w, math_name, res)
    out = {"workload": w["name"], "math": math_name, "value": res["value"],
           "ms_per_step": res["ms_per_step"], "timed_steps": res["timed_steps"],
           "train_ms": res.get("train_ms"), "preprocess_ms": res.get("pre_ms")}
    if "e2e" in res:
        out["e2e"] = {k: res["e2e"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")}
    if roof:
        out["roofline_frac_tensor"] = roof["frac"]
    if roof_hbm:
        out["roofline_frac_hbm"] = roof_hbm["frac"]
    return out


def main():
    if len(sys.argv) >= 4 and sys.argv[1] == "--_sampler":
        return sampler_main(int(sys.argv[2]), sys.argv[3])
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOAD_NAMES),
                    help="default: v1 on one GPU, the sweep on several")
    ap.add_argument("--math", default=os.environ.get("ALEGPU_MATH", "tf32"), choices=["fp32", "bf16", "tf32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline workload / math only")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--min-seconds", type=float, default=MIN_TIMED_S)
    ap.add_argument("--sampler", default="process", choices=["process", "thread", "none"])
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.workload is None:
        args.workload = "v1" if max(world, args.gpus) == 1 else "sweep"
    if args.impl == "reference":
        return run_reference(args)

    import alegpu
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = uid = None
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
    w = load_workload(args.workload, world)
    use_graph = not args.no_graph
    sampler = ClockSampler(local, args.sampler) if rank == 0 and args.sampler != "none" else None
    H = Harness(dist, sampler)
    extras = {}

    run = Run(w, args.math, local, 1 if world > 1 else world, 0 if world > 1 else rank, None if world > 1 else uid,
              use_graph, seed_rank=rank)
    if world > 1:
        # the SAME per-GPU slice on one GPU with no exchange, measured in this run on every rank's own GPU:
        # the base the N-GPU value is to be read against (the driver's N = 1 run is configs/v1.yaml)
        solo = Harness(None, None).measure(run, args.steps, args.warmup, min_s=min(1.0, args.min_seconds),
                                           e2e=False, stage_steps=4)
        extras["single_gpu_same_slice"] = {"value": H.max_over_ranks(-solo["value"]) * -1.0,
                                           "ms_per_step": H.max_over_ranks(solo["ms_per_step"]),
                                           "note": "world_size 1 learner on this rank's 512-env slice, slowest rank"}
        run.L.close()
        run.pre_ms, run.train_ms = [], []
        run.world = world
        import synth
        run.L = alegpu.Learner(w["H"], 4, run.mb, {"bf16": alegpu.MATH_BF16, "tf32": alegpu.MATH_TF32,
                                                   "fp32": alegpu.MATH_FP32}[args.math], w["lr"],
                               device=local, world_size=world, rank=rank, nccl_unique_id=uid)
        run.L.set_params(synth.orthogonal_params(w["H"], 4, seed=42))
        run.stream = run.L.stream
    res = H.measure(run, args.steps, args.warmup, min_s=args.min_seconds)
    comm_mode = run.L.comm_mode()
    run.close()

    # extras at N = 1: the other tensor-core precision on the headline workload, and the other two
    # benchmarked configurations (short runs; the headline is the line's `value`)
    if world == 1 and not args.no_extras:
        other = "bf16" if args.math == "tf32" else "tf32"
        r2 = Run(w, other, local, 1, 0, None, use_graph, seed_rank=0)
        extras[other] = brief(w, other, H.measure(r2, args.steps, args.warmup, min_s=1.0))
        r2.close()
        for name in ("v0", "sweep", "v1"):
            if name == w["name"]:
                continue
            w2 = load_workload(name, 1)
            r2 = Run(w2, args.math, local, 1, 0, None, use_graph, seed_rank=0)
            extras[name] = brief(w2, args.math, H.measure(r2, min(args.steps, 10), args.warmup, min_s=1.0))
            r2.close()
    if sampler is not None:
        sampler.stop()

    if rank != 0:  # the learner (its NCCL communicators included) was closed gracefully above; torch's own
        sys.stdout.flush()  # process group is left to process exit
        os._exit(0)
    roof, roof_hbm = rooflines(w, args.math, res)
    cfg = shape_config(w, world)
    cfg.update({"cuda_graph": use_graph, "math": args.math, "gradient_exchange": comm_mode,
                "l2": "inputs larger than L2 (raw frames %.0f MB per step)" % (w["T"] * run.slab / 1e6)})
    line = {"metric": METRIC, "value": res["value"], "unit": "agent-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"],
            "repeats": res["repeats"], "timed_steps": res["timed_steps"],
            "timed_region_s": res["timed_region_s"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {"bf16": "bf16", "tf32": "tf32", "fp32": "f32"}[args.math],
            "data": "synthetic", "config": cfg, "e2e": res["e2e"], "gpu_launches": res["launches"],
            "clocks": sampler.summary() if sampler is not None else
            {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]},
            "roofline": roof if roof else roof_hbm, "roofline_hbm": roof_hbm,
            "stages_ms": {"preprocess": res["pre_ms"], "train": res["train_ms"]}, "also": extras}
    if not args.no_cpu_baseline and world == 1:  # reported on rank 0 at N=1 only
        threads = os.cpu_count() or 1
        E = cpu_sample_envs(w, 8192)
        v, desc = cpu_reference_iteration(w, threads, E, iters=3, warm=1)
        line["cpu_baseline"] = {"value": v, "unit": "agent-steps/s", "cores": threads,
                                "kind": "reference", "sample": desc}
    print(json.dumps(line))
    sys.stdout.flush()
    if world > 1:
        os._exit(0)  # torch's process group is not torn down collectively; nothing of ours is alive


if __name__ == "__main__":
    try:
        main()
    except BaseException:
        import traceback
        traceback.print_exc()
        sys.stdout.flush()
        sys.stderr.flush()
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            os._exit(1)  # a failed rank must not leave the others waiting in a collective
        raise
