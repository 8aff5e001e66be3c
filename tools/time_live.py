"""Device timing of the live-path kernels (resize_gray.cu) against the HBM roofline.
Algorithmic bytes: stand-alone resize w*h + nw*nh per image (40 656 B at 210x160 -> 84x84);
fused step 2*33 600 + 28 224 + 21 168 = 116 592 B per env step (all-reset steps carry nothing)."""
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ale-libtorch-ppo_b200"))
import alegpu  # noqa: E402

DA = alegpu.DeviceArray
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {}


def time_ms(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    L = alegpu.lib()
    e0, e1, ms = C.c_void_p(), C.c_void_p(), C.c_float()
    alegpu._check(L.alegpu_event_create(C.byref(e0)))
    alegpu._check(L.alegpu_event_create(C.byref(e1)))
    alegpu._check(L.alegpu_event_record(e0, None))
    for _ in range(iters):
        fn()
    alegpu._check(L.alegpu_event_record(e1, None))
    alegpu._check(L.alegpu_event_synchronize(e1))
    alegpu._check(L.alegpu_event_elapsed_ms(e0, e1, C.byref(ms)))
    return ms.value / iters


def main():
    rng = np.random.default_rng(0)
    out = {}
    for E in (8, 512, 4096):
        fr = DA.from_numpy(rng.integers(0, 256, (E, 2, 210, 160), dtype=np.uint8))
        starts = DA.from_numpy(np.zeros(E, np.uint8))
        a, b = DA.zeros((E, 4, 84, 84), np.uint8), DA.zeros((E, 4, 84, 84), np.uint8)
        step = lambda: alegpu.preprocess_gray_step(fr.ptr, fr.ptr + 33600, starts, a, b,
                                                   frame_env_stride=67200, E=E)
        ms = time_ms(step)
        out[f"gray_step_E{E}"] = dict(us=ms * 1e3, GBps=116592 * E / ms / 1e6)
        n = 2 * E
        o = DA.zeros((n, 84, 84), np.uint8)
        rs = lambda: alegpu._check(alegpu.lib().alegpu_resize_grayscale(
            C.c_void_p(fr.ptr), C.c_int64(n), 160, 210, 84, 84, C.c_void_p(o.ptr), None, None))
        ms = time_ms(rs)
        out[f"resize_n{n}"] = dict(us=ms * 1e3, GBps=40656 * n / ms / 1e6)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
