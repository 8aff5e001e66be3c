"""tools/make_golden.py — regenerate tests/golden/*.npz by running the REFERENCE's own code
(oracle/_ref/libref_oracle.so = /root/reference/src/ai/{gae,buffer,vision}.cc,
ppo/{losses,train}.cc compiled unmodified, see oracle/Makefile) on the seeded inputs of
tests/cases.py.  Only runnable where /root/reference exists (make -C oracle ref); the fixtures
it writes are committed so the tests need neither the reference nor torch.

    python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ale-libtorch-ppo_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import cases  # noqa: E402
from oracle import ref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def main():
    assert ref.available(), "build oracle/_ref first: make -C oracle ref"
    ref.set_num_threads(1)
    os.makedirs(OUT, exist_ok=True)
    # 1. the six GAE known-answer cases, through the reference's gae()
    kat = {}
    for name, c in cases.GAE_KATS.items():
        kat[name] = ref.gae(c["rewards"], c["values"], c["next_values"], c["terminals"],
                            c["truncations"], c["episode_starts"], cases.GAMMA, cases.LAMBDA)
    np.savez(os.path.join(OUT, "gae_kats.npz"), **kat)
    # 2. Buffer add x T + get on a seeded rollout (reward clamp + gae + returns + masks)
    s = cases.gae_case()
    b = ref.buffer_roundtrip(s["actions"], s["rewards"], s["terminals"], s["truncations"],
                             s["episode_starts"], s["logits"], s["values"], s["next_values"],
                             cases.GAMMA, cases.LAMBDA)
    np.savez(os.path.join(OUT, "buffer_get.npz"), **b)
    # 3. preprocess: 2-frame max + libtorch area resize + gray (+ rounding), and a stack sequence
    fr = cases.frames_case()
    g, q = ref.preprocess(fr[:, 0], fr[:, 1])
    stack = np.zeros((1, 4, 84, 84), np.uint8)
    seq = []
    for i, st in enumerate([1, 0, 0, 1, 0]):
        stack = ref.stack_update(stack, q[i:i + 1], np.array([st], np.uint8))
        seq.append(stack.copy())
    np.savez_compressed(os.path.join(OUT, "preprocess.npz"), gray_f32=g, gray_u8=q,
                        stacks=np.concatenate(seq))
    # 4. losses::compute (+ autograd grads) and normalize_logits
    lc = cases.loss_case()
    lo = ref.loss(lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"], lc["values"],
                  lc["returns"], lc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    lo["loss"] = np.float32(lo["loss"])
    lo["log_softmax"] = ref.normalize_logits(lc["logits"])
    np.savez(os.path.join(OUT, "loss.npz"), **lo)
    # 5. network forward / backward / clip+Adam / train() at hidden_size 32 (configs/debug.yaml)
    H = 32
    params = cases.net_params(H)
    tc = cases.train_case()
    L = ref.Learner(H, 4, 42, cases.LR)
    assert L.P == params.size
    L.set_params(params)
    logits, values = L.forward(tc["obs"][:6])
    c1, c2, c3, fc = L.activations(tc["obs"][:2])
    grads, loss, _, _ = L.grads(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"],
                                tc["returns"], tc["masks"], cases.CLIP, cases.VCOEF, cases.ECOEF)
    offs = cases.param_offsets(H)
    idx = cases.sample_idx(params.size)
    tn1 = L.apply_grads(grads, cases.MAX_NORM)
    p1 = L.get_params()
    tn2 = L.apply_grads(grads * 0.01, cases.MAX_NORM)  # second step: un-clipped branch
    p2 = L.get_params()
    m2, v2, st2 = L.get_adam()
    np.savez(os.path.join(OUT, "net_h32.npz"), logits=logits, values=values, conv1=c1[:, :, ::5, ::5],
             conv2=c2, conv3=c3, fc=fc, loss=np.float32(loss),
             grad_norms=np.array([np.linalg.norm(grads[offs[i]:offs[i + 1]].astype(np.float64))
                                  for i in range(12)]),
             grad_samples=grads[idx], total_norm=np.float32([tn1, tn2]), p1=p1[idx], p2=p2[idx],
             m2=m2[idx], v2=v2[idx], step=np.int64(st2))
    # 6. train(): 2 epochs x 2 minibatches of 8
    L2 = ref.Learner(H, 4, 42, cases.LR)
    L2.set_params(params)
    met = L2.train(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                   tc["masks"], 2, 2, cases.CLIP, cases.VCOEF, cases.ECOEF, cases.MAX_NORM)
    pf = L2.get_params()
    np.savez(os.path.join(OUT, "train_h32.npz"), params=pf[idx],
             param_norms=np.array([np.linalg.norm(pf[offs[i]:offs[i + 1]].astype(np.float64))
                                   for i in range(12)]), **met)
    bench_shapes()
    print("wrote", sorted(os.listdir(OUT)))


def bench_shapes():
    """7. the benchmarked shapes through the reference's own code (H = 512): a whole v0 train() and
    one v1 / one sweep minibatch (forward, loss, backward).  The fixtures keep per-tensor norms and
    2048 samples per tensor (a full gradient is 6.7 MB)."""
    ref.set_num_threads(os.cpu_count() or 1)
    offs = cases.param_offsets(512)
    samp = cases.tensor_samples(offs)
    flat_idx = np.concatenate(samp)
    out = {}
    for name, c in cases.BENCH_SHAPES.items():
        params, tc = cases.bench_shape_case(name)
        L = ref.Learner(512, 4, 42, cases.LR)
        L.set_params(params)
        if "epochs" in c:
            met = L.train(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"], tc["returns"],
                          tc["masks"], c["epochs"], c["nmb"], c["clip"], c["vcoef"], c["ecoef"],
                          cases.MAX_NORM)
            pf = L.get_params()
            out[name + ".params"] = pf[flat_idx]
            out[name + ".update"] = (pf - params)[flat_idx]
            out[name + ".update_norms"] = np.array(
                [np.linalg.norm((pf - params)[offs[i]:offs[i + 1]].astype(np.float64)) for i in range(12)])
            out[name + ".loss"] = met["loss"][:, :, 0]
            out[name + ".clipped_gradients"] = met["clipped_gradients"]
            for k in ("clipped_losses", "value_losses", "entropies", "ratio"):
                out[name + "." + k] = met[k][:, :, :32]
        else:
            g, loss, lg, v = L.grads(tc["obs"], tc["actions"], tc["log_probs"], tc["advantages"],
                                     tc["returns"], tc["masks"], c["clip"], c["vcoef"], c["ecoef"])
            out[name + ".grads"] = g[flat_idx]
            out[name + ".grad_norms"] = np.array(
                [np.linalg.norm(g[offs[i]:offs[i + 1]].astype(np.float64)) for i in range(12)])
            out[name + ".loss"] = np.float32(loss)
            out[name + ".logits"], out[name + ".values"] = lg, v
    np.savez_compressed(os.path.join(OUT, "bench_shapes.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "bench_shapes":
        assert ref.available(), "build oracle/_ref first: make -C oracle ref"
        bench_shapes()
    else:
        main()
