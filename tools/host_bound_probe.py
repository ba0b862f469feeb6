"""Is the GPT-2 step's backward bound by the GPU or by the host thread that launches it? Per step: host wall time of the
forward call, of loss.backward() (returns when the autograd thread has enqueued everything) and of the tail, next to the
GPU's own time for the same three spans (CUDA events). If host backward time ~ GPU backward time, every microsecond a
gradient hook spends on the autograd thread is a microsecond of step time.

    python tools/host_bound_probe.py [--steps 10]         (B2_LAND_LATE=1: gradients are copied into the flat buffer after
                                                           backward, on the main thread, instead of inside the hooks)
"""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--stock", action="store_true")
    a = ap.parse_args()
    import b200_train as b200
    import workloads as TR
    dev = torch.device("cuda", 0)
    model = TR.make_gpt2(dtype=torch.bfloat16, seed=0).to(dev).train()
    if a.stock:
        opt = torch.optim.AdamW(TR.hf_param_groups(model, 1e-3), lr=2e-4)
        clip = torch.nn.utils.clip_grad_norm_
    else:
        b200.patch_causal_lm(model)
        opt = b200.AdamW(TR.hf_param_groups(model, 1e-3), lr=2e-4)
        clip = b200.clip_grad_norm_
    ids, lab = TR.synthetic_hf_batch(8, 1024, 50257, 1234, device=dev)
    rows = []
    for s in range(a.steps + 3):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        torch.cuda.synchronize()
        t0 = time.perf_counter(); ev[0].record()
        opt.zero_grad(set_to_none=True)
        loss = model(input_ids=ids, labels=lab).loss
        t1 = time.perf_counter(); ev[1].record()
        loss.backward()
        t2 = time.perf_counter(); ev[2].record()
        clip(model.parameters(), 0.3)
        opt.step()
        t3 = time.perf_counter(); ev[3].record()
        torch.cuda.synchronize()
        t4 = time.perf_counter()
        if s >= 3:
            rows.append((1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]),
                         ev[2].elapsed_time(ev[3]), 1e3 * (t4 - t0)))
    m = [sum(r[i] for r in rows) / len(rows) for i in range(7)]
    print(f"{'stock' if a.stock else 'b200'} LAND_LATE={os.environ.get('B2_LAND_LATE', '0')}: host fwd {m[0]:.2f} bwd {m[1]:.2f} tail {m[2]:.2f} ms | "
          f"GPU fwd {m[3]:.2f} bwd {m[4]:.2f} tail {m[5]:.2f} ms | step wall (synchronised every step) {m[6]:.2f} ms")


if __name__ == "__main__":
    main()
