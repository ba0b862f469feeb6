#!/usr/bin/env python
"""bench.py — train tokens/sec of the data-parallel step on N B200s of one box (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W                 (N>1: launched by torch.distributed.run, one rank/GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W   (the reference's CPU path on the host cores)

Workload (config.workload): BASELINE configs[1] — GPT-2 small (124M, random init from GPT2Config()), full fine-tune,
seq 1024, bf16 parameters/gradients/AdamW state (the reference's full-FT dtype regime), 8 sequences
per GPU per step, synthetic wikitext-2-shaped token batches with a -100 padded tail. One step = forward (stock HF
backbone) -> fused LM-head+cross-entropy (tcgen05) -> backward -> [N>1: peer-memory gradient all-reduce] ->
clip_grad_norm_(0.3) -> fused AdamW(lr 2e-4, wd 1e-3), i.e. the reference's step tail with our three kernels in it.
Weak scaling: the per-GPU batch is fixed as N grows.

Prints ONE JSON line (rank 0). `value` = tokens/s with the batch already resident in HBM; `e2e` = the same step
driven through the public API from pinned host buffers, H2D copy and loss D2H inside the timed region.
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))

B_PER_GPU, SEQ, MAX_NORM, LR, WD = 8, 1024, 0.3, 2e-4, 1e-3
METRIC = "train tokens/sec/box"
F1_TRAFFIC_KEPT, F1_TRAFFIC_RECOMPUTE = 4.70e9, 3.18e9      # measured DRAM bytes per step of the F1 family (see f1_roofline)
# --workload: the default is the configuration BASELINE.json's metric is quoted on at N=1 (configs[1]); llama-1b is
# configs[2] (Llama-3.2-1B shape, seq 2048, 2 sequences x 4 accumulation steps per GPU, ref: fine_tune_config.json:13-14)
WORKLOADS = {
    "gpt2-124m": dict(desc="gpt2-124m full fine-tune, seq 1024, bf16 (BASELINE configs[1])", b=8, t=1024, ga=1),
    "llama-1b": dict(desc="llama-3.2-1b shape full fine-tune, seq 2048, bf16, 2 sequences x 4 accumulation steps per GPU "
                          "(BASELINE configs[2])", b=2, t=2048, ga=4),
    "llama-8b": dict(desc="llama-3-8b shape full fine-tune, 128k-vocab untied LM head, seq 2048, bf16, 1 sequence x 2 "
                          "accumulation steps per GPU (BASELINE configs[3])", b=1, t=2048, ga=2),
}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=d["hbm_gbs"], bf16_tflops=d["bf16_tflops"], bf16_tflops_sustained=d.get("bf16_tflops_sustained"),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.path = tempfile.mktemp(prefix="b2_clocks_")
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows = []
        with open(self.path) as f:
            for line in f:
                parts = [x.strip() for x in line.split(",")]
                if len(parts) >= 7:
                    rows.append(parts)
        os.unlink(self.path)
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        pw = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(rows[0][1]) if rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(rows), "power_w_max": max(pw) if pw else None}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path (stock torch/transformers call sequence, oracle/torch_ref.py)
    on the box's host cores, all threads, on a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ.pop("OMP_NUM_THREADS", None)       # torchrun sets it to 1; this arm uses every host thread it can
    import torch
    from oracle import torch_ref as TR
    torch.set_num_threads(os.cpu_count() or 1)
    steps = max(1, min(args.steps, 3))
    warm = max(1, min(args.warmup, 1))
    r = TR.cpu_gpt2_baseline(steps=steps, warmup=warm, B=1, T=SEQ, dtype=torch.bfloat16)
    gloo = configs0_gloo_baseline(TR)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["tokens_per_s"], "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "gpt2-124m full fine-tune, seq 1024, bf16 (BASELINE configs[1])",
                   "note": "reference = stock torch/transformers step (HF loss, clip_grad_norm_, torch.optim.AdamW) on host cores; "
                           "bounded sample; ray/trl/accelerate are not installed so the loop body is restated in oracle/torch_ref.py"},
        "cpu_baseline": {"value": r["tokens_per_s"], "unit": "tokens/s", "cores": r["threads"], "kind": "port", "sample": r["sample"],
                         "configs0_gloo": gloo},
        "e2e": {"value": r["tokens_per_s"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def configs0_gloo_baseline(TR):
    """BASELINE configs[0], the reference's own CPU-runnable case (ref: ray-jobs/pytorch_llm_ray.py:346-350,362-376):
    small GPT, TorchTrainer(ScalingConfig(num_workers=2, use_gpu=False)) => gloo, through the trainer shim, at the two
    thread settings SURVEY.md §8(d) names: Ray-faithful 1 thread per worker, and cores/2 per worker. A different model
    than the headline workload, so it is reported NEXT TO the GPT-2 sample, not as `value`."""
    out = {"host_cores": os.cpu_count(), "workers": 2, "backend": "gloo", "launcher": "b200_train.train.TorchTrainer (ray is not installed)"}
    # under torchrun this process carries RANK / WORLD_SIZE of the GPU job: the trainer must start its own two workers
    # (TORCHELASTIC_USE_AGENT_STORE in particular: with it set, the workers' env:// rendezvous would wait for the torchrun
    # agent's store on a port nobody serves; OMP_NUM_THREADS=1 is torchrun's default and would cap the workers' threads)
    drop = [k for k in os.environ if k.startswith("TORCHELASTIC_") or k in (
        "RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_WORLD_SIZE", "GROUP_RANK", "ROLE_RANK",
        "ROLE_NAME", "ROLE_WORLD_SIZE", "GROUP_WORLD_SIZE", "OMP_NUM_THREADS", "TORCH_NCCL_ASYNC_ERROR_HANDLING")]
    saved = {k: os.environ.pop(k) for k in drop}
    try:
        _configs0_runs(TR, out)
    finally:
        os.environ.update(saved)
    return out


def _configs0_runs(TR, out):
    for name, threads in (("threads_1_per_worker", 1), ("threads_half_cores_per_worker", max(1, (os.cpu_count() or 2) // 2))):
        try:
            g = TR.cpu_gloo_small_gpt_baseline(steps=3, warmup=1, world=2, threads=threads)
            out[name] = {"value": g["tokens_per_s"], "unit": "tokens/s", "ms_per_step": g["ms_per_step"],
                         "threads_per_worker": threads, "sample": g["sample"]}
        except Exception as e:  # noqa: BLE001 - a baseline that cannot run is reported, it does not sink the bench line
            out[name] = {"error": repr(e)[:300]}


# ------------------------------------------------------------------------------------------------ our arm
def f1_roofline(torch, b200, model, device, peaks, iters=5):
    """Times the fused LM-head+CE kernel family (forward + backward: 4 launches of the tcgen05 mainloop) alone, with CUDA
    events on the launching stream and an L2 flush between iterations. Algorithmic flops = 6*N*V*D."""
    W = model.get_output_embeddings().weight
    V, D = W.shape
    N = B_PER_GPU * SEQ
    g = torch.Generator(device="cpu").manual_seed(7)
    H = (torch.randn(N, D, generator=g) * 0.5).to(device, torch.bfloat16).requires_grad_(True)
    Wl = W.detach().clone().requires_grad_(True)
    labels = torch.randint(0, V, (B_PER_GPU, SEQ), generator=g).to(device)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=device)
    times = []
    # The GPU is kept busy (a spin kernel of ~1 ms) while the host enqueues the family behind it, and the first event is
    # recorded AFTER that kernel in stream order: the interval is then the kernels running back to back, as they do inside
    # the GPU-bound training step where the host is a step ahead — not the host's launch latency on an idle GPU.
    spin = int(2e6)
    for it in range(iters + 2):
        flush.fill_(it & 0xff)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(spin)
        e0.record()
        loss = b200.fused_lm_head_ce(H.view(B_PER_GPU, SEQ, D), Wl, labels)
        loss.backward()
        e1.record()
        torch.cuda.synchronize()
        H.grad = None
        Wl.grad = None
        if it >= 2:
            times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    flops = 6.0 * N * V * D
    achieved = flops / (ms * 1e-3) / 1e12
    # the four launches of the family, each timed alone (2*N*V*D flop per launch)
    from b200_train import ops
    y = labels.reshape(-1)
    Hd, Wd = H.detach(), Wl.detach()
    r = ops.lmhead_ce_fwd(Hd, Wd, y, shift_T=SEQ)
    ldd = (V + 63) // 64 * 64          # row pitch: a multiple of 128 bytes
    dl = torch.empty(N, ldd, dtype=torch.bfloat16, device=device)
    dH = torch.empty(N, D, dtype=torch.bfloat16, device=device)
    dW = torch.empty(V, D, dtype=torch.bfloat16, device=device)
    gs = torch.full((1,), 1.0 / N, device=device)
    from b200_train import loss as L
    keep = L._KEEP_LOGITS and N <= L.DEFAULT_CHUNK_ROWS and N * ldd * 2 <= L._KEEP_LOGITS_MAX_BYTES   # the route the loss takes
    recompute = lambda: ops.lmhead_ce_bwd_dlogits(Hd, Wd, y, r.lse, gs, 0, N, dl, shift_T=SEQ)   # noqa: E731
    if keep:
        route = (("forward", lambda: ops.lmhead_ce_fwd(Hd, Wd, y, shift_T=SEQ, keep_logits=dl)),
                 ("dlogits", lambda: ops.ce_dlogits_inplace_(dl, y, r.lse, gs, 0, N, V, shift_T=SEQ)))
    else:
        route = (("forward", lambda: ops.lmhead_ce_fwd(Hd, Wd, y, shift_T=SEQ)), ("dlogits", recompute))
    parts = {}
    for name, fn in route + (("dH", lambda: ops.gemm_bf16(dl, Wd, dH, N, D, V, b_mn_major=True)),
                             ("dW", lambda: ops.gemm_bf16(dl, Hd, dW, V, D, N, a_mn_major=True, b_mn_major=True)),
                             ("dlogits_recompute_route", recompute)):
        ts = []
        for it in range(iters + 1):
            flush.fill_(it & 0xff)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(spin)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            if it >= 1:
                ts.append(e0.elapsed_time(e1))
        t = sum(ts) / len(ts)
        if keep and name == "dlogits":      # not a GEMM: one read + one write of the [N, V] bf16 matrix
            parts[name] = {"ms": t, "gbs": 4.0 * N * V / (t * 1e-3) / 1e9, "bound": "hbm"}
        else:
            parts[name] = {"ms": t, "tflops": 2.0 * N * V * D / (t * 1e-3) / 1e12}
    other = parts.pop("dlogits_recompute_route")
    if keep:
        kernel = ("umma_kernel family per step: fused LM-head+CE forward that keeps its bf16 logits, in-place soft-max "
                  "gradient (HBM-bound, no tensor cores), dH GEMM, dW GEMM (timed together through autograd; "
                  "6*N*V*D algorithmic flop = executed flop)")
    else:
        kernel = ("umma_kernel family per step: fused LM-head+CE forward, dlogits recompute, dH GEMM, dW GEMM "
                  "(timed together through autograd; 6*N*V*D algorithmic flop, 8*N*V*D executed)")
    return {"bound": "tensor", "achieved": achieved, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
            "frac": achieved / peaks["bf16_tflops"],
            # dram read+write summed over the family's launches from ncu --set full (profiles/README.md): measured for the
            # recompute route only (3.13e9); the kept-logits route has not been captured yet
            # dram__bytes_read + dram__bytes_write summed over the family's launches, one `ncu --set full` capture per route
            # (profiles/r2_f1_ncu_metrics_final.json): kept route 0.89 (forward+store) + 1.59 (in-place gradient) + 1.14 (dH incl.
            # split-K reduce) + 1.09 (dW) GB; recompute route 0.10 + 0.86 + 1.14 + 1.09 GB. Algorithmic (no N*V term): 0.347 GB
            # (captured at the GPT-2 head, N = 8192, D = 768, V = 50 257: null for the other workloads' head shapes)
            "traffic": (F1_TRAFFIC_KEPT if keep else F1_TRAFFIC_RECOMPUTE) if (N, D, V) == (8192, 768, 50257) else None,
            "traffic_recompute_route": F1_TRAFFIC_RECOMPUTE if (N, D, V) == (8192, 768, 50257) else None,
            "route": "kept-logits" if keep else "recompute",
            "kernel": kernel,
            "ms": ms, "algorithmic_flops": flops, "launches": parts, "dlogits_recompute_route": other,
            # the same 6*N*V*D over the SUM of the four launch durations (no host gaps, no combine/finalize/reduce helpers)
            "achieved_sum_of_launches": flops / (sum(v["ms"] for v in parts.values()) * 1e-3) / 1e12,
            "peak_source": peaks["source"] + ", burst bf16 GEMM",
            "timing": "CUDA events on the launching stream, 256 MB L2 flush before every iteration, synchronize after; a ~1 ms "
                      "spin kernel ahead of the first event lets the host enqueue the family while the GPU is busy, so the "
                      "interval holds the kernels back to back (as in the GPU-bound step), not host launch latency"}


def f3_roofline(torch, ops, device, peaks, n=1 << 27, iters=5):
    """Fused AdamW alone on a flat pure-bf16 buffer larger than L2 (14 B/element algorithmic)."""
    p = torch.randn(n, device=device).to(torch.bfloat16)
    g = (torch.randn(n, device=device) * 0.01).to(torch.bfloat16)
    m = torch.zeros(n, dtype=torch.bfloat16, device=device)
    v = torch.zeros(n, dtype=torch.bfloat16, device=device)
    times = []
    for it in range(iters + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(int(2e5))                           # (see f1_roofline: no host launch latency inside the interval)
        e0.record()
        ops.adamw_(p, g, m, v, lr=LR, weight_decay=WD, step=it + 1)
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    gbs = 14.0 * n / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
            "traffic": 1.823e9, "kernel": "adamw_kernel (pure bf16, n=2^27; 14 B/element = 1.879e9 algorithmic bytes)", "ms": ms}


def f2_parity_check(torch, dist, model, optimizer, ops, batch, B):
    """N > 1, before anything is timed: is what the peer-memory kernels leave in the flat gradient buffer the mean over
    ranks? One un-timed backward under no_sync() gives this rank's LOCAL gradients; NCCL — as the checker only — forms
    their mean in fp32; then the wrapper's own bucketed reduction (model.reduce_gradients(): same kernels, same buckets
    as the timed steps) runs and is compared: max error in bf16 ulps, identical bits on every rank, and the fused sum
    of squares against b2_sumsq of the reduced buffer. With a sharded optimizer only this rank's slices are compared."""
    rank, world = dist.get_rank(), dist.get_world_size()
    ids, lab = batch
    optimizer.zero_grad(set_to_none=True)
    with model.no_sync():
        model(input_ids=ids[:B], labels=lab[:B]).loss.backward()
    res = {"f2": "ok", "max_ulp": 0.0, "max_err_over_tol": 0.0, "kernel": None, "identical_across_ranks": True, "sumsq_rel_err": 0.0,
           "sharded": bool(model.shard_optimizer), "buckets": len(model._buckets)}
    CH = 1 << 26                                           # elements per piece: the checker must not need GBs of its own
    refs = []
    for buf in model.flat_buffers:
        buf.install_grads()
        ref = torch.empty_like(buf.grad)
        mag = torch.empty_like(buf.grad)                   # mean over ranks of |g|: the scale of the fp32 summation error
        for a in range(0, buf.numel, CH):
            piece = buf.grad[a:a + CH].float()
            absp = piece.abs()
            dist.all_reduce(piece)                         # NCCL: the checker
            dist.all_reduce(absp)
            ref[a:a + CH] = (piece * (1.0 / world)).to(buf.dtype)
            mag[a:a + CH] = (absp * (1.0 / world)).to(buf.dtype)
        refs.append((ref, mag))
    model.reduce_gradients()
    torch.cuda.synchronize()
    for buf, (ref, mag) in zip(model.flat_buffers, refs):
        res["kernel"] = model.comm.kernel_name(buf.grad)
        segs = [(lo, hi) for lo, hi, _ in buf.shards] if buf.shards is not None else [(0, buf.numel)]
        mant_bits = 8 if buf.dtype == torch.bfloat16 else 24
        chk = torch.zeros(1, dtype=torch.int64, device=buf.grad.device)
        direct = 0.0
        for lo, hi in segs:
            for a in range(lo, hi, CH):
                b = min(a + CH, hi)
                got, want = buf.grad[a:b].float(), ref[a:b].float()
                ulp = torch.ldexp(torch.ones_like(want), torch.frexp(want.abs().clamp_min(1e-30))[1] - mant_bits)
                # one rounding of the result, plus what a different fp32 summation ORDER can move the sum by when the
                # addends cancel (the sum is far smaller than they are): W fp32 roundings at the addends' magnitude
                slack = mag[a:b].float() * (world * 2.0 ** -23)
                res["max_ulp"] = max(res["max_ulp"], float(((got - want).abs() / ulp).max()))
                res["max_err_over_tol"] = max(res["max_err_over_tol"], float(((got - want).abs() / (ulp + slack)).max()))
                chk += buf.grad[a:b].view(torch.int16 if buf.dtype == torch.bfloat16 else torch.int32).to(torch.int64).sum()
            if hi > lo:
                direct += float(ops.sumsq(buf.grad[lo:hi]).item())
        if buf.shards is None:                             # replicated result: every rank must hold the same bits
            all_chk = [torch.empty_like(chk) for _ in range(world)]
            dist.all_gather(all_chk, chk)
            res["identical_across_ranks"] &= all(int(c) == int(all_chk[0]) for c in all_chk)
            fused = float(buf.reduced_sumsq.double().sum())
        else:
            fused = float(buf.reduced_sumsq[rank])
        res["sumsq_rel_err"] = max(res["sumsq_rel_err"], abs(fused - direct) / max(direct, 1e-30))
    del refs
    # W = 2: the fp32 sum of two values is exact, so the kernels must agree with NCCL bit for bit (max_ulp == 0). W > 2:
    # the summation order differs (ours: rank order; NVLS: inside the switch; NCCL: ring), so the bound is one rounding
    # of the result plus the fp32 reassociation error at the magnitude of the addends (max_err_over_tol <= 1); max_ulp,
    # measured at the magnitude of the RESULT, is reported too and exceeds 1 only where the ranks' gradients cancel
    ok = (res["max_ulp"] == 0.0 if world == 2 else res["max_err_over_tol"] <= 1.001) \
        and res["identical_across_ranks"] and res["sumsq_rel_err"] < 1e-4
    flag = torch.tensor([0 if ok else 1], device=ids.device)
    dist.all_reduce(flag)
    if int(flag):
        res["f2"] = "mismatch"
    optimizer.zero_grad(set_to_none=True)
    return res


def f2_roofline(torch, dist, comm, device, n, label, iters=8):
    """The peer-memory all-reduce alone on an n-element bf16 symmetric buffer: CUDA events on the launching stream, max
    over ranks, NCCL's all_reduce timed the same way beside it. Algorithmic bytes per GPU per direction = 2 (W-1)/W n s
    (the busbw convention; with the in-switch NVLS reduction each GPU actually moves about n s, so > 100 % of the link
    figure is possible and is labelled by `kernel`)."""
    world = dist.get_world_size()
    t = comm.alloc(n, torch.bfloat16, device)
    t.normal_(0, 1e-2)
    ref = torch.empty(n, dtype=torch.bfloat16, device=device).normal_(0, 1e-2)

    def timed(fn):
        ts = []
        for it in range(iters + 2):
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            if it >= 2:
                ts.append(e0.elapsed_time(e1))
        ms = torch.tensor([sum(ts) / len(ts)], device=device)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    ms = timed(lambda: comm.allreduce_(t, want_sumsq=True))
    ms_nccl = timed(lambda: dist.all_reduce(ref))
    bytes_dir = 2.0 * (world - 1) / world * n * 2
    gbs = bytes_dir / (ms * 1e-3) / 1e9
    return {"bound": "nvlink", "achieved": gbs, "peak": 900.0, "unit": "GB/s", "frac": gbs / 900.0,
            "peak_source": "NVLink 5 nominal, per direction per GPU (a plain peer copy from a kernel sustains ~775 here: profiles/r1_nvlink_raw_copy_probe.txt)",
            "traffic": None, "ms": ms, "algorithmic_bytes": bytes_dir, "world": world,
            "kernel": f"peer_allreduce ({comm.kernel_name(t)}) + fused 1/W scale + sum of squares, {label}",
            "nccl_ms": ms_nccl, "nccl_busbw_gbs": bytes_dir / (ms_nccl * 1e-3) / 1e9}


def stock_gpu_steps(torch, TR, device, vocab=None, steps=3, warmup=2):
    """Informational: the reference's stock torch path (HF loss, clip_grad_norm_, torch.optim.AdamW) on THIS GPU, same
    model / batch shape, a few steps. vocab=50304 pads the vocabulary to a multiple of 64 — cuBLAS runs the V = 50257
    head at ~100 TFLOP/s, which flatters any comparison with it."""
    kw = {} if vocab is None else {"vocab_size": vocab}
    model = TR.make_gpt2(dtype=torch.bfloat16, seed=0, **kw).to(device).train()
    opt = torch.optim.AdamW(TR.hf_param_groups(model, WD), lr=LR)
    ids, lab = TR.synthetic_hf_batch(B_PER_GPU, SEQ, 50257, 1234, device=device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for s in range(warmup + steps):
        if s == warmup:
            torch.cuda.synchronize()
            e0.record()
        opt.zero_grad(set_to_none=True)
        model(input_ids=ids, labels=lab).loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), MAX_NORM)
        opt.step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    import b200_train as b200
    import workloads as TR                   # model factories + synthetic inputs, shared with the reference arm
    from b200_train import _lib, ops

    peaks = load_peaks()
    global B_PER_GPU, SEQ
    wl = WORKLOADS[args.workload]
    B_PER_GPU, SEQ, GA = wl["b"], wl["t"], wl["ga"]
    if args.workload == "gpt2-124m":
        model = TR.make_gpt2(dtype=torch.bfloat16, seed=0)
    else:
        model = TR.make_llama(dtype=torch.bfloat16, seed=0, device=device,
                              **(TR.LLAMA_8B if args.workload == "llama-8b" else {}))
    model.train()
    stock = args.impl == "stock-gpu"       # informational third arm: the reference's stock torch path on THIS GPU
    if stock:
        model = model.to(device)
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank])
        inner = model.module if hasattr(model, "module") else model
        optimizer = torch.optim.AdamW(TR.hf_param_groups(inner, WD), lr=LR)
        clip = torch.nn.utils.clip_grad_norm_
    else:
        b200.patch_causal_lm(model)
        # .to(device); peer-memory DDP when world > 1 (--shard-optimizer: reduce-scatter -> AdamW on the shard -> all-gather)
        psk = None
        if args.shard_optimizer:
            psk = {"shard_optimizer": True, "overlap_param_sync": args.shard_optimizer >= 2}
        model = b200.prepare_model(model, parallel_strategy_kwargs=psk)
        inner = model.module if hasattr(model, "module") else model
        optimizer = b200.AdamW(TR.hf_param_groups(inner, WD), lr=LR)
        clip = b200.clip_grad_norm_
    V = inner.config.vocab_size

    n_batches = 4
    host = [TR.synthetic_hf_batch(B_PER_GPU * GA, SEQ, V, 1234 + 97 * rank + i) for i in range(n_batches)]
    pinned = [(a.pin_memory(), b.pin_memory()) for a, b in host]
    resident = [(a.to(device), b.to(device)) for a, b in host]
    no_sync = getattr(model, "no_sync", None)

    def step(ids, labels):
        optimizer.zero_grad(set_to_none=True)
        if GA == 1:
            loss = model(input_ids=ids, labels=labels).loss
            loss.backward()
        else:
            # HF Trainer's accumulation: every micro-batch loss is sum-CE / (non-ignored labels of the whole step),
            # gradient sync only on the last micro-batch ($SP/transformers/trainer.py:2109-2159; DDP.no_sync)
            num_items = (labels[:, 1:] != -100).sum()
            loss = 0.0
            for k in range(GA):
                sl = slice(k * B_PER_GPU, (k + 1) * B_PER_GPU)
                ctx = no_sync() if (no_sync is not None and k < GA - 1) else contextlib.nullcontext()
                with ctx:
                    l = model(input_ids=ids[sl], labels=labels[sl], num_items_in_batch=num_items).loss
                    l.backward()
                loss = loss + l.detach()
        clip(model.parameters(), MAX_NORM)
        optimizer.step()
        return loss

    def timed(batches_on_host):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = _lib.launch_count
        e0.record()
        last = None
        for s in range(args.steps):
            if batches_on_host:
                a, b = pinned[s % n_batches]
                ids, lab = a.to(device, non_blocking=True), b.to(device, non_blocking=True)
                last = float(step(ids, lab).item())          # D2H of the step's loss, every step
            else:
                ids, lab = resident[s % n_batches]
                last = step(ids, lab)
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), _lib.launch_count - l0, (last if isinstance(last, float) else float(last.item()))

    for s in range(max(args.warmup, 3)):
        step(*resident[s % n_batches])
    torch.cuda.synchronize()
    parity = None
    if world > 1 and not stock:
        parity = f2_parity_check(torch, dist, model, optimizer, ops, resident[0], B_PER_GPU)
        if parity["f2"] != "ok":
            if rank == 0:
                print(json.dumps({"error": "peer-memory gradient reduction disagrees with the NCCL check", "parity_check": parity}), flush=True)
            raise SystemExit(3)
        step(*resident[0])                                  # back to a normal step before timing
        torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats(device)             # peak memory of the timed steps, not of the checker above

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_total, launches, last_loss = timed(False)
    ms_e2e, _, _ = timed(True)
    mem_gb = torch.cuda.max_memory_allocated(device) / 1e9
    clocks = sampler.stop() if sampler else None      # sampled across both timed regions

    trace = None
    if world > 1 and not stock and os.environ.get("B2_DDP_TRACE") == "1":
        step(*resident[0])
        trace = model.trace_report()                         # rank 0's view of the last backward's bucket timeline
    tokens_per_step = world * B_PER_GPU * GA * SEQ
    value = tokens_per_step * args.steps / (ms_total * 1e-3)
    e2e = tokens_per_step * args.steps / (ms_e2e * 1e-3)

    roof = cpu = extra = None
    f2 = []
    if world > 1 and not stock:
        # F2 alone (all ranks take part): on a buffer the size of this model's flat gradients, and on configs[4]'s 2^30 bf16
        n_model = sum(b.numel for b in model.flat_buffers)
        f2.append(f2_roofline(torch, dist, model.comm, device, n_model, f"n = {n_model} bf16 (this model's flat gradient buffer)"))
        if not args.no_f2_large:
            f2.append(f2_roofline(torch, dist, model.comm, device, 1 << 30, "n = 2^30 bf16 (BASELINE configs[4])"))
    if rank == 0 and not stock:
        roof = f1_roofline(torch, b200, inner, device, peaks)
        extra = [f3_roofline(torch, ops, device, peaks)] + f2
    if world > 1:
        dist.barrier()
    stock_ms = None
    if rank == 0 and world == 1 and not stock and args.workload == "gpt2-124m" and not args.no_stock:
        del optimizer
        stock_ms = {"stock_gpu_ms_per_step": stock_gpu_steps(torch, TR, device),
                    "stock_gpu_ms_per_step_vocab_padded_to_50304": stock_gpu_steps(torch, TR, device, vocab=50304),
                    "note": "the reference's stock torch path on this GPU, 3 steps after 2 warm-up (informational; "
                            "the padded-vocabulary run takes cuBLAS's misaligned-N slowdown out of the comparison)"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not stock and args.workload == "gpt2-124m":
        # the one place this arm touches oracle/: the reported CPU baseline (the reference's stock call sequence)
        from oracle import torch_ref as REF
        torch.set_num_threads(os.cpu_count() or 1)
        r = REF.cpu_gpt2_baseline(steps=2, warmup=1, B=1, T=SEQ, dtype=torch.bfloat16)
        cpu = {"value": r["tokens_per_s"], "unit": "tokens/s", "cores": r["threads"], "kind": "port", "sample": r["sample"],
               "configs0_gloo": configs0_gloo_baseline(REF)}
    if rank == 0:
        line = {
            **({"impl": "stock-gpu"} if stock else {}),
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": wl["desc"], "global_batch": world * B_PER_GPU * GA, "grad_accum": GA,
                       "seq_len": SEQ, "parallelism": f"dp{world}",
                       "optimizer": "AdamW lr 2e-4 wd 1e-3, clip 0.3, pure-bf16 state"
                                    + (", sharded over ranks (reduce-scatter / AdamW on the shard / all-gather"
                                       + (", gather under the next forward)" if args.shard_optimizer >= 2 else ")") if args.shard_optimizer and world > 1 else ""),
                       "l2": "no explicit flush between steps: one step streams params+grads+AdamW state (1.0 GB) and >4 GB of "
                             "activations, far beyond the 126 MB L2; the roofline micro-runs flush 256 MB between iterations",
                       "last_loss": last_loss},
            "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": 2 * B_PER_GPU * GA * SEQ * 8, "d2h_bytes_per_step": 4,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roof, "rooflines_other": extra, "cpu_baseline": cpu,
            "parity_check": parity, "stock_gpu": stock_ms, "ddp_trace": trace,
            # torch's caching allocator only: at N > 1 the flat parameter and gradient buffers (2 x P x s bytes) live in
            # symmetric memory outside it; the AdamW state, activations and the kept logits are inside
            "max_memory_allocated_gb": mem_gb,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        if hasattr(model, "close"):
            model.close()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "stock-gpu"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stock", action="store_true", help="skip the informational stock-torch-on-this-GPU steps")
    ap.add_argument("--no-f2-large", action="store_true", help="N > 1: skip the 2^30-element all-reduce microbenchmark")
    ap.add_argument("--shard-optimizer", type=int, default=0, choices=[0, 1, 2],
                    help="N > 1: 1 = reduce-scatter / AdamW on the shard / all-gather; 2 = also overlap the all-gather with the next forward")
    ap.add_argument("--workload", default="gpt2-124m", choices=sorted(WORKLOADS))
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch ourselves one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__), "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup), "--impl", args.impl, "--workload", args.workload, "--shard-optimizer", str(args.shard_optimizer)]
        cmd += (["--no-cpu-baseline"] if args.no_cpu_baseline else []) + (["--no-f2-large"] if args.no_f2_large else [])
        raise SystemExit(subprocess.call(cmd))
    run_ours(args)


if __name__ == "__main__":
    main()
