"""The reference's step-tail CALL SEQUENCE on stock PyTorch — TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs import this. It holds
no kernels of ours: everything here is torch / transformers as the reference invokes them, restated because the
reference scripts cannot be imported on the GPU box (/root/reference is absent there, and `ray`, `trl`, `accelerate`
are not installed anywhere — SURVEY.md §8c).

* models and synthetic inputs come from the neutral `workloads` module at the repository root (re-exported here).
* `basic_llm_step` is the loop body ref :272-279; `hf_causal_lm_step` is the HF Trainer step that
  ref: ray-jobs/fine_tune_llama_ray.py:333 reaches ($SP/transformers/trainer.py:1867-1936, 2493), with the
  hyper-parameters of ref: ray-jobs/fine_tune_config.json:12-29.
"""
import os
import time

import torch
import torch.nn as nn

from workloads import (BasicLLMRestated, HF_LOOP, REF_LOOP, REF_MAIN, SMALL_GPT, disable_dropout,  # noqa: F401
                       hf_param_groups, lr_lambda_factory, make_gpt2, make_llama, synthetic_char_stream,
                       synthetic_hf_batch, windows)


# ------------------------------------------------------------------------------------------------ the step, stock
def basic_llm_step(model, optimizer, scheduler, criterion, inputs, targets, vocab, max_norm=1.0):
    """ref: ray-jobs/pytorch_llm_ray.py:272-279, verbatim call sequence. Returns (loss tensor, grad-norm tensor)."""
    optimizer.zero_grad(set_to_none=True)
    outputs = model(inputs)
    loss = criterion(outputs.reshape(-1, vocab), targets.reshape(-1))
    loss.backward()
    norm = torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=max_norm)
    optimizer.step()
    if scheduler is not None:
        scheduler.step()
    return loss.detach(), norm


def hf_causal_lm_step(model, optimizer, micro_batches, max_norm=0.3, no_sync=None):
    """One optimizer step of HF Trainer over `micro_batches` = [(input_ids, labels), ...] (gradient accumulation):
    num_items_in_batch = number of non-ignored SHIFTED labels over all micro-batches ($SP/transformers/trainer.py:
    2109-2159); each micro-batch loss = sum-CE / num_items ($SP/transformers/loss/loss_utils.py:28-67); backward;
    clip_grad_norm_(max_norm) (:2493); optimizer.step(); zero_grad. Returns (summed loss, grad norm)."""
    num_items = sum(int((lab[..., 1:] != -100).sum()) for _, lab in micro_batches)
    total = 0.0
    for i, (ids, lab) in enumerate(micro_batches):
        ctx = no_sync() if (no_sync is not None and i < len(micro_batches) - 1) else _null()
        with ctx:
            out = model(input_ids=ids, labels=lab, num_items_in_batch=num_items) if len(micro_batches) > 1 \
                else model(input_ids=ids, labels=lab)
            out.loss.backward()
        total = total + out.loss.detach()
    norm = torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm)
    optimizer.step()
    optimizer.zero_grad(set_to_none=True)
    return total, norm


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# ------------------------------------------------------------------------------------------------ CPU baselines
def cpu_gpt2_baseline(steps=2, warmup=1, B=1, T=1024, dtype=torch.float32, threads=None, seed=0):
    """The reference path for BASELINE configs[1] on host cores: stock HF GPT-2 124M forward+loss, backward,
    clip_grad_norm_(0.3), torch.optim.AdamW(lr 2e-4, wd 1e-3) — a bounded sample (B sequences of T tokens per step
    instead of 8). Returns dict(tokens_per_s, ms_per_step, cores, threads, sample)."""
    if threads:
        torch.set_num_threads(threads)
    model = make_gpt2(dtype=dtype, seed=seed)
    model.train()
    disable_dropout(model)
    opt = torch.optim.AdamW(hf_param_groups(model, HF_LOOP["weight_decay"]), lr=HF_LOOP["lr"])
    times = []
    for s in range(warmup + steps):
        ids, lab = synthetic_hf_batch(B, T, model.config.vocab_size, 1234 + s)
        t0 = time.perf_counter()
        hf_causal_lm_step(model, opt, [(ids, lab)], HF_LOOP["max_norm"])
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    return dict(tokens_per_s=B * T / (ms / 1e3), ms_per_step=ms, cores=os.cpu_count(), threads=torch.get_num_threads(),
                sample=f"GPT-2 124M {str(dtype).split('.')[-1]} on CPU, B={B} x T={T} per step, {steps} steps after {warmup} warm-up")


def small_gpt_gloo_loop(config):
    """train_loop_per_worker of BASELINE configs[0]: the reference's loop (ref: ray-jobs/pytorch_llm_ray.py:121-312, body
    :272-279) on its small GPT with `use_gpu=False` => gloo, written against the `ray.train` surface exactly as the
    reference writes it — here that surface is the b200_train.train shim (ray is not installed), whose prepare_model
    returns torch's own DistributedDataParallel for CPU workers. Stock torch arithmetic throughout."""
    import torch.distributed as dist
    from b200_train import train
    torch.set_num_threads(config["threads"])
    ctx = train.get_context()
    rank, world = ctx.get_world_rank(), ctx.get_world_size()
    V = 283
    stream = synthetic_char_stream(vocab=V)
    torch.manual_seed(0)
    model = train.torch.prepare_model(BasicLLMRestated(V, **SMALL_GPT))          # ref :230
    assert next(model.parameters()).device.type == "cpu"
    opt = torch.optim.AdamW(model.parameters(), lr=REF_LOOP["lr"], weight_decay=REF_LOOP["weight_decay"])   # ref :236
    crit = nn.CrossEntropyLoss()                                                   # ref :237
    B, T = REF_LOOP["batch_size_per_worker"], REF_LOOP["dataset_seq_len"]
    g = torch.Generator().manual_seed(0)
    perm = torch.randperm(16000, generator=g)[rank::world]            # DistributedSampler(shuffle=True, seed=0), ref :206-216
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lr_lambda_factory(len(perm) // B))
    times = []
    steps, warmup = config["steps"], config["warmup"]
    for s in range(warmup + steps):
        x, y = windows(stream, perm[s * B:(s + 1) * B], T)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        basic_llm_step(model, opt, sched, crit, x, y, V, REF_LOOP["max_norm"])
        if world > 1:
            dist.barrier()
        if s >= warmup:
            times.append(time.perf_counter() - t0)
    train.report({"sec_per_step": sum(times) / len(times), "threads": torch.get_num_threads()})


def cpu_gloo_small_gpt_baseline(steps=5, warmup=2, world=2, threads=None):
    """BASELINE configs[0]: the reference loop (BasicLLM small GPT, synthetic wikitext-2-shaped chars, DDP over gloo,
    num_workers=2, use_gpu=False) timed on host cores, launched through TorchTrainer(ScalingConfig(num_workers=2,
    use_gpu=False)) of the trainer shim (ref: ray-jobs/pytorch_llm_ray.py:346-350,362-376). Returns dict(tokens_per_s, ...)."""
    from b200_train.train import ScalingConfig, TorchConfig, TorchTrainer
    threads = threads or max(1, (os.cpu_count() or 2) // world)
    res = TorchTrainer(small_gpt_gloo_loop, train_loop_config={"steps": steps, "warmup": warmup, "threads": threads},
                       scaling_config=ScalingConfig(num_workers=world, use_gpu=False),
                       torch_config=TorchConfig(backend="gloo", timeout_s=240)).fit()
    if res.error is not None or not res.metrics:
        raise RuntimeError(f"CPU gloo baseline failed: {res.error!r}")
    sec = res.metrics["sec_per_step"]
    B, T = REF_LOOP["batch_size_per_worker"], REF_LOOP["dataset_seq_len"]
    return dict(tokens_per_s=world * B * T / sec, ms_per_step=1e3 * sec, cores=os.cpu_count(), threads=threads * world,
                threads_per_worker=threads, workers=world,
                sample=f"BasicLLM small GPT fp32 (BASELINE configs[0]), TorchTrainer shim, {world} gloo workers x {threads} "
                       f"threads, B={B} x T={T} per worker, {steps} steps after {warmup} warm-up")
