"""The reference's Llama fine-tuning job (ray-jobs/fine_tune_llama_ray.py) on the B200 step tail, full fine-tune regime.

    python examples/fine_tune_llama_b200.py --num-workers 2 --steps 8            # 2 GPUs of this box, Llama-3.2-1B shape
    python examples/fine_tune_llama_b200.py --num-workers 1 --tiny --steps 4     # smoke-sized model

The reference builds `SFTConfig(...)` from ray-jobs/fine_tune_config.json and calls `SFTTrainer(...).train()` inside
Ray's `train_loop_per_worker` (ref :197-333). What that call executes on the hot path is HF Trainer's inner loop
($SP/transformers/trainer.py:1867-1936, 2109-2159, 2493): per optimizer step GRADIENT_ACCUMULATION_STEPS micro-batches
whose losses are summed cross-entropies divided by the step's non-ignored label count, DDP synchronising on the last one
only, clip_grad_norm_(MAX_GRAD_NORM), optimizer.step(), lr_scheduler.step(), model.zero_grad().

* If `trl` and `accelerate` are importable, the real SFTTrainer drives the step with our pieces plugged in through its own
  extension points: `optimizers=(b200_train.AdamW, scheduler)`, the patched model forward (fused LM-head loss), and
  `RayTrainReportCallback` / `prepare_trainer` from b200_train.train.huggingface.transformers.
* They are not installed in this image (no network), so the default path below runs that inner loop itself, with the same
  hyper-parameter names as fine_tune_config.json, on synthetic token batches (the gretel SQL dataset and the Llama
  tokenizer cannot be downloaded) and a randomly initialised model of the Llama-3.2-1B shape (`--tiny` shrinks it).
  QLoRA (bitsandbytes / peft) is outside the scope of the step tail; this is the USE_QLORA=false regime.
"""
import argparse
import importlib.util
import math
import os
import sys
import tempfile

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))

# same keys as ref: ray-jobs/fine_tune_config.json:12-29 (the ones the step tail reads)
CONFIG = {
    "NUM_TRAIN_EPOCHS": 1, "PER_DEVICE_TRAIN_BATCH_SIZE": 2, "GRADIENT_ACCUMULATION_STEPS": 4, "LEARNING_RATE": 2e-4,
    "WEIGHT_DECAY": 0.001, "LR_SCHEDULER_TYPE": "cosine", "MAX_GRAD_NORM": 0.3, "WARMUP_RATIO": 0.03,
    "LOGGING_STEPS": 2, "MAX_SEQ_LENGTH": 1024,
}


def cosine_with_warmup(total_steps, warmup_ratio):
    """transformers.get_cosine_schedule_with_warmup as a LambdaLR factor (lr_scheduler_type "cosine")."""
    warm = math.ceil(total_steps * warmup_ratio)

    def fn(step):
        if step < warm:
            return float(step) / float(max(1, warm))
        prog = float(step - warm) / float(max(1, total_steps - warm))
        return max(0.0, 0.5 * (1.0 + math.cos(math.pi * prog)))
    return fn


def train_loop_per_worker(config):
    import b200_train
    from b200_train import train
    import workloads as TR

    rank = train.get_context().get_world_rank()
    world_size = train.get_context().get_world_size()
    device = train.torch.get_device()                                      # ref :223 (device_map)
    shape = dict(vocab_size=4099, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                 num_key_value_heads=2, head_dim=32, max_position_embeddings=2048) if config["tiny"] else {}
    model = TR.make_llama(dtype=torch.bfloat16, seed=0, device=device, **shape)           # ref :240 (torch_dtype bf16)
    model.config.use_cache = False                                          # ref :241
    model.train()
    b200_train.patch_causal_lm(model)                                       # lm_head + ForCausalLMLoss -> fused kernel
    model = train.torch.prepare_model(model)                                # what accelerate's DDP wrap does in SFTTrainer
    inner = model.module if hasattr(model, "module") else model
    optimizer = b200_train.AdamW(TR.hf_param_groups(inner, config["WEIGHT_DECAY"]), lr=config["LEARNING_RATE"])
    total = config["max_steps"]
    scheduler = torch.optim.lr_scheduler.LambdaLR(optimizer, cosine_with_warmup(total, config["WARMUP_RATIO"]))
    ga, bs, T = config["GRADIENT_ACCUMULATION_STEPS"], config["PER_DEVICE_TRAIN_BATCH_SIZE"], config["MAX_SEQ_LENGTH"]
    V = inner.config.vocab_size

    if importlib.util.find_spec("trl") and importlib.util.find_spec("accelerate") and config.get("use_sft_trainer"):
        return _run_with_sft_trainer(config, inner, optimizer, scheduler, train)

    no_sync = getattr(model, "no_sync", None)
    last = None
    for step in range(total):
        ids, labels = TR.synthetic_hf_batch(bs * ga, T, V, seed=1000 * rank + step, device=device)
        num_items = (labels[:, 1:] != -100).sum()                          # trainer.py:2109-2159 (num_items_in_batch)
        loss = 0.0
        for k in range(ga):
            sl = slice(k * bs, (k + 1) * bs)
            ctx = no_sync() if (no_sync is not None and k < ga - 1) else _null()
            with ctx:
                out = model(input_ids=ids[sl], labels=labels[sl], num_items_in_batch=num_items)
                out.loss.backward()
            loss = loss + out.loss.detach()
        norm = b200_train.clip_grad_norm_(model.parameters(), config["MAX_GRAD_NORM"])    # trainer.py:2493
        optimizer.step()
        scheduler.step()
        model.zero_grad()                                                   # HF Trainer drops gradients this way
        if (step + 1) % config["LOGGING_STEPS"] == 0 or step + 1 == total:
            last = {"loss": float(loss), "grad_norm": float(norm), "learning_rate": scheduler.get_last_lr()[0],
                    "step": step + 1}
            if rank == 0:
                print(f"step {step + 1}/{total}  loss {last['loss']:.4f}  grad_norm {last['grad_norm']:.3f}  "
                      f"lr {last['learning_rate']:.2e}", flush=True)
    ckpt = None
    if rank == 0:                                                           # ref :336-360 saves the model from rank 0
        d = tempfile.mkdtemp(prefix="b2_sft_")
        torch.save(inner.state_dict(), os.path.join(d, "model.pth"))
        torch.save(optimizer.state_dict() if not getattr(model, "shard_optimizer", False) else {}, os.path.join(d, "optimizer.pth"))
        ckpt = train.Checkpoint.from_directory(d)
    train.report(last or {}, checkpoint=ckpt)
    if hasattr(model, "close"):
        model.close()


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def _run_with_sft_trainer(config, model, optimizer, scheduler, train):
    """With trl + accelerate present: the reference's own SFTTrainer call (ref :297-333), our optimizer and loss inside."""
    from datasets import Dataset
    from trl import SFTConfig, SFTTrainer
    from b200_train.train.huggingface.transformers import RayTrainReportCallback, prepare_trainer
    import workloads as TR
    ids, _ = TR.synthetic_hf_batch(64, config["MAX_SEQ_LENGTH"], model.config.vocab_size, seed=7)
    ds = Dataset.from_dict({"input_ids": ids.tolist()})
    args = SFTConfig(max_steps=config["max_steps"], per_device_train_batch_size=config["PER_DEVICE_TRAIN_BATCH_SIZE"],
                     gradient_accumulation_steps=config["GRADIENT_ACCUMULATION_STEPS"], max_grad_norm=config["MAX_GRAD_NORM"],
                     logging_steps=config["LOGGING_STEPS"], bf16=True, report_to=[], output_dir=tempfile.mkdtemp())
    trainer = SFTTrainer(model=model, args=args, train_dataset=ds, optimizers=(optimizer, scheduler),
                         callbacks=[RayTrainReportCallback()])
    prepare_trainer(trainer).train()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--num-workers", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--tiny", action="store_true")
    ap.add_argument("--seq-len", type=int, default=None)
    ap.add_argument("--storage", default=None)
    ap.add_argument("--use-sft-trainer", action="store_true")
    args = ap.parse_args()
    from b200_train.train import RunConfig, ScalingConfig, TorchConfig, TorchTrainer
    cfg = dict(CONFIG, max_steps=args.steps, tiny=args.tiny, use_sft_trainer=args.use_sft_trainer)
    if args.seq_len:
        cfg["MAX_SEQ_LENGTH"] = args.seq_len
    elif args.tiny:
        cfg["MAX_SEQ_LENGTH"] = 128
    trainer = TorchTrainer(                                                 # ref :445-457
        train_loop_per_worker, train_loop_config=cfg,
        scaling_config=ScalingConfig(num_workers=args.num_workers, use_gpu=True, resources_per_worker={"GPU": 1}),
        run_config=RunConfig(name="fine_tune_llama_b200", storage_path=args.storage or tempfile.mkdtemp()),
        torch_config=TorchConfig(backend="nccl"))
    result = trainer.fit()
    if result.error is not None:
        raise SystemExit(f"training failed: {result.error!r}")
    print("RESULT", result.metrics, "checkpoint:", result.checkpoint.path if result.checkpoint else None)


if __name__ == "__main__":
    main()
