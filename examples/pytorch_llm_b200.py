"""The reference's BasicLLM job (ray-jobs/pytorch_llm_ray.py) on the B200 step tail — same loop shape, our drop-ins.

    python examples/pytorch_llm_b200.py --num-workers 2 --steps 20        # 2 GPUs of this box
    python examples/pytorch_llm_b200.py --num-workers 1 --bf16            # 1 GPU, bf16 model -> fused LM-head loss

What changed with respect to the reference's train_loop_per_worker (ref :122-313) is only WHERE four things come from:
`train` / TorchTrainer (b200_train.train instead of ray.train), `AdamW`, `CrossEntropyLoss`, `clip_grad_norm_`
(b200_train instead of torch). Data: the synthetic wikitext-2-shaped character stream of SURVEY.md §8(d) replaces the
GCS-FUSE file (no network here); the model is the restated BasicLLM, bit-identical to the reference class
(oracle/make_golden.py). In bf16 mode the last Linear is replaced by the fused LM-head + cross-entropy kernel.
"""
import argparse
import math
import os
import sys
import tempfile

import torch
from torch.optim.lr_scheduler import LambdaLR

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def train_loop_per_worker(config):
    import b200_train
    from b200_train import train
    import workloads as TR                       # model definition + synthetic data

    session = train.get_context()
    world_rank, world_size = session.get_world_rank(), session.get_world_size()
    device = train.torch.get_device()
    vocab = config["vocab_size"]
    seq_len, bs = config["dataset_seq_len"], config["batch_size_per_worker"]

    all_token_ids = TR.synthetic_char_stream(config["n_tokens"], vocab)
    loader = b200_train.ResidentWindowLoader(all_token_ids, seq_len, bs, device, shuffle=True, num_replicas=world_size,
                                             rank=world_rank, limit=16000 if config["test_run"] else None)
    num_batches = min(len(loader), config["max_steps"])

    torch.manual_seed(0)
    model = TR.BasicLLMRestated(vocab, config["embed_dim"], config["num_heads"], config["num_layers"], config["hidden_dim"],
                                config["model_max_seq_len"])
    if config["bf16"]:
        model = model.to(torch.bfloat16)
    model = train.torch.prepare_model(model)
    inner = model.module if hasattr(model, "module") else model

    stock = config.get("stock", False)         # --stock: the reference's own torch calls, for a side-by-side timing
    optimizer = (torch.optim.AdamW if stock else b200_train.AdamW)(model.parameters(), lr=config["lr"], weight_decay=0.01)
    criterion = torch.nn.CrossEntropyLoss() if stock else b200_train.CrossEntropyLoss()
    clip = torch.nn.utils.clip_grad_norm_ if stock else b200_train.clip_grad_norm_
    total = config["num_epochs"] * num_batches
    warm = int(total * config["warmup_steps_ratio"])

    def lr_lambda(step):
        if step < warm:
            return float(step) / float(max(1, warm))
        if step < total:
            prog = float(step - warm) / float(max(1, total - warm))
            return config["min_lr_ratio"] + (1.0 - config["min_lr_ratio"]) * 0.5 * (1.0 + math.cos(math.pi * prog))
        return config["min_lr_ratio"]
    scheduler = LambdaLR(optimizer, lr_lambda)

    step = 0
    for epoch in range(config["num_epochs"]):
        loader.sampler.set_epoch(epoch)
        for batch_idx, (inputs, targets) in enumerate(loader):
            if batch_idx >= num_batches:
                break
            if batch_idx == config["timing_warmup"]:
                t_start = torch.cuda.Event(enable_timing=True); t_start.record()
            optimizer.zero_grad(set_to_none=True)
            if config["bf16"] and not stock:
                # hidden states -> fused LM head + CE (logits never materialised); labels are pre-shifted by the dataset
                hidden = inner.features(inputs) if not hasattr(model, "module") else _features(model, inputs)
                loss = b200_train.fused_lm_head_ce(hidden, inner.fc_out.weight, targets, inner.fc_out.bias, shift=False)
            else:
                outputs = model(inputs)
                loss = criterion(outputs.reshape(-1, vocab), targets.reshape(-1))
            loss.backward()
            clip(model.parameters(), max_norm=1.0)
            optimizer.step()
            scheduler.step()
            step += 1
            if world_rank == 0 and (batch_idx % config["log_frequency_batches"] == 0 or batch_idx == num_batches - 1):
                print(f"Epoch {epoch + 1}, Batch {batch_idx + 1}/{num_batches}, Step {step}, Loss: {loss.item():.4f}, "
                      f"LR: {scheduler.get_last_lr()[0]:.2e}", flush=True)
        if num_batches > config["timing_warmup"]:
            t_end = torch.cuda.Event(enable_timing=True); t_end.record(); torch.cuda.synchronize()
            timed = num_batches - config["timing_warmup"]
            ms = t_start.elapsed_time(t_end) / timed
            if world_rank == 0:
                print(f"TIMING {ms:.2f} ms/step, {world_size * bs * seq_len / ms * 1e3:.0f} tokens/s over {timed} steps "
                      f"({'stock torch' if stock else 'b200_train'}, {world_size} GPU(s))", flush=True)
        metrics = {"loss": loss.item(), "epoch": epoch + 1, "learning_rate_epoch_end": scheduler.get_last_lr()[0],
                   "global_step_epoch_end": step}
        with tempfile.TemporaryDirectory() as tmp:
            checkpoint = None
            osd = optimizer.state_dict()                       # (collective when the optimizer state is sharded)
            if world_rank == 0:
                torch.save(inner.state_dict(), os.path.join(tmp, "model.pth"))
                torch.save(osd, os.path.join(tmp, "optimizer.pth"))
                torch.save(scheduler.state_dict(), os.path.join(tmp, "scheduler.pth"))
                checkpoint = train.Checkpoint.from_directory(tmp)
            train.report(metrics=metrics, checkpoint=checkpoint)


def _features(ddp_model, inputs):
    """Run the wrapped module's feature extractor through the wrapper's forward-side bookkeeping."""
    ddp_model.forward_prepare()
    return ddp_model.module.features(inputs)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--num-workers", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--small", action="store_true", help="BASELINE configs[0] small GPT instead of the reference's 1.2 B model")
    ap.add_argument("--stock", action="store_true", help="use torch's AdamW / CrossEntropyLoss / clip_grad_norm_ (timing comparison)")
    ap.add_argument("--storage", default=os.path.join(tempfile.gettempdir(), "b2_runs"))
    args = ap.parse_args()
    from b200_train.train import CheckpointConfig, RunConfig, ScalingConfig, TorchConfig, TorchTrainer
    big = dict(embed_dim=2048, num_layers=24, num_heads=16, hidden_dim=8192)
    small = dict(embed_dim=256, num_layers=4, num_heads=8, hidden_dim=1024)
    cfg = {"lr": 3e-4, "batch_size_per_worker": 16, "num_epochs": 1, **(small if args.small else big),
           "model_max_seq_len": 1024, "dataset_seq_len": 256, "log_frequency_batches": 20, "warmup_steps_ratio": 0.05,
           "min_lr_ratio": 0.01, "test_run": True, "vocab_size": 283, "n_tokens": 20000, "max_steps": args.steps,
           "bf16": args.bf16, "stock": args.stock, "timing_warmup": min(5, max(0, args.steps - 2))}
    trainer = TorchTrainer(train_loop_per_worker, train_loop_config=cfg,
                           scaling_config=ScalingConfig(num_workers=args.num_workers, use_gpu=True, resources_per_worker={"GPU": 1}),
                           run_config=RunConfig(name="b200_basicllm", storage_path=args.storage,
                                                checkpoint_config=CheckpointConfig(num_to_keep=1, checkpoint_score_attribute="loss",
                                                                                   checkpoint_score_order="min")),
                           torch_config=TorchConfig(backend="nccl"))
    result = trainer.fit()
    print("RESULT", result.metrics, "checkpoint:", result.checkpoint.path if result.checkpoint else None, flush=True)
    if result.error:
        raise result.error


if __name__ == "__main__":
    main()
