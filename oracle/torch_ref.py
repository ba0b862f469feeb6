"""The reference's step-tail CALL SEQUENCE on stock PyTorch — TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs import this. It holds
no kernels of ours: everything here is torch / transformers as the reference invokes them, restated because the
reference scripts cannot be imported on the GPU box (/root/reference is absent there, and `ray`, `trl`, `accelerate`
are not installed anywhere — SURVEY.md §8c).

* `BasicLLMRestated` restates the reference's model class (ref: ray-jobs/pytorch_llm_ray.py:57-105) with identical
  module/parameter names; oracle/make_golden.py loads the reference's own class next to it and checks that a
  shared state_dict gives bit-identical outputs (tests/golden/basicllm_*.npz pin that).
* `basic_llm_step` is the loop body ref :272-279; `hf_causal_lm_step` is the HF Trainer step that
  ref: ray-jobs/fine_tune_llama_ray.py:333 reaches ($SP/transformers/trainer.py:1867-1936, 2493), with the
  hyper-parameters of ref: ray-jobs/fine_tune_config.json:12-29.
* the synthetic-data makers implement SURVEY.md §8(d) (wikitext-2-shaped char stream; HF batches with a padded tail).
"""
import math
import os
import time

import numpy as np
import torch
import torch.nn as nn


# ------------------------------------------------------------------------------------------------ model (restated)
class _PositionalEncoding(nn.Module):
    """ref: ray-jobs/pytorch_llm_ray.py:57-73 — sinusoidal table registered as buffer `pe` [1, max_len, d]."""

    def __init__(self, d_model, max_len=5000):
        super().__init__()
        pos = torch.arange(max_len, dtype=torch.float).unsqueeze(1)
        div = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
        pe = torch.zeros(max_len, d_model)
        pe[:, 0::2] = torch.sin(pos * div)
        pe[:, 1::2] = torch.cos(pos * (div[:-1] if (d_model % 2 and div.size(0) > 1) else div))
        self.register_buffer("pe", pe.unsqueeze(0))

    def forward(self, x):
        return x + self.pe[:, :x.size(1), :]


class BasicLLMRestated(nn.Module):
    """ref: ray-jobs/pytorch_llm_ray.py:75-105. Embedding * sqrt(d) + sinusoidal PE + dropout + TransformerEncoder
    (post-norm, GELU, batch_first, additive causal float mask) + Linear(d, vocab) WITH bias, untied."""

    def __init__(self, vocab_size, embed_dim, num_heads, num_layers, hidden_dim, max_seq_len=1024, dropout=0.1):
        super().__init__()
        self.embed_dim = embed_dim
        self.token_embedding = nn.Embedding(vocab_size, embed_dim)
        self.positional_encoding = _PositionalEncoding(embed_dim, max_seq_len)
        layer = nn.TransformerEncoderLayer(d_model=embed_dim, nhead=num_heads, dim_feedforward=hidden_dim,
                                           dropout=dropout, batch_first=True, activation="gelu")
        self.transformer_decoder = nn.TransformerEncoder(layer, num_layers)
        self.fc_out = nn.Linear(embed_dim, vocab_size)
        self.dropout = nn.Dropout(dropout)
        self.max_seq_len = max_seq_len

    @staticmethod
    def causal_mask(sz, device):
        keep = torch.ones(sz, sz, device=device).tril() == 1
        return torch.zeros(sz, sz, device=device).masked_fill(~keep, float("-inf"))

    def features(self, src):
        """Everything up to (not including) fc_out: the hidden states the LM head consumes."""
        x = self.token_embedding(src) * math.sqrt(self.embed_dim)
        x = self.dropout(self.positional_encoding(x))
        return self.transformer_decoder(x, mask=self.causal_mask(src.size(1), src.device))

    def forward(self, src):
        return self.fc_out(self.features(src))


def disable_dropout(model):
    """Deterministic training passes for parity work: zero every nn.Dropout AND the attention-probability dropout
    that nn.MultiheadAttention keeps as a plain float attribute."""
    for m in model.modules():
        if isinstance(m, nn.Dropout):
            m.p = 0.0
        if isinstance(m, nn.MultiheadAttention):
            m.dropout = 0.0
    return model


SMALL_GPT = dict(embed_dim=256, num_heads=8, num_layers=4, hidden_dim=1024, max_seq_len=1024)   # BASELINE configs[0]
REF_MAIN = dict(embed_dim=2048, num_heads=16, num_layers=24, hidden_dim=8192, max_seq_len=1024)  # ref :324-344
REF_LOOP = dict(lr=3e-4, weight_decay=0.01, max_norm=1.0, batch_size_per_worker=16, dataset_seq_len=256,
                warmup_steps_ratio=0.05, min_lr_ratio=0.01)
HF_LOOP = dict(lr=2e-4, weight_decay=1e-3, max_norm=0.3, grad_accum=4, per_device_batch=2)       # fine_tune_config.json


# ------------------------------------------------------------------------------------------------ synthetic data
def synthetic_char_stream(n_tokens=16_256 + 1, vocab=283, seed=1234):
    """Wikitext-2-shaped char stream: Zipf(s=1) over `vocab` symbols. With test_run the
    reference only ever touches the first 16 000 windows (ref :198-201), i.e. the first 16 256 tokens."""
    g = torch.Generator().manual_seed(seed)
    w = 1.0 / torch.arange(1, vocab + 1, dtype=torch.float64)
    return torch.multinomial(w / w.sum(), n_tokens, replacement=True, generator=g)


def windows(stream, idx, T):
    """TextDataset.__getitem__ (ref :116-119) for a batch of start indices: inputs and PRE-SHIFTED targets."""
    ar = torch.arange(T)
    return stream[idx[:, None] + ar[None, :]], stream[idx[:, None] + ar[None, :] + 1]


def synthetic_hf_batch(B, T, vocab, seed, pad_tail=True, device="cpu"):
    """input_ids uniform over `vocab`; labels = input_ids with a right-padded tail of length U[0, T/4] set to -100
    (what SFT's collator does to padding)."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, vocab, (B, T), generator=g)
    labels = ids.clone()
    if pad_tail:
        tails = torch.randint(0, T // 4 + 1, (B,), generator=g)
        for b in range(B):
            if tails[b]:
                labels[b, T - int(tails[b]):] = -100
    return ids.to(device), labels.to(device)


# ------------------------------------------------------------------------------------------------ the step, stock
def lr_lambda_factory(total_steps, warmup_ratio=0.05, min_lr_ratio=0.01):
    """ref :240-258."""
    warm = int(total_steps * warmup_ratio)
    decay = total_steps - warm

    def fn(step):
        if step < warm:
            return float(step) / float(max(1, warm))
        if step < total_steps:
            prog = float(step - warm) / float(max(1, decay))
            return min_lr_ratio + (1.0 - min_lr_ratio) * 0.5 * (1.0 + math.cos(math.pi * prog))
        return min_lr_ratio
    return fn


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


def hf_param_groups(model, weight_decay):
    """HF Trainer's two groups: decay for everything except biases and norm weights
    ($SP/transformers/trainer.py:1143-1170)."""
    decay, no_decay = [], []
    norm_types = (nn.LayerNorm,)
    norm_param_ids = set()
    for m in model.modules():
        if isinstance(m, norm_types) or m.__class__.__name__.lower().endswith("rmsnorm"):
            for p in m.parameters(recurse=False):
                norm_param_ids.add(id(p))
    seen = set()
    for n, p in model.named_parameters():
        if not p.requires_grad or id(p) in seen:
            continue
        seen.add(id(p))
        (no_decay if (n.endswith("bias") or id(p) in norm_param_ids) else decay).append(p)
    return [{"params": decay, "weight_decay": weight_decay}, {"params": no_decay, "weight_decay": 0.0}]


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


def make_gpt2(dtype=torch.bfloat16, seed=0, **overrides):
    """GPT-2 small (124M) random init from GPT2Config() (BASELINE configs[1]); overrides shrink it for tests."""
    from transformers import GPT2Config, GPT2LMHeadModel
    torch.manual_seed(seed)
    cfg = GPT2Config(**overrides)
    cfg.use_cache = False
    return GPT2LMHeadModel(cfg).to(dtype)


def make_llama(dtype=torch.bfloat16, seed=0, **overrides):
    """Llama-3.2-1B shape by default (BASELINE configs[2]); `tie_word_embeddings=False, hidden_size=4096, ...` gives 8B."""
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    kw = dict(vocab_size=128256, hidden_size=2048, intermediate_size=8192, num_hidden_layers=16,
              num_attention_heads=32, num_key_value_heads=8, head_dim=64, rope_theta=500000.0,
              tie_word_embeddings=True, max_position_embeddings=4096)
    kw.update(overrides)
    cfg = LlamaConfig(**kw)
    cfg.use_cache = False
    return LlamaForCausalLM(cfg).to(dtype)


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


def _gloo_worker(rank, world, port, steps, warmup, threads, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    V = 283
    stream = synthetic_char_stream(vocab=V)
    torch.manual_seed(0)
    model = nn.parallel.DistributedDataParallel(BasicLLMRestated(V, **SMALL_GPT))      # prepare_model, use_gpu=False
    opt = torch.optim.AdamW(model.parameters(), lr=REF_LOOP["lr"], weight_decay=REF_LOOP["weight_decay"])
    crit = nn.CrossEntropyLoss()
    B, T = REF_LOOP["batch_size_per_worker"], REF_LOOP["dataset_seq_len"]
    g = torch.Generator().manual_seed(0)
    perm = torch.randperm(16000, generator=g)[rank::world]            # DistributedSampler(shuffle=True, seed=0)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lr_lambda_factory(len(perm) // B))
    times = []
    for s in range(warmup + steps):
        x, y = windows(stream, perm[s * B:(s + 1) * B], T)
        dist.barrier()
        t0 = time.perf_counter()
        basic_llm_step(model, opt, sched, crit, x, y, V, REF_LOOP["max_norm"])
        dist.barrier()
        if s >= warmup:
            times.append(time.perf_counter() - t0)
    if rank == 0:
        q.put(sum(times) / len(times))
    dist.destroy_process_group()


def cpu_gloo_small_gpt_baseline(steps=5, warmup=2, world=2, threads=None):
    """BASELINE configs[0]: the reference loop (BasicLLM small GPT, synthetic wikitext-2-shaped chars, DDP over gloo,
    num_workers=2, use_gpu=False) timed on host cores. Returns dict(tokens_per_s, ...)."""
    import socket
    import torch.multiprocessing as mp
    threads = threads or max(1, (os.cpu_count() or 2) // world)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    mp.start_processes(_gloo_worker, args=(world, port, steps, warmup, threads, q), nprocs=world, join=True,
                       start_method="spawn")
    sec = q.get()
    B, T = REF_LOOP["batch_size_per_worker"], REF_LOOP["dataset_seq_len"]
    return dict(tokens_per_s=world * B * T / sec, ms_per_step=1e3 * sec, cores=os.cpu_count(), threads=threads * world,
                sample=f"BasicLLM small GPT fp32, {world} gloo workers x {threads} threads, B={B} x T={T} per worker")
