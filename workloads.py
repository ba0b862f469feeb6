"""Workload definitions shared by bench.py (both arms), the tests and examples/: model factories with random
initialisation, and the synthetic inputs of SURVEY.md §8(d). Nothing here computes the hot path — that is either the
product (gke-ray-train_b200/) or the checker (oracle/).

* `BasicLLMRestated` restates the reference's model class (ref: ray-jobs/pytorch_llm_ray.py:57-105) with identical
  module/parameter names; oracle/make_golden.py loads the reference's own class next to it and checks that a shared
  state_dict gives bit-identical outputs (tests/golden/basicllm_*.npz pin that).
* `make_gpt2` / `make_llama` build the HF models of BASELINE configs[1]-[3] from their configs (no checkpoints: there
  is no network); `hf_param_groups` reproduces HF Trainer's decay / no-decay split.
* `synthetic_char_stream`, `windows`, `synthetic_hf_batch` are the seeded synthetic inputs.
"""
import math

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


# ------------------------------------------------------------------------------------------------ schedules, param groups, HF models
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


def make_gpt2(dtype=torch.bfloat16, seed=0, **overrides):
    """GPT-2 small (124M) random init from GPT2Config() (BASELINE configs[1]); overrides shrink it for tests."""
    from transformers import GPT2Config, GPT2LMHeadModel
    torch.manual_seed(seed)
    cfg = GPT2Config(**overrides)
    cfg.use_cache = False
    return GPT2LMHeadModel(cfg).to(dtype)


def make_llama(dtype=torch.bfloat16, seed=0, device=None, **overrides):
    """Llama-3.2-1B shape by default (BASELINE configs[2]); `tie_word_embeddings=False, hidden_size=4096, ...` gives 8B.
    `device`: build (and randomly initialise) the parameters directly there — an 8 B-parameter model takes minutes to
    initialise on host cores and seconds on the GPU."""
    import contextlib
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    kw = dict(vocab_size=128256, hidden_size=2048, intermediate_size=8192, num_hidden_layers=16,
              num_attention_heads=32, num_key_value_heads=8, head_dim=64, rope_theta=500000.0,
              tie_word_embeddings=True, max_position_embeddings=4096)
    kw.update(overrides)
    cfg = LlamaConfig(**kw)
    cfg.use_cache = False
    with (torch.device(device) if device is not None else contextlib.nullcontext()):
        model = LlamaForCausalLM(cfg)
    return model.to(dtype)


LLAMA_8B = dict(hidden_size=4096, intermediate_size=14336, num_hidden_layers=32, num_attention_heads=32,
                num_key_value_heads=8, head_dim=128, tie_word_embeddings=False)      # BASELINE configs[3]
