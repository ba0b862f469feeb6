"""CPU oracle for the data-parallel training-step tail — TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package. The product path (gke-ray-train_b200/) never does, and has no CPU fallback.

What is restated here, in numpy, is the arithmetic of the third-party calls the reference makes on its
step tail. The reference itself (ericehanley/gke-ray-train) holds no arithmetic of its own and NO tests,
golden vectors or fixtures for this path — the algorithm lives in un-vendored
dependencies: torch (pinned 2.3.0 by the reference, ref: a3-mega/gke-ray-cluster-setup.sh:74; 2.11.0 is
what is installed here and what the oracle is checked against) and transformers (pinned 4.50.0; 5.5.0 here).
The oracle is therefore pinned against outputs of the reference's own call sequence run in this container
on stock torch: tests/golden/*.npz, produced by oracle/make_golden.py, which imports BasicLLM / TextDataset
from /root/reference/ray-jobs/pytorch_llm_ray.py and drives the loop body ref :225-279 unchanged.

Each function cites the reference call site (ref:) and the library source whose arithmetic it follows
($SP = site-packages).
"""
import math

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------------------------------
# bf16 helpers (numpy has no bfloat16): values are carried as float32 that are exactly representable.
def bf16_round(x):
    """Round-to-nearest-even float32 -> bfloat16 -> float32 (what torch's .to(bfloat16) does)."""
    x = np.ascontiguousarray(x, dtype=F32)
    u = x.view(np.uint32).astype(np.uint64)
    rounded = ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)
    out = rounded.view(F32).copy()
    nan = np.isnan(x)
    if nan.any():
        out[nan] = np.nan
    return out.reshape(x.shape)


def bf16_ulp(x):
    """Spacing of bfloat16 at |x| (8 significant bits)."""
    x = np.abs(np.asarray(x, dtype=np.float64))
    e = np.floor(np.log2(np.maximum(x, 2.0 ** -126)))
    return 2.0 ** (e - 7)


def f32_ulp(x):
    return np.spacing(np.abs(np.asarray(x, dtype=F32))).astype(np.float64)


# --------------------------------------------------------------------------------------------------
# Cross-entropy
def shift_labels(labels, ignore_index=-100):
    """HF causal-LM shift: labels = pad(labels, (0, 1), value=ignore_index)[..., 1:].
    ref: ray-jobs/fine_tune_llama_ray.py:333 -> $SP/transformers/loss/loss_utils.py:58-61."""
    labels = np.asarray(labels, dtype=np.int64)
    out = np.full_like(labels, ignore_index)
    out[..., :-1] = labels[..., 1:]
    return out


def log_softmax_rows(logits):
    """fp32 log-softmax over the last axis, max-subtracted (ATen host_softmax / vec_log_softmax)."""
    x = np.asarray(logits, dtype=F32)
    m = x.max(axis=-1, keepdims=True)
    z = x - m
    lse = np.log(np.exp(z, dtype=F32).sum(axis=-1, keepdims=True, dtype=F32))
    return z - lse, (m + lse)[..., 0]


def cross_entropy(logits, labels, ignore_index=-100, reduction="mean", num_items=None):
    """nn.CrossEntropyLoss()(logits2d, labels1d) — ref: ray-jobs/pytorch_llm_ray.py:237,275; and the
    fixed_cross_entropy of $SP/transformers/loss/loss_utils.py:28-42 (reduction 'sum' / num_items when
    num_items_in_batch is given). Returns (loss, n_valid, lse[N]).
    Out-of-range labels raise IndexError like torch's CPU kernel; an all-ignored batch gives NaN for
    'mean' and 0 for 'sum'."""
    logits = np.asarray(logits, dtype=F32)
    labels = np.asarray(labels, dtype=np.int64).reshape(-1)
    N, V = logits.shape
    assert labels.shape[0] == N
    valid = labels != ignore_index
    bad = valid & ((labels < 0) | (labels >= V))
    if bad.any():
        raise IndexError(f"Target {int(labels[bad][0])} is out of bounds.")
    logp, lse = log_softmax_rows(logits)
    picked = -logp[np.nonzero(valid)[0], labels[valid]]
    total = picked.sum(dtype=np.float64)
    n_valid = int(valid.sum())
    if num_items is not None:
        loss = total / num_items
    elif reduction == "sum":
        loss = total
    else:
        loss = total / n_valid if n_valid else float("nan")
    return F32(loss), n_valid, lse


def cross_entropy_grad(logits, labels, grad_scale, ignore_index=-100):
    """d loss / d logits for loss = grad_scale * sum_valid(lse - z): (softmax - onehot) * grad_scale on
    non-ignored rows, 0 elsewhere (ATen nll_loss_backward + log_softmax_backward)."""
    logits = np.asarray(logits, dtype=F32)
    labels = np.asarray(labels, dtype=np.int64).reshape(-1)
    logp, _ = log_softmax_rows(logits)
    d = np.exp(logp, dtype=F32)
    valid = labels != ignore_index
    rows = np.nonzero(valid)[0]
    d[rows, labels[valid]] -= 1.0
    d[~valid] = 0.0
    return (d * F32(grad_scale)).astype(F32)


def lmhead_ce(H, W, labels, bias=None, shift_T=0, ignore_index=-100, reduction="mean", num_items=None,
              bf16_logits=True):
    """lm_head + causal-LM loss as the reference's bf16 path computes it:
    logits = (H @ W.T [+ b]) rounded to bf16 ($SP/transformers/models/llama/modeling_llama.py:486-491, a bf16
    nn.Linear), upcast .float() ($SP/transformers/loss/loss_utils.py:55), labels shifted inside the loss when
    shift_T > 0 (:58-61), CE with ignore_index (:28-42). ref: ray-jobs/fine_tune_llama_ray.py:333;
    BasicLLM.fc_out (+bias, fp32, bf16_logits=False, shift_T=0) ref: ray-jobs/pytorch_llm_ray.py:87,104,275.
    H [N, D], W [V, D] given as float32 arrays (bf16-representable for the bf16 path)."""
    H = np.asarray(H, dtype=F32)
    W = np.asarray(W, dtype=F32)
    logits = H @ W.T
    if bias is not None:
        logits = logits + np.asarray(bias, dtype=F32)[None, :]
    if bf16_logits:
        logits = bf16_round(logits)
    labels = np.asarray(labels, dtype=np.int64)
    if shift_T:
        labels = shift_labels(labels.reshape(-1, shift_T), ignore_index)
    labels = labels.reshape(-1)
    loss, n_valid, lse = cross_entropy(logits, labels, ignore_index, reduction, num_items)
    return dict(loss=loss, n_valid=n_valid, lse=lse, logits=logits, labels=labels)


def lmhead_ce_grads(H, W, labels, grad_scale, bias=None, shift_T=0, ignore_index=-100, bf16_logits=True):
    """Backward of lmhead_ce for loss = grad_scale * sum_valid(...): returns dH, dW, db (float32).
    In the bf16 path torch rounds dlogits to bf16 before the two GEMMs (autograd of a bf16 nn.Linear)."""
    r = lmhead_ce(H, W, labels, bias, shift_T, ignore_index, "sum", None, bf16_logits)
    d = cross_entropy_grad(r["logits"], r["labels"], grad_scale, ignore_index)
    if bf16_logits:
        d = bf16_round(d)
    H = np.asarray(H, dtype=F32)
    W = np.asarray(W, dtype=F32)
    return d @ W, d.T @ H, d.sum(axis=0), d


# --------------------------------------------------------------------------------------------------
# clip_grad_norm_
def clip_grad_norm(grads, max_norm):
    """torch.nn.utils.clip_grad_norm_(params, max_norm) — ref: ray-jobs/pytorch_llm_ray.py:277 (1.0);
    ref: ray-jobs/fine_tune_config.json:19 (0.3). $SP/torch/nn/utils/clip_grad.py:91-117,165-183:
    total = || (||g_k||_2)_k ||_2 ; coef = clamp(max_norm / (total + 1e-6), max=1) ; g_k *= coef.
    Returns (total_norm, coef, clipped grads)."""
    norms = np.array([np.sqrt((np.asarray(g, dtype=F32).astype(np.float64) ** 2).sum()) for g in grads])
    total = F32(np.sqrt((norms ** 2).sum()))
    coef = F32(max_norm) / (total + F32(1e-6))
    coef = F32(min(float(coef), 1.0))
    return total, coef, [(np.asarray(g, dtype=F32) * coef).astype(F32) for g in grads]


# --------------------------------------------------------------------------------------------------
# AdamW
def _fma(a, b, c):
    """float32 fused multiply-add emulated through float64 (the product of two float32 is exact there)."""
    return (np.asarray(a, dtype=np.float64) * np.asarray(b, dtype=np.float64) + np.asarray(c, dtype=np.float64)).astype(F32)


def adamw_step(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, step, state_bf16=False,
               use_fma=True):
    """One torch.optim.AdamW step on one tensor — ref: ray-jobs/pytorch_llm_ray.py:236,278.
    Follows $SP/torch/optim/adam.py `_single_tensor_adam` (:347-552) with decoupled weight decay:
      p *= 1 - lr*wd ; m.lerp_(g, 1-b1) ; v.mul_(b2).addcmul_(g, g, value=1-b2) ;
      denom = sqrt(v)/sqrt(1-b2^t) + eps ; p.addcdiv_(m, denom, value=-(lr/(1-b1^t))).
    Scalars are formed in Python doubles and narrowed to float32 inside the kernels, as torch does.
    state_bf16=True models the reference's pure-bf16 full fine-tune: p, g, m, v are bf16 and
    every in-place op rounds its result to bf16. Returns new (p, m, v) as float32 arrays."""
    rnd = bf16_round if state_bf16 else (lambda x: np.asarray(x, dtype=F32))
    p = np.asarray(p, dtype=F32); g = np.asarray(g, dtype=F32); m = np.asarray(m, dtype=F32); v = np.asarray(v, dtype=F32)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = lr / bc1
    bc2_sqrt = math.sqrt(bc2)
    w1 = F32(1.0 - beta1)
    w2 = F32(1.0 - beta2)
    p = rnd(p * F32(1.0 - lr * weight_decay))
    m = rnd(_fma(w1, g - m, m) if use_fma else m + w1 * (g - m))
    v = rnd(v * F32(beta2))
    v = rnd(_fma(w2, (g * g).astype(F32), v) if use_fma else v + w2 * (g * g))
    denom = rnd(np.sqrt(v, dtype=F32))
    denom = rnd(denom / F32(bc2_sqrt))
    denom = rnd(denom + F32(eps))
    q = (m / denom).astype(F32)
    p = rnd(_fma(F32(-step_size), q, p) if use_fma else p + F32(-step_size) * q)
    return p, m, v


# --------------------------------------------------------------------------------------------------
# DDP gradient averaging
def ddp_average(grads_per_rank):
    """What DistributedDataParallel leaves in .grad after backward: each rank's bucket is divided by the world
    size as it is copied in, then summed across ranks (c10d Reducer::mark_variable_ready_dense + allreduce;
    installed by prepare_model, ref: ray-jobs/pytorch_llm_ray.py:230,276).
    grads_per_rank: list over ranks of float32 arrays of equal shape."""
    W = len(grads_per_rank)
    acc = np.zeros_like(np.asarray(grads_per_rank[0], dtype=F32))
    for g in grads_per_rank:
        acc = acc + (np.asarray(g, dtype=F32) / F32(W))
    return acc


# --------------------------------------------------------------------------------------------------
# learning-rate schedule (host scalar)
def lr_lambda(step, total_steps, warmup_ratio=0.05, min_lr_ratio=0.01):
    """ref: ray-jobs/pytorch_llm_ray.py:240-258 — linear warm-up then cosine decay to a floor."""
    warm = int(total_steps * warmup_ratio)
    decay = total_steps - warm
    if step < warm:
        return float(step) / float(max(1, warm))
    if step < total_steps:
        prog = float(step - warm) / float(max(1, decay))
        return min_lr_ratio + (1.0 - min_lr_ratio) * 0.5 * (1.0 + math.cos(math.pi * prog))
    return min_lr_ratio
