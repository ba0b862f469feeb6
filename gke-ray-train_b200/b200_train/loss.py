"""Loss side of the step tail.

* `CrossEntropyLoss` — drop-in for `nn.CrossEntropyLoss()` as used at ref: ray-jobs/pytorch_llm_ray.py:237,275
  (materialised fp32 logits `[N, V]`, class-index targets, mean reduction, ignore_index -100).
* `fused_lm_head_ce` / `FusedLMHeadCE` / `patch_causal_lm` — the LM head and the causal-LM loss in one tcgen05
  kernel for bf16 models, replacing `lm_head(hidden)` + `ForCausalLMLoss` that `trainer.train()` reaches
  (ref: ray-jobs/fine_tune_llama_ray.py:333 -> $SP/transformers/models/llama/modeling_llama.py:486-491 ->
  $SP/transformers/loss/loss_utils.py:45-67).
"""
import os

import torch
import torch.nn as nn

from . import ops
from ._lib import B2Error
from .flat import flat_of

DEFAULT_CHUNK_ROWS = 8192
# Keep the forward's bf16 logits for the backward when they fit one chunk (B2_KEEP_LOGITS=0: always recompute them;
# B2_KEEP_LOGITS_MAX_MB bounds the buffer, default 4096 MiB — it replaces the d(logits) chunk buffer of the same size,
# but lives from the head's forward to its backward instead of only during the backward).
_KEEP_LOGITS = os.environ.get("B2_KEEP_LOGITS", "1") != "0"
_KEEP_LOGITS_MAX_BYTES = int(float(os.environ.get("B2_KEEP_LOGITS_MAX_MB", "4096")) * (1 << 20))


def _gscale(grad_out, result, reduction, num_items):
    """Device scalar multiplying (softmax - onehot): grad_output / denominator. No host sync."""
    g = grad_out.detach().to(torch.float32).reshape(1)
    if num_items is not None:
        if isinstance(num_items, torch.Tensor):
            return g / num_items.to(g.device, torch.float32).reshape(1)
        return g / float(num_items)
    if reduction == "sum":
        return g.clone()
    return g / result.n_valid.to(torch.float32)


def _pick_loss(result, reduction, num_items):
    if num_items is not None:
        if isinstance(num_items, torch.Tensor):
            return result.loss[0] / num_items.to(result.loss.device, torch.float32)
        return result.loss[0] / float(num_items)
    return result.loss[0] if reduction == "sum" else result.loss[1]


class _CERowsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index, reduction, check):
        x = logits if logits.stride(-1) == 1 else logits.contiguous()
        r = ops.ce_rows_fwd(x, labels, 0, ignore_index)
        if check and int(r.err.item()):
            raise IndexError("Target out of bounds for cross_entropy (label outside [0, V) and != ignore_index)")
        ctx.save_for_backward(x, labels, r.lse, r.n_valid)
        ctx.ignore_index, ctx.reduction = ignore_index, reduction
        ctx.result = r
        return _pick_loss(r, reduction, None).clone()

    @staticmethod
    def backward(ctx, grad_out):
        x, labels, lse, n_valid = ctx.saved_tensors
        gs = _gscale(grad_out, ctx.result, ctx.reduction, None)
        d = ops.ce_rows_bwd(x, labels, lse, gs, 0, ctx.ignore_index)
        return d, None, None, None, None


class CrossEntropyLoss(nn.Module):
    """`criterion = CrossEntropyLoss(); loss = criterion(outputs.reshape(-1, V), targets.reshape(-1))`.

    check_labels=True makes an out-of-range target raise IndexError like torch (this costs one host sync per
    call); with False the loss comes back NaN and `last_result.err` is set instead."""

    def __init__(self, weight=None, size_average=None, ignore_index=-100, reduce=None, reduction="mean",
                 label_smoothing=0.0, check_labels=False):
        super().__init__()
        if weight is not None or label_smoothing != 0.0 or size_average is not None or reduce is not None:
            raise B2Error("b200_train.CrossEntropyLoss implements what the reference uses: no class weights, "
                          "no label smoothing, reduction 'mean' or 'sum'")
        if reduction not in ("mean", "sum"):
            raise B2Error("reduction must be 'mean' or 'sum'")
        self.ignore_index, self.reduction, self.check_labels = ignore_index, reduction, check_labels

    def forward(self, input, target):
        if input.dim() != 2 or target.dim() != 1:
            raise B2Error("expected logits [N, V] and targets [N] (the reference reshapes to exactly this)")
        return _CERowsFn.apply(input, target.contiguous(), self.ignore_index, self.reduction, self.check_labels)


class _LMHeadCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hidden, weight, bias, labels, shift_T, ignore_index, reduction, num_items, round_bf16, chunk_rows,
                check_labels=False, weight_shared=False):
        H = hidden if hidden.stride(-1) == 1 else hidden.contiguous()
        bias32 = None if bias is None else bias.detach().to(torch.float32).contiguous()
        # One chunk of rows and a gradient wanted: let the forward leave its bf16 logits in the buffer the backward would
        # otherwise fill by a second pass over the tensor cores; the backward then rewrites them in place (HBM-bound).
        N, V = H.shape[0], weight.shape[0]
        ldd = (V + 63) // 64 * 64          # row pitch: a multiple of 128 bytes
        kept = None
        if (_KEEP_LOGITS and round_bf16 and 0 < N <= chunk_rows and any(ctx.needs_input_grad[:3])
                and N * ldd * 2 <= _KEEP_LOGITS_MAX_BYTES):
            kept = torch.empty(N, ldd, dtype=torch.bfloat16, device=H.device)
        r = ops.lmhead_ce_fwd(H, weight, labels, bias=bias32, shift_T=shift_T, ignore_index=ignore_index,
                              round_logits_bf16=round_bf16, keep_logits=kept)
        ctx.kept = kept
        if check_labels and int(r.err.item()):          # one host sync; torch raises here, by default we poison the loss
            raise IndexError("Target out of bounds for the LM-head loss (label outside [0, V) and != ignore_index)")
        ctx.save_for_backward(H, weight, labels, r.lse, r.n_valid)
        ctx.bias32, ctx.has_bias = bias32, bias is not None
        ctx.bias_dtype = None if bias is None else bias.dtype
        ctx.cfg = (shift_T, ignore_index, reduction, num_items, round_bf16, chunk_rows)
        # where the weight's gradient can be written without passing through a fresh [V, D] tensor (see backward)
        ctx.w_param = weight if (flat_of(weight) is not None and weight.dtype == torch.bfloat16) else None
        ctx.w_shared = bool(weight_shared)
        ctx.result = r
        return _pick_loss(r, reduction, num_items).clone()

    @staticmethod
    def backward(ctx, grad_out):
        H, W, labels, lse, n_valid = ctx.saved_tensors
        shift_T, ignore_index, reduction, num_items, round_bf16, chunk_rows = ctx.cfg
        N, D = H.shape
        V = W.shape[0]
        gs = _gscale(grad_out, ctx.result, reduction, num_items)
        ldd = (V + 63) // 64 * 64          # row pitch: a multiple of 128 bytes
        chunk = min(N, chunk_rows)
        kept, ctx.kept = ctx.kept, None
        if kept is None and getattr(ctx, "kept_used", False):
            raise B2Error("this LM-head loss was already back-propagated once: its kept logits were turned into the "
                          "gradient in place. Set B2_KEEP_LOGITS=0 to back-propagate the same graph twice")
        ctx.kept_used = kept is not None
        dl = kept if kept is not None else torch.empty(chunk, ldd, dtype=torch.bfloat16, device=H.device)
        need_dH, need_dW = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        dH = torch.empty(N, D, dtype=torch.bfloat16, device=H.device) if need_dH else None
        n_chunks = (N + chunk - 1) // chunk
        dW = None
        # The weight lives in a flat buffer (b200_train.AdamW / prepare_model): its gradient goes STRAIGHT into the
        # parameter's slice of the flat gradient buffer instead of a fresh [V, D] tensor that autograd would then add or
        # copy there (two to five extra passes over a 77 MB - 1 GB matrix).
        #  * weight shared with the input embedding (GPT-2, Llama-3.2-1B; ref: fine_tune_llama_ray.py:230-241 loads such
        #    models): the GEMM writes (first micro-batch) or accumulates (later ones) into the slice, `.grad` is pointed
        #    at it, and autograd gets no gradient from here — the embedding's own gradient is then added IN PLACE by
        #    AccumulateGrad, which also fires the parameter's hooks (bucket ready, optimizer bookkeeping).
        #  * weight used here only: while the parameter has no gradient yet, the GEMM writes into the slice and autograd
        #    is handed a fresh alias of it, which AccumulateGrad adopts without a copy; once a gradient exists (later
        #    micro-batches) the ordinary path runs.
        direct = None
        if need_dW and ctx.w_param is not None and W.shape == ctx.w_param.shape:
            f = flat_of(ctx.w_param)
            if f is not None:
                f[0].flush_landing()             # (a gradient of this parameter still waiting to be copied in lands first)
                view = f[0].grad_view(f[0]._index[id(ctx.w_param)])
                have = ctx.w_param.grad is not None and ctx.w_param.grad.data_ptr() == view.data_ptr()
                if ctx.w_shared or not have:
                    direct = (view, have)
        if need_dW:
            # one chunk: write bf16 directly; several: accumulate in fp32 and round once at the end, like the
            # single fp32-accumulated GEMM of the reference's autograd
            if direct is not None and n_chunks == 1:
                dW = direct[0]
            else:
                dW = torch.empty(V, D, dtype=torch.bfloat16 if n_chunks == 1 else torch.float32, device=H.device)
        db = torch.empty(V, dtype=torch.float32, device=H.device) if ctx.has_bias else None
        for c in range(n_chunks):
            r0 = c * chunk
            rows = min(chunk, N - r0)
            # one call per chunk: d(logits) (in place from the kept logits, or recomputed), dH, dW, db
            # dH[rows, D] = dL[rows, V] @ W[V, D];  dW[V, D] += dL^T[V, rows] @ H[rows, D];  db[V] += column sums of dL
            ops.lmhead_ce_bwd(H, W, labels, lse, gs, r0, rows, dl, kept=kept is not None,
                              dH=dH[r0:r0 + rows] if need_dH else None, dW=dW if need_dW else None,
                              dW_accumulate=c > 0 or (direct is not None and dW is direct[0] and direct[1]),
                              db=db, db_accumulate=c > 0, bias=ctx.bias32, shift_T=shift_T, ignore_index=ignore_index,
                              round_logits_bf16=round_bf16)
        if direct is not None:
            view, have = direct
            if dW is not view:                       # several chunks: the fp32 sum is rounded once, into the slice
                if have:
                    view.add_(dW)
                else:
                    view.copy_(dW)
            if ctx.w_shared:
                ctx.w_param.grad = view
                dW = None
            else:
                dW = view.view(view.shape)           # a fresh alias nobody else holds: AccumulateGrad keeps it as .grad
        elif dW is not None and dW.dtype != W.dtype:
            dW = dW.to(W.dtype)
        if db is not None:
            db = db.to(ctx.bias_dtype)
        return dH, dW, db, None, None, None, None, None, None, None, None, None


def fused_lm_head_ce(hidden, weight, labels, bias=None, *, shift=True, ignore_index=-100, reduction="mean",
                     num_items_in_batch=None, round_logits_bf16=True, chunk_rows=DEFAULT_CHUNK_ROWS, check_labels=False,
                     weight_shared=False):
    """loss = CE(hidden @ weight.T (+bias), labels) without the fp32 [N, V] logits / log-probability tensors of the stock path
    (at most a bf16 copy of the logits lives from this forward to its backward, see _KEEP_LOGITS above).

    hidden [..., T, D] bf16, weight [V, D] bf16, labels [..., T] int64. shift=True applies the causal shift inside
    (row t is scored against labels[t+1], the last position ignored) exactly as HF's ForCausalLMLoss; shift=False
    expects pre-shifted labels (the reference's TextDataset, ref: ray-jobs/pytorch_llm_ray.py:116-119).
    num_items_in_batch: HF gradient-accumulation normaliser — the loss is sum / num_items_in_batch.
    check_labels=True raises IndexError on an out-of-range label like torch (one host sync); otherwise the loss is NaN.
    weight_shared=True promises that `weight` ALSO receives a gradient from elsewhere in the same graph (a tied input
    embedding): the head's share is then accumulated directly into the parameter's flat gradient slice (see backward)."""
    if hidden.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
        raise B2Error("fused_lm_head_ce takes bfloat16 hidden states and weight (use CrossEntropyLoss for fp32 logits)")
    T = hidden.shape[-2] if hidden.dim() >= 2 else 1
    H2 = hidden.reshape(-1, hidden.shape[-1])
    y = labels.reshape(-1).contiguous()
    if y.device != H2.device:
        y = y.to(H2.device)
    return _LMHeadCEFn.apply(H2, weight, bias, y, T if shift else 0, ignore_index, reduction, num_items_in_batch,
                             round_logits_bf16, chunk_rows, check_labels, weight_shared)


class FusedLMHeadCE(nn.Module):
    """Module form: owns nothing, borrows the LM head's weight (tied weights keep working)."""

    def __init__(self, lm_head, shift=True, ignore_index=-100, reduction="mean", round_logits_bf16=True,
                 chunk_rows=DEFAULT_CHUNK_ROWS, weight_shared=False):
        super().__init__()
        self.lm_head = lm_head
        self.weight_shared = weight_shared
        self.shift, self.ignore_index, self.reduction = shift, ignore_index, reduction
        self.round_logits_bf16, self.chunk_rows = round_logits_bf16, chunk_rows

    def forward(self, hidden, labels, num_items_in_batch=None, weight_shared=None):
        # The head's weight is read here without going through lm_head.forward: if a data-parallel wrapper is still
        # all-gathering updated parameters under this forward pass, wait for the range that holds it.
        f = flat_of(self.lm_head.weight)
        gate = getattr(f[0], "param_gate", None) if f is not None else None
        if gate is not None:
            gate(f[1])
        return fused_lm_head_ce(hidden, self.lm_head.weight, labels, getattr(self.lm_head, "bias", None),
                                shift=self.shift, ignore_index=self.ignore_index, reduction=self.reduction,
                                num_items_in_batch=num_items_in_batch, round_logits_bf16=self.round_logits_bf16,
                                chunk_rows=self.chunk_rows,
                                weight_shared=self.weight_shared if weight_shared is None else weight_shared)


def patch_causal_lm(model, chunk_rows=DEFAULT_CHUNK_ROWS):
    """Re-routes `model(input_ids, labels=...)` of a HF *ForCausalLM / GPT2LMHeadModel through the fused kernel:
    the backbone runs as before, the logits are never formed, and the returned output carries `loss` (logits=None).
    Calls without labels are untouched. Returns the same model object."""
    from transformers.modeling_outputs import CausalLMOutputWithPast

    head = model.get_output_embeddings()
    backbone = model.base_model
    if head is None or backbone is model:
        raise B2Error("patch_causal_lm: model has no separate backbone / output embedding")
    emb = model.get_input_embeddings()
    tied = emb is not None and getattr(emb, "weight", None) is head.weight
    fused = FusedLMHeadCE(head, shift=True, chunk_rows=chunk_rows)
    orig_forward = model.forward

    def forward(input_ids=None, labels=None, num_items_in_batch=None, **kw):
        if labels is None or head.weight.dtype != torch.bfloat16:
            if num_items_in_batch is not None:
                kw["num_items_in_batch"] = num_items_in_batch
            return orig_forward(input_ids=input_ids, labels=labels, **kw)
        kw.pop("return_dict", None)
        kw.pop("logits_to_keep", None)
        out = backbone(input_ids=input_ids, **kw)
        hidden = out[0]
        # tied head: the embedding lookup of THIS forward also back-propagates into the weight (only when token ids, not
        # ready-made embeddings, came in)
        loss = fused(hidden, labels, num_items_in_batch=num_items_in_batch,
                     weight_shared=tied and input_ids is not None and emb.weight.requires_grad)
        return CausalLMOutputWithPast(loss=loss, logits=None, past_key_values=getattr(out, "past_key_values", None),
                                      hidden_states=getattr(out, "hidden_states", None),
                                      attentions=getattr(out, "attentions", None))

    model.forward = forward
    # kept out of the module tree on purpose: registering it would add `_b2_fused_head.*` keys to model.state_dict()
    # and break checkpoint compatibility with the unpatched model (ref: pytorch_llm_ray.py:301 saves state_dict())
    object.__setattr__(model, "_b2_fused_head", fused)
    return model
