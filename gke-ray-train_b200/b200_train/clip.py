"""b200_train.clip_grad_norm_ — drop-in for torch.nn.utils.clip_grad_norm_.

ref: ray-jobs/pytorch_llm_ray.py:277 `torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0)`; HF Trainer
with MAX_GRAD_NORM 0.3 (ref: ray-jobs/fine_tune_config.json:19). Semantics: $SP/torch/nn/utils/clip_grad.py:50-183.

One sum-of-squares launch per flat gradient buffer, one scalar kernel for the coefficient, and — when the
parameters belong to a b200_train.AdamW — NO pass to scale the gradients: the coefficient stays on the device and
the optimizer kernel multiplies by it as it reads each gradient. Otherwise the gradients are scaled in place.
Returns the total norm as a 0-dim device tensor, like torch (no host sync).

Deviation from torch, and how to switch it off: with a b200_train.AdamW attached, `.grad` still holds the UNCLIPPED
values after this call (torch scales them in place, $SP/torch/nn/utils/clip_grad.py:165-183); only the optimizer step sees
the clipped ones. Code that reads or logs gradients between clip and step should pass `defer_to_optimizer=False` (one
extra read+write pass over the gradients, exactly torch's behaviour). A second deferred call before step() composes with
the first one the way two in-place clips would: it is given the norm of the already-clipped gradients.
"""
import torch

from . import ops
from ._lib import B2Error
from .flat import buffers_of, flat_of


@torch.no_grad()
def clip_grad_norm_(parameters, max_norm, norm_type=2.0, error_if_nonfinite=False, foreach=None, *,
                    defer_to_optimizer=True):
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    params = [p for p in parameters if p.grad is not None]
    if float(norm_type) != 2.0:
        raise B2Error("b200_train.clip_grad_norm_ implements the 2-norm the reference uses")
    if not params:
        return torch.tensor(0.0)
    device = params[0].grad.device
    # fast path (the usual call: every parameter of the model, all of them living in the same flat buffers, in buffer
    # order): a few identity checks instead of rebuilding the parameter -> buffer map on every step
    f0 = flat_of(params[0])
    bufs = loose = None
    if f0 is not None and len(params) == len(f0[0].params) and all(a is b for a, b in zip(params, f0[0].params)):
        bufs, loose, whole_all = [f0[0]], [], True
    else:
        bufs, loose = buffers_of(params)
        whole_all = False
    covered = None if whole_all else {id(p) for p in params}
    pieces = []
    for buf in bufs:
        whole = whole_all or all(id(p) in covered for p in buf.params)
        reduced = getattr(buf, "reduced_sumsq", None)
        if whole and reduced is not None:
            pieces.append(reduced)          # already produced by the all-reduce kernel's epilogue: no extra pass
            continue
        part = torch.empty(1, dtype=torch.float32, device=device)
        if whole:
            buf.install_grads()
            ops.sumsq(buf.grad, out=part)                       # padding is zero: the whole buffer in one launch
        else:                                                   # a subset of the buffer's parameters
            first = True
            for i, p in enumerate(buf.params):
                if id(p) in covered:
                    ops.sumsq(buf.grad_view(i).reshape(-1), out=part, accumulate=not first)
                    first = False
        pieces.append(part)
    if loose:
        part = torch.empty(1, dtype=torch.float32, device=device)
        first = True
        for p in loose:
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            ops.sumsq(g.reshape(-1), out=part, accumulate=not first)
            first = False
        pieces.append(part)
    parts = pieces[0] if len(pieces) == 1 else torch.cat(pieces)
    # an earlier clip whose coefficient is still waiting for the optimizer: these gradients count as already scaled by it
    earlier = None
    for buf in bufs:
        if buf.pending_coef is not None:
            earlier = buf.pending_coef
            break
    if earlier is not None:
        parts = parts * (earlier * earlier)
    norm, coef = ops.clip_coef(parts, max_norm)
    if earlier is not None:
        coef = coef * earlier
    if error_if_nonfinite and not bool(torch.isfinite(norm).item()):
        raise RuntimeError(f"The total norm of order {norm_type} for gradients from `parameters` is non-finite, "
                           "so it cannot be clipped.")
    for buf in bufs:
        whole = whole_all or all(id(p) in covered for p in buf.params)
        if whole and buf.optimizer_attached and defer_to_optimizer:
            buf.pending_coef = coef                             # consumed by the next AdamW.step()
        elif whole:
            ops.scale_(buf.grad, coef)
            buf.pending_coef = None                             # (an earlier deferred coefficient is part of `coef`)
        else:
            for i, p in enumerate(buf.params):
                if id(p) in covered:
                    ops.scale_(buf.grad_view(i).reshape(-1), coef)
    for p in loose:
        if p.grad.is_contiguous():
            ops.scale_(p.grad.view(-1), coef)
        else:
            g = p.grad.contiguous()
            ops.scale_(g.view(-1), coef)
            p.grad.copy_(g)
    return norm.reshape(())
