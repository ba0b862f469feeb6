"""Tensor-level wrappers over the C ABI (include/b2train.h). PyTorch is used for device memory and
streams only; every function here enqueues hand-written sm_100a kernels and nothing else.
"""
import torch

from . import _lib
from ._lib import B2Error, call, dtype_code, ptr, require_cuda, stream_ptr

_scratch = {}


def _sumsq_scratch(device):
    """Zero-initialised scratch for b2_sumsq, one per (device, stream) so overlapping calls never share it."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None:
        buf = torch.zeros(_lib.lib.b2_sumsq_scratch_bytes() // 4, dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


def _flat(t, what):
    if not t.is_contiguous():
        raise B2Error(f"{what} must be contiguous")
    return t


def adamw_(p, g, m, v, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, step, gscale=None):
    """In-place fused AdamW on flat (contiguous) tensors. `step` is the 1-based step count."""
    require_cuda(p, g, m, v, gscale)
    n = p.numel()
    if not (g.numel() == n and m.numel() == n and v.numel() == n):
        raise B2Error("adamw_: p, g, m, v must have the same number of elements")
    if m.dtype != v.dtype:
        raise B2Error("adamw_: exp_avg and exp_avg_sq must share a dtype")
    for t, w in ((p, "p"), (g, "g"), (m, "m"), (v, "v")):
        _flat(t, w)
    with torch.cuda.device(p.device):
        call("b2_adamw", ptr(p), ptr(g), ptr(m), ptr(v), n, dtype_code(p.dtype), dtype_code(g.dtype),
             dtype_code(m.dtype), float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), int(step),
             ptr(gscale), stream_ptr())


def adamw_grouped_(p, g, m, v, chunk_group, groups, gscale=None):
    """Fused AdamW over a flat buffer holding several param_groups. `chunk_group`: uint8 device tensor, one entry
    per 64 elements; `groups`: list of dicts(lr, beta1, beta2, eps, weight_decay, step)."""
    import ctypes
    require_cuda(p, g, m, v, chunk_group, gscale)
    n = p.numel()
    if chunk_group.dtype != torch.uint8 or chunk_group.numel() * 64 != n:
        raise B2Error("adamw_grouped_: chunk_group must be uint8 with one entry per 64 elements")
    arr = (_lib.AdamWGroup * len(groups))()
    for i, h in enumerate(groups):
        arr[i] = _lib.AdamWGroup(float(h["lr"]), float(h["beta1"]), float(h["beta2"]), float(h["eps"]),
                                 float(h["weight_decay"]), int(h["step"]))
    with torch.cuda.device(p.device):
        call("b2_adamw_grouped", ptr(p), ptr(g), ptr(m), ptr(v), n, dtype_code(p.dtype), dtype_code(g.dtype),
             dtype_code(m.dtype), ptr(chunk_group), ctypes.addressof(arr), len(groups), ptr(gscale), stream_ptr())


def sumsq(g, out=None, accumulate=False):
    """Sum of squares of a flat tensor -> 1-element fp32 device tensor (no host sync)."""
    require_cuda(g)
    _flat(g, "g")
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=g.device)
        accumulate = False
    with torch.cuda.device(g.device):
        call("b2_sumsq", ptr(g), g.numel(), dtype_code(g.dtype), ptr(_sumsq_scratch(g.device)), ptr(out),
             int(bool(accumulate)), stream_ptr())
    return out


def clip_coef(sumsq_parts, max_norm):
    """(total_norm, coef) device scalars from a fp32 vector of sum-of-squares partials."""
    require_cuda(sumsq_parts)
    if sumsq_parts.dtype != torch.float32:
        raise B2Error("clip_coef: partials must be float32")
    norm = torch.empty(1, dtype=torch.float32, device=sumsq_parts.device)
    coef = torch.empty(1, dtype=torch.float32, device=sumsq_parts.device)
    with torch.cuda.device(sumsq_parts.device):
        call("b2_clip_coef", ptr(sumsq_parts), sumsq_parts.numel(), float(max_norm), ptr(norm), ptr(coef), stream_ptr())
    return norm, coef


def scale_(g, coef):
    require_cuda(g, coef)
    _flat(g, "g")
    with torch.cuda.device(g.device):
        call("b2_scale", ptr(g), g.numel(), dtype_code(g.dtype), ptr(coef), stream_ptr())
    return g


class CEResult:
    """Device-side results of a cross-entropy forward: nothing here forces a host sync."""
    __slots__ = ("lse", "zt", "loss", "n_valid", "err")

    def __init__(self, lse, zt, loss, n_valid, err):
        self.lse, self.zt, self.loss, self.n_valid, self.err = lse, zt, loss, n_valid, err

    @property
    def loss_sum(self):
        return self.loss[0]

    @property
    def loss_mean(self):
        return self.loss[1]


def _ce_outputs(N, device):
    lse = torch.empty(max(N, 1), dtype=torch.float32, device=device)
    zt = torch.zeros(max(N, 1), dtype=torch.float32, device=device)
    loss = torch.empty(2, dtype=torch.float32, device=device)
    n_valid = torch.empty(1, dtype=torch.int64, device=device)
    err = torch.empty(1, dtype=torch.int32, device=device)
    return lse, zt, loss, n_valid, err


def _check_labels(labels, N):
    if labels.dtype != torch.int64:
        raise B2Error("labels must be int64")
    if labels.numel() != N:
        raise B2Error(f"labels has {labels.numel()} elements, expected {N}")
    return _flat(labels, "labels")


def ce_rows_fwd(logits, labels, shift_T=0, ignore_index=-100):
    """Cross-entropy over materialised logits [N, V] (row stride may exceed V)."""
    require_cuda(logits, labels)
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise B2Error("ce_rows_fwd: logits must be [N, V] with unit column stride")
    N, V = logits.shape
    ld = logits.stride(0) if N > 1 else max(V, logits.stride(0))
    _check_labels(labels, N)
    lse, zt, loss, n_valid, err = _ce_outputs(N, logits.device)
    with torch.cuda.device(logits.device):
        call("b2_ce_rows_fwd", ptr(logits), dtype_code(logits.dtype), N, V, ld, ptr(labels), int(shift_T),
             int(ignore_index), ptr(lse), ptr(zt), ptr(loss), ptr(n_valid), ptr(err), stream_ptr())
    return CEResult(lse, zt, loss, n_valid, err)


def ce_rows_bwd(logits, labels, lse, gscale, shift_T=0, ignore_index=-100, out=None):
    require_cuda(logits, labels, lse, gscale)
    N, V = logits.shape
    ld = logits.stride(0) if N > 1 else max(V, logits.stride(0))
    if out is None:
        out = torch.empty_strided(logits.shape, logits.stride(), dtype=logits.dtype, device=logits.device)
    elif out.stride() != logits.stride() or out.dtype != logits.dtype:
        raise B2Error("ce_rows_bwd: out must match logits' dtype and strides")
    with torch.cuda.device(logits.device):
        call("b2_ce_rows_bwd", ptr(logits), dtype_code(logits.dtype), N, V, ld, ptr(labels), int(shift_T),
             int(ignore_index), ptr(lse), ptr(gscale), ptr(out), stream_ptr())
    return out


def _check_hw(H, W, bias):
    require_cuda(H, W, bias)
    if H.dtype != torch.bfloat16 or W.dtype != torch.bfloat16:
        raise B2Error("the fused LM-head kernel takes bfloat16 hidden states and weight")
    if H.dim() != 2 or W.dim() != 2 or H.stride(1) != 1 or W.stride(1) != 1 or H.shape[1] != W.shape[1]:
        raise B2Error("expected H [N, D] and W [V, D] with unit inner stride")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != W.shape[0] or not bias.is_contiguous()):
        raise B2Error("bias must be a contiguous float32 [V] tensor")


def lmhead_ce_fwd(H, W, labels, bias=None, shift_T=0, ignore_index=-100, round_logits_bf16=True, keep_logits=None):
    """Fused logits = H @ W.T (+bias) and cross-entropy. The logits are not materialised, unless `keep_logits`
    (bf16 [N, ld >= V], ld % 8 == 0) is given: then the bf16-rounded logits are also left there, for
    ce_dlogits_inplace_ (a backward without a second tensor-core pass)."""
    _check_hw(H, W, bias)
    N, D = H.shape
    V = W.shape[0]
    _check_labels(labels, N)
    lse, zt, loss, n_valid, err = _ce_outputs(N, H.device)
    with torch.cuda.device(H.device):
        ws_bytes = _lib.lib.b2_lmhead_ce_workspace_bytes(N)
        ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=H.device)
        if keep_logits is None:
            call("b2_lmhead_ce_fwd", ptr(H), H.stride(0), ptr(W), W.stride(0), ptr(bias), ptr(labels), N, D, V,
                 int(shift_T), int(ignore_index), int(bool(round_logits_bf16)), ptr(ws), ws_bytes, ptr(lse), ptr(zt),
                 ptr(loss), ptr(n_valid), ptr(err), stream_ptr())
        else:
            if not round_logits_bf16:
                raise B2Error("keep_logits stores bf16 logits: it needs round_logits_bf16=True")
            _check_logits_buffer(keep_logits, N, V, H.device)
            call("b2_lmhead_ce_fwd_keep", ptr(H), H.stride(0), ptr(W), W.stride(0), ptr(bias), ptr(labels), N, D, V,
                 int(shift_T), int(ignore_index), ptr(ws), ws_bytes, ptr(lse), ptr(zt), ptr(loss), ptr(n_valid),
                 ptr(err), ptr(keep_logits), keep_logits.stride(0), stream_ptr())
    return CEResult(lse, zt, loss, n_valid, err)


def _check_logits_buffer(t, rows, V, device):
    if (t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1 or t.shape[0] < rows or t.shape[1] < V
            or t.stride(0) % 8 or t.device != device):
        raise B2Error("logits buffer must be bfloat16 [>= rows, >= V] on the same device, unit inner stride, "
                      "row pitch a multiple of 8")


def ce_dlogits_inplace_(logits, labels, lse, gscale, row0, rows, V, shift_T=0, ignore_index=-100):
    """Turns the kept bf16 logits of rows [row0, row0+rows) (`logits` starts at row row0) into d(logits), in place —
    the same values lmhead_ce_bwd_dlogits would write."""
    require_cuda(logits, labels, lse, gscale)
    _check_logits_buffer(logits, rows, V, logits.device)
    with torch.cuda.device(logits.device):
        call("b2_ce_dlogits_inplace", ptr(logits), logits.stride(0), ptr(labels), int(row0), int(rows), int(V),
             int(shift_T), int(ignore_index), ptr(lse), ptr(gscale), stream_ptr())
    return logits


def lmhead_ce_bwd_dlogits(H, W, labels, lse, gscale, row0, rows, out, bias=None, shift_T=0, ignore_index=-100,
                          round_logits_bf16=True):
    """d(logits) for rows [row0, row0+rows) into `out` (bf16 [>=rows, ldd])."""
    _check_hw(H, W, bias)
    if out.dtype != torch.bfloat16 or out.stride(1) != 1:
        raise B2Error("dlogits buffer must be bfloat16 with unit inner stride")
    D, V = H.shape[1], W.shape[0]
    Hc = H[row0:row0 + rows]
    with torch.cuda.device(H.device):
        call("b2_lmhead_ce_bwd_dlogits", ptr(Hc), H.stride(0), ptr(W), W.stride(0), ptr(bias), ptr(labels), int(row0),
             int(rows), D, V, int(shift_T), int(ignore_index), int(bool(round_logits_bf16)), ptr(lse), ptr(gscale),
             ptr(out), out.stride(0), stream_ptr())
    return out


def gemm_bf16(A, B, C, M, N, K, a_mn_major=False, b_mn_major=False, accumulate=False, split_k=True):
    """C[M,N] (+)= A[M,K] @ B[N,K]^T. A is stored [M, lda] (K-major) or [K, lda] (MN-major); same for B.
    split_k: hand the kernel fp32 scratch so it may split K when the output alone cannot fill the SMs."""
    require_cuda(A, B, C)
    if A.dtype != torch.bfloat16 or B.dtype != torch.bfloat16:
        raise B2Error("gemm_bf16: A and B must be bfloat16")
    if A.stride(1) != 1 or B.stride(1) != 1 or C.stride(1) != 1:
        raise B2Error("gemm_bf16: unit inner stride required")
    with torch.cuda.device(A.device):
        ws, ws_bytes = None, 0
        tiles = ((M + 127) // 128) * ((N + 255) // 256)
        if split_k and tiles < 12 * 148 and K >= 2048:
            ws_bytes = _lib.lib.b2_gemm_workspace_bytes(int(M), int(N))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=A.device)
        call("b2_gemm_bf16", ptr(A), int(a_mn_major), A.stride(0), ptr(B), int(b_mn_major), B.stride(0), ptr(C),
             dtype_code(C.dtype), C.stride(0), int(M), int(N), int(K), int(bool(accumulate)), ptr(ws), ws_bytes,
             stream_ptr())
    return C


def lmhead_ce_bwd(H, W, labels, lse, gscale, row0, rows, dlogits, *, kept, dH=None, dW=None, dW_accumulate=False, db=None,
                  db_accumulate=False, bias=None, shift_T=0, ignore_index=-100, round_logits_bf16=True):
    """The LM-head backward of rows [row0, row0+rows) in one call (b2_lmhead_ce_bwd): d(logits) into `dlogits` (bf16
    [>= rows, ldd]; with kept=True it holds those rows' kept logits and is rewritten in place), then dH (bf16 [rows, D]),
    dW (bf16 or fp32 [V, D], optionally accumulated) and db (fp32 [V], optionally accumulated); None skips an output."""
    _check_hw(H, W, bias)
    D, V = H.shape[1], W.shape[0]
    _check_logits_buffer(dlogits, rows, V, H.device)
    require_cuda(dH, dW, db, lse, gscale, labels)
    if dH is not None and (dH.dtype != torch.bfloat16 or dH.stride(1) != 1 or dH.shape[0] < rows or dH.shape[1] != D):
        raise B2Error("lmhead_ce_bwd: dH must be bfloat16 [rows, D] with unit inner stride")
    if dW is not None and (dW.stride(1) != 1 or tuple(dW.shape) != (V, D)):
        raise B2Error("lmhead_ce_bwd: dW must be [V, D] with unit inner stride")
    if db is not None and (db.dtype != torch.float32 or db.numel() != V or not db.is_contiguous()):
        raise B2Error("lmhead_ce_bwd: db must be a contiguous float32 [V] tensor")
    Hc = H[row0:row0 + rows]
    with torch.cuda.device(H.device):
        ws, ws_bytes = None, 0
        tiles = ((rows + 127) // 128) * ((D + 255) // 256)
        if dH is not None and tiles < 12 * 148 and V >= 2048:
            ws_bytes = _lib.lib.b2_gemm_workspace_bytes(int(rows), int(D))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=H.device)
        call("b2_lmhead_ce_bwd", ptr(Hc), H.stride(0), ptr(W), W.stride(0), ptr(bias), ptr(labels), int(row0), int(rows),
             D, V, int(shift_T), int(ignore_index), int(bool(round_logits_bf16)), ptr(lse), ptr(gscale), ptr(dlogits),
             dlogits.stride(0), int(bool(kept)), ptr(dH), 0 if dH is None else dH.stride(0), ptr(dW),
             0 if dW is None else dtype_code(dW.dtype), 0 if dW is None else dW.stride(0), int(bool(dW_accumulate)),
             ptr(db), int(bool(db_accumulate)), ptr(ws), ws_bytes, stream_ptr())


def land_copy(srcs, dst, dst_offsets):
    """Copies the contiguous tensors `srcs` into the flat tensor `dst` at element offsets `dst_offsets` with one launch
    (b2_land_copy). All tensors share dst's dtype; sources must be 16-byte aligned (fresh allocations are)."""
    import ctypes
    n = len(srcs)
    if n == 0:
        return
    require_cuda(dst, *srcs)
    es = dst.element_size()
    ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in srcs])
    offs = (ctypes.c_size_t * n)(*[int(o) * es for o in dst_offsets])
    sizes = (ctypes.c_size_t * n)(*[t.numel() * es for t in srcs])
    with torch.cuda.device(dst.device):
        call("b2_land_copy", ptrs, offs, sizes, n, ptr(dst), stream_ptr())
