"""Flat parameter / gradient storage.

The fused kernels (gradient sum-of-squares, AdamW, peer-memory all-reduce) run over one contiguous buffer per
dtype instead of one launch per parameter. `flatten_params` re-homes every parameter's storage into such a
buffer (the parameter objects stay the same, so `model.parameters()`, `state_dict()` and autograd are
unaffected) and installs `.grad` tensors that are views of a matching flat gradient buffer — the arrangement
DistributedDataParallel calls `gradient_as_bucket_view` ($SP/torch/nn/parallel/distributed.py:655), which
prepare_model installs for the reference (ref: ray-jobs/pytorch_llm_ray.py:230).
"""
import torch

import os

# measurement aid (tools/host_bound_probe.py): leave the gradients where autograd put them until install_grads() runs
_LAND_LATE = os.environ.get("B2_LAND_LATE") == "1"
# gradients that wait in their own allocations are moved into the flat buffer once this many bytes have gathered (bounds
# the extra memory; B2_LAND_FLUSH_MB=0 copies each one at once, the round-2 behaviour before the gather kernel)
_LAND_FLUSH_BYTES = int(float(os.environ.get("B2_LAND_FLUSH_MB", "64")) * (1 << 20))

ALIGN = 64  # elements; every parameter starts on a 64-element boundary (>= 128 B), padding stays zero
# The buffer length is a multiple of 64 * lcm(1..8) elements, so that for any world size up to 8 the per-rank slices of
# the peer-memory collectives (and of a sharded optimizer) are equal-sized and start on 64-element boundaries.
TOTAL_MULTIPLE = 64 * 840


class FlatBuffer:
    """One dtype's worth of parameters laid out back to back."""

    def __init__(self, dtype, device, params, alloc=None):
        self.dtype, self.device = dtype, device
        self.params = list(params)
        self.offsets = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
        self.numel = (off + TOTAL_MULTIPLE - 1) // TOTAL_MULTIPLE * TOTAL_MULTIPLE
        self.comm = None
        alloc = alloc or (lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev))
        self.data = alloc(self.numel, dtype, device)
        self.grad = alloc(self.numel, dtype, device)
        self.data.zero_()
        self.grad.zero_()
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.data[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p._b2_flat = (self, o)
        self._views = None
        self._index = {id(p): i for i, p in enumerate(self.params)}
        self._stolen, self._stolen_bytes, self._stolen_set = [], 0, set()   # gradients waiting for the next b2_land_copy launch
        for i, p in enumerate(self.params):
            if p.grad is not None:             # a gradient that already exists moves into the buffer; none stays none
                self.grad_view(i).copy_(p.grad)
                p.grad = self.grad_view(i)
        for p in self.params:
            # registered before any wrapper's / optimizer's own hooks: by the time those run, the gradient is in place
            p.register_post_accumulate_grad_hook(self._land)
        # set by clip_grad_norm_ when an attached optimizer will apply the coefficient inside its kernel
        self.pending_coef = None
        self.optimizer_attached = False
        # exp_avg / exp_avg_sq, created by the optimizer
        self.exp_avg = None
        self.exp_avg_sq = None
        # [(begin, end, state_offset)] element ranges this rank updates when a data-parallel wrapper shards the optimizer
        self.shards = None
        self.after_shard_update = None       # wrapper's callback: (shard index) -> all-gather of that range

    def grad_view(self, i):
        """The slice of the flat gradient buffer that is parameter i's .grad (one persistent tensor object per parameter:
        building 100+ views on every call would cost the host about a millisecond per step)."""
        if self._views is None:
            self._views = [self.grad[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, self.offsets)]
        return self._views[i]

    def _land(self, p):
        """Post-accumulate-grad hook. After `zero_grad(set_to_none=True)` (ours, the model's, HF Trainer's) a parameter
        has no .grad, so autograd hands over the freshly computed gradient tensor instead of adding it to something; here
        that tensor is copied into the parameter's slice of the flat buffer — one read and one write — and released;
        the copies are gathered: a gradient waits in its own allocation until 64 MB have accumulated (or someone needs the
        buffer) and one b2_land_copy launch then moves them all, instead of one small copy kernel per parameter.
        (Zero-filling the buffer up front and letting autograd add into it costs a fill plus a read-modify-write: four
        passes over the gradients instead of two.) Later micro-batches of an accumulation find .grad already pointing
        into the buffer and autograd adds in place."""
        if _LAND_LATE:
            return
        i = self._index.get(id(p))
        if i is None:                          # a copy of one of our parameters that inherited the hook: not ours
            return
        want = self.grad_view(i)
        g = p.grad
        if g is want:
            return
        if g is None or g.data_ptr() == want.data_ptr():
            p.grad = want
            return
        if self._can_queue(g, p):
            # one launch for many gradients: this one waits (still owned by the parameter) until enough bytes have
            # gathered, the data-parallel wrapper is about to reduce its bucket, or the buffer is next read as a whole
            if i in self._stolen_set:          # queued by an earlier micro-batch; autograd has added to it in place
                return
            self._stolen.append(i)
            self._stolen_set.add(i)
            self._stolen_bytes += g.numel() * g.element_size()
            if self._stolen_bytes >= _LAND_FLUSH_BYTES:
                self.flush_landing()
        else:
            want.copy_(g)
            p.grad = want

    def _can_queue(self, g, p):
        """A gradient the gather kernel can take: on the device, the buffer's dtype, contiguous, 16-byte aligned."""
        return (g.is_cuda and g.dtype == self.dtype and g.is_contiguous() and g.data_ptr() % 16 == 0
                and g.numel() == p.numel())

    def flush_landing(self):
        """Copies every gradient that waits in its own allocation into its slice of the flat buffer — one b2_land_copy
        launch instead of one copy per parameter — and points the parameters' .grad at the slices."""
        if not self._stolen:
            return
        from . import ops
        idx, self._stolen, self._stolen_bytes = self._stolen, [], 0
        self._stolen_set.clear()
        srcs, offs, keep = [], [], []
        for i in idx:
            p = self.params[i]
            g, want = p.grad, self.grad_view(i)
            if g is None or g is want or g.data_ptr() == want.data_ptr():
                continue                       # dropped or re-pointed since it was queued
            srcs.append(g)
            offs.append(self.offsets[i])
            keep.append(i)
        ops.land_copy(srcs, self.grad, offs)
        for i in keep:
            self.params[i].grad = self.grad_view(i)

    def install_grads(self):
        """Make every parameter's .grad its slice of the flat gradient buffer, as the kernels that run over the whole
        buffer (sum of squares, all-reduce, AdamW) expect: a gradient that lives elsewhere (assigned by hand, or produced
        before the buffer existed) is copied in; a parameter WITHOUT a gradient gets a zeroed slice (it contributes
        nothing to the norm; the optimizer skips it, see AdamW._table_for)."""
        self.flush_landing()
        moved = 0
        self.grad_view(0) if self.params else None
        for p, want in zip(self.params, self._views or ()):
            g = p.grad
            if g is want:
                continue
            if g is None:
                want.zero_()
                p.grad = want
            elif g.data_ptr() != want.data_ptr():
                want.copy_(g)
                p.grad = want
                moved += 1
            else:
                p.grad = want          # same memory through another tensor object: adopt ours
        return moved

    def ensure_grads(self):
        """Kept for callers of the round-1 API: gradients dropped by a `zero_grad(set_to_none=True)` no longer need to be
        re-installed before a backward pass — `_land` moves each one into the buffer as it is produced."""
        return

    def zero_grad(self):
        """`optimizer.zero_grad(set_to_none=True)` (ref: ray-jobs/pytorch_llm_ray.py:273): the gradients are dropped, as
        in torch — the flat buffer is NOT cleared; each slice is overwritten when its gradient lands."""
        self._stolen, self._stolen_bytes = [], 0
        self._stolen_set.clear()
        for p in self.params:
            p.grad = None
        self.pending_coef = None


def flat_of(p):
    """(FlatBuffer, offset) if `p` lives in a flat buffer. A deep-copied parameter carries a copy of the attribute but
    not the storage, so the claim is verified against the buffer."""
    f = getattr(p, "_b2_flat", None)
    if f is None:
        return None
    buf, off = f
    if p.data_ptr() != buf.data.data_ptr() + off * buf.data.element_size():
        del p._b2_flat
        return None
    return f


def flatten_params(params, alloc=None):
    """Flattens the given parameters (those not already flattened) by (device, dtype). Returns the list of
    FlatBuffers that now hold them, in first-appearance order."""
    seen, groups, order = set(), {}, []
    existing = []
    for p in params:
        if id(p) in seen:
            continue
        seen.add(id(p))
        if not p.requires_grad:
            continue
        f = flat_of(p)
        if f is not None:
            if f[0] not in existing:
                existing.append(f[0])
            continue
        key = (p.device, p.dtype)
        if key not in groups:
            groups[key] = []
            order.append(key)
        groups[key].append(p)
    out = list(existing)
    for key in order:
        out.append(FlatBuffer(key[1], key[0], groups[key], alloc=alloc))
    return out


def buffers_of(params):
    """Distinct FlatBuffers behind `params`, plus the parameters that are not flattened."""
    bufs, loose, seen = [], [], set()
    for p in params:
        if id(p) in seen:
            continue
        seen.add(id(p))
        f = flat_of(p)
        if f is None:
            loose.append(p)
        elif f[0] not in bufs:
            bufs.append(f[0])
    return bufs, loose
