"""prepare_model and the data-parallel wrapper it installs.

Mirror of `ray.train.torch.prepare_model(model, move_to_device=True, parallel_strategy="ddp",
parallel_strategy_kwargs=None)` as the reference calls it (ref: ray-jobs/pytorch_llm_ray.py:230): move the model to
this worker's device and, when world_size > 1, wrap it for data parallelism; with one worker the bare module comes
back (so `.module` only exists when world_size > 1 — same as Ray, cf. ref :301).

The wrapper plays the role of torch.nn.parallel.DistributedDataParallel ($SP/torch/nn/parallel/distributed.py:655):
`.module`, `forward`, `no_sync()`, rank-0 weights broadcast at construction. What differs is the data path: the
gradients live in ONE flat buffer per dtype in symmetric memory, and ranges of it ("buckets") are averaged in place over
NVLink by our peer-memory kernel (b2_peer_allreduce / b2_peer_allreduce_mc) while the rest of backward still runs — no
bucket copy-in, no NCCL all-reduce, no copy-out.
"""
import contextlib
import os

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.autograd import Variable

from ._lib import B2Error
from .flat import flatten_params
from .symm import PeerComm


def get_device():
    """`ray.train.torch.get_device()` (ref: ray-jobs/pytorch_llm_ray.py:128): this worker's GPU."""
    if not torch.cuda.is_available():
        return torch.device("cpu")
    return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())


def plan_buckets(offsets, total, esize, cap_bytes):
    """Cuts a flat buffer (parameter start offsets `offsets`, `total` elements, `esize` bytes each) into contiguous
    ranges from its END, each closed as soon as it holds at least `cap_bytes`; the last one takes whatever is left at the
    front. Returns [(start, numel, [parameter indices])] in launch order (last parameters first)."""
    out, end, members = [], total, []
    for i in range(len(offsets) - 1, -1, -1):
        members.append(i)
        start = offsets[i] if i > 0 else 0
        if (end - start) * esize >= cap_bytes or i == 0:
            out.append((start, end - start, members))
            end, members = start, []
    return out


class _Bucket:
    """A contiguous element range of one flat gradient buffer, reduced by one kernel launch."""
    __slots__ = ("buf", "offset", "numel", "nparams", "pending")

    def __init__(self, buf, offset, numel, nparams):
        self.buf, self.offset, self.numel, self.nparams, self.pending = buf, offset, numel, nparams, nparams


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, process_group=None, broadcast_buffers=True,
                 bucket_cap_mb=None, overlap=True, shard_optimizer=False, **unused):
        super().__init__()
        if not dist.is_initialized():
            raise B2Error("DistributedDataParallel needs an initialised torch.distributed process group")
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise B2Error("DistributedDataParallel: module has no parameter that requires a gradient")
        device = params[0].device
        if device.type != "cuda":
            raise B2Error("b200_train.DistributedDataParallel runs on CUDA devices only (no CPU path); "
                          "for use_gpu=False the trainer shim wraps torch's own DDP instead")
        torch.cuda.set_device(device)
        self.module = module
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.comm = PeerComm(device, process_group)
        self.flat_buffers = flatten_params(params, alloc=self.comm.alloc)   # (not `buffers`: that is nn.Module's method)
        self.shard_optimizer = bool(shard_optimizer)
        # DDP's _sync_module_states: every rank starts from rank 0's weights and buffers (bootstrap over NCCL)
        for buf in self.flat_buffers:
            dist.broadcast(buf.data, src=0, group=process_group)
            buf.comm = self.comm
            buf.reduced_sumsq = None
            if self.shard_optimizer:
                # ZeRO-1 style: rank r reduces, clips and updates only slice r of the flat buffer, then
                # pushes its slice of the updated parameters to every peer; exp_avg / exp_avg_sq exist for that slice only
                buf.shard_range = self.comm.slice_bounds(buf.numel, buf.dtype)
            buf._sumsq_store = torch.empty(self.world_size, dtype=torch.float32, device=device)
        for b in module.buffers():
            if b.is_cuda:
                dist.broadcast(b, src=0, group=process_group)
        self.require_backward_grad_sync = True
        self._callback_queued = False
        # Buckets, like torch DDP's ($SP/torch/nn/parallel/distributed.py:831-833), but as ranges of the
        # flat buffer: parameters are laid out in registration order and their gradients arrive roughly in reverse, so
        # buckets are cut from the END of each buffer and launched in that fixed order on a side stream while the rest
        # of backward still runs.
        # (B2_DDP_OVERLAP / B2_DDP_BUCKET_MB / B2_DDP_MAX_BLOCKS override the arguments for measurements)
        self.overlap = bool(int(os.environ.get("B2_DDP_OVERLAP", int(bool(overlap))))) and not self.shard_optimizer
        # measured on GPT-2 124M / 2 GPUs (profiles/README.md): 100 MiB buckets on <= 96 CTAs beat torch's 25 MiB default
        bucket_cap_mb = float(os.environ.get("B2_DDP_BUCKET_MB", 100 if bucket_cap_mb is None else bucket_cap_mb))
        self._overlap_blocks = int(os.environ.get("B2_DDP_MAX_BLOCKS", 96))
        cap_bytes = int(bucket_cap_mb * 1024 * 1024) if self.overlap else (1 << 62)      # no overlap: one launch per buffer
        self._buckets, self._bucket_of = [], {}
        for buf in self.flat_buffers:
            for start, numel, members in plan_buckets(buf.offsets, buf.numel, buf.data.element_size(), cap_bytes):
                for j in members:
                    self._bucket_of[id(buf.params[j])] = len(self._buckets)
                self._buckets.append(_Bucket(buf, start, numel, len(members)))
        self._next_bucket = 0
        self._comm_stream = torch.cuda.Stream(device=device) if self.overlap else None
        for p in params:
            p.register_post_accumulate_grad_hook(self._on_grad)

    # ------------------------------------------------------------------ backward-side plumbing
    def _launch_bucket(self, b, tail=False):
        """Peer-memory all-reduce of one bucket. With overlap it runs on the side stream, ordered after everything the
        backward stream has enqueued so far, on a capped grid so it shares the SMs with the remaining backward kernels."""
        first = self._next_bucket == 0 or b.buf is not self._buckets[self._next_bucket - 1].buf
        mode = 1 if first else 2          # sum-of-squares slots: restart on a buffer's first bucket, then keep adding
        if self.overlap:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                self.comm.allreduce_(b.buf.grad, want_sumsq=mode, offset=b.offset, numel=b.numel,
                                     max_blocks=0 if tail else self._overlap_blocks)      # nothing left to overlap with: whole GPU
        else:
            self.comm.allreduce_(b.buf.grad, want_sumsq=mode, offset=b.offset, numel=b.numel,
                                 scatter_only=self.shard_optimizer)

    def _launch_ready(self, force=False):
        while self._next_bucket < len(self._buckets):
            b = self._buckets[self._next_bucket]
            if b.pending > 0 and not force:
                break
            last_of_buf = (self._next_bucket + 1 == len(self._buckets)
                           or self._buckets[self._next_bucket + 1].buf is not b.buf)
            self._launch_bucket(b, tail=force)
            self._next_bucket += 1
            if last_of_buf:
                # the W per-rank partials of this buffer are complete once this launch retires: keep a copy, the slots
                # are reused by the next buffer / step
                stream = self._comm_stream if self.overlap else torch.cuda.current_stream()
                with torch.cuda.stream(stream):
                    b.buf._sumsq_store.copy_(self.comm._parts)
                b.buf.reduced_sumsq = b.buf._sumsq_store

    def _on_grad(self, p):
        if not self.require_backward_grad_sync:
            return
        if not self._callback_queued:
            self._callback_queued = True
            Variable._execution_engine.queue_callback(self._finalize_backward)
        k = self._bucket_of.get(id(p))
        if k is not None:
            self._buckets[k].pending -= 1
            if self.overlap:
                self._launch_ready()

    def _finalize_backward(self):
        """Runs once per backward pass, after the last gradient has been accumulated."""
        self._callback_queued = False
        for buf in self.flat_buffers:
            buf.install_grads()            # a stray .grad (allocated by autograd after a foreign zero_grad) is copied in
        self._launch_ready(force=True)     # whatever is left (all of it without overlap; unused parameters' buckets)
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
        self._next_bucket = 0
        for b in self._buckets:
            b.pending = b.nparams

    # ------------------------------------------------------------------ nn.Module surface
    def forward(self, *args, **kwargs):
        self.forward_prepare()
        return self.module(*args, **kwargs)

    def forward_prepare(self):
        """Forward-side bookkeeping (gradient views in place, stale reductions forgotten). forward() calls it; call it
        yourself only when you drive `self.module` piecewise (e.g. its feature extractor + a fused loss)."""
        for buf in self.flat_buffers:
            buf.ensure_grads()
            buf.reduced_sumsq = None

    @contextlib.contextmanager
    def no_sync(self):
        """Gradient accumulation: backward passes inside this context only accumulate locally
        ($SP/torch/nn/parallel/distributed.py:1474-1500; GRADIENT_ACCUMULATION_STEPS, ref: fine_tune_config.json:14)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def state_dict(self, *args, **kwargs):
        # same key layout as torch DDP: entries are prefixed with "module."
        return super().state_dict(*args, **kwargs)

    def close(self):
        self.comm.close()


def prepare_model(model, move_to_device=True, parallel_strategy="ddp", parallel_strategy_kwargs=None):
    parallel_strategy_kwargs = parallel_strategy_kwargs or {}
    device = get_device()
    if isinstance(move_to_device, torch.device):
        device, move_to_device = move_to_device, True
    if device.type == "cuda":
        torch.cuda.set_device(device)
    if move_to_device:
        model = model.to(device)
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if world > 1 and parallel_strategy:
        if parallel_strategy != "ddp":
            raise B2Error(f"parallel_strategy={parallel_strategy!r}: the reference uses 'ddp' only")
        if device.type != "cuda":
            return nn.parallel.DistributedDataParallel(model, **parallel_strategy_kwargs)   # use_gpu=False: stock path
        return DistributedDataParallel(model, **parallel_strategy_kwargs)
    return model
