"""prepare_model and the data-parallel wrapper it installs.

Mirror of `ray.train.torch.prepare_model(model, move_to_device=True, parallel_strategy="ddp",
parallel_strategy_kwargs=None)` as the reference calls it (ref: ray-jobs/pytorch_llm_ray.py:230): move the model to
this worker's device and, when world_size > 1, wrap it for data parallelism; with one worker the bare module comes
back (so `.module` only exists when world_size > 1 — same as Ray, cf. ref :301).

The wrapper plays the role of torch.nn.parallel.DistributedDataParallel ($SP/torch/nn/parallel/distributed.py:655):
`.module`, `forward`, `no_sync()`, rank-0 weights broadcast at construction. What differs is the data path: the
gradients live in ONE flat buffer per dtype in symmetric memory, and when backward finishes a single peer-memory
kernel per buffer (b2_peer_allreduce) averages them over NVLink — no bucket copy-in, no NCCL all-reduce, no copy-out.
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


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, process_group=None, broadcast_buffers=True,
                 shard_optimizer=False, **unused):
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
        self.buffers = flatten_params(params, alloc=self.comm.alloc)
        self.shard_optimizer = bool(shard_optimizer)
        # DDP's _sync_module_states: every rank starts from rank 0's weights and buffers (bootstrap over NCCL)
        for buf in self.buffers:
            dist.broadcast(buf.data, src=0, group=process_group)
            buf.comm = self.comm
            buf.reduced_sumsq = None
        for b in module.buffers():
            if b.is_cuda:
                dist.broadcast(b, src=0, group=process_group)
        self.require_backward_grad_sync = True
        self._callback_queued = False
        for p in params:
            p.register_post_accumulate_grad_hook(self._on_grad)

    # ------------------------------------------------------------------ backward-side plumbing
    def _on_grad(self, _p):
        if self._callback_queued or not self.require_backward_grad_sync:
            return
        self._callback_queued = True
        Variable._execution_engine.queue_callback(self._finalize_backward)

    def _finalize_backward(self):
        """Runs once per backward pass, after the last gradient has been accumulated."""
        self._callback_queued = False
        for buf in self.buffers:
            if buf.install_grads():
                pass  # a stray .grad was copied into the flat buffer
            parts = self.comm.allreduce_(buf.grad, want_sumsq=True, scatter_only=False)
            buf.reduced_sumsq = parts.clone()      # `world` floats; clip_grad_norm_ reuses them instead of re-reading

    # ------------------------------------------------------------------ nn.Module surface
    def forward(self, *args, **kwargs):
        for buf in self.buffers:
            missing = [i for i, p in enumerate(buf.params) if p.grad is None]
            if missing:
                # someone dropped the gradients (zero_grad(set_to_none=True) of a stock optimizer): hand autograd
                # zero-filled views of the flat buffer again so it accumulates in place
                if len(missing) == len(buf.params):
                    buf.grad.zero_()
                else:
                    for i in missing:
                        buf.grad_view(i).zero_()
                buf.install_grads()
            buf.reduced_sumsq = None
        return self.module(*args, **kwargs)

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
