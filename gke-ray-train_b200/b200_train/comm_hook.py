"""Gradient-bucket communication hook for STOCK torch DistributedDataParallel.

The reference gets its data parallelism from `ray.train.torch.prepare_model` (ref: ray-jobs/pytorch_llm_ray.py:230), which
wraps the model in torch's own DDP. That wrapper already has a plug-in point for the gradient exchange:
`DistributedDataParallel.register_comm_hook(state, hook)` with `hook(state, bucket: dist.GradBucket) ->
torch.futures.Future[Tensor]` ($SP/torch/nn/parallel/distributed.py:1987-2067). This module is the
drop-in at THAT boundary, for jobs that want to keep Ray's / torch's wrapper untouched:

    model = ray.train.torch.prepare_model(model)                       # stock DDP
    state = b200_train.PeerAllreduceState()                            # after the process group exists
    model.register_comm_hook(state, b200_train.peer_allreduce_hook)

Each bucket is averaged over the ranks by the same peer-memory kernel the `b200_train.prepare_model` path uses
(b2_peer_allreduce / b2_peer_allreduce_mc, include/b2train.h) instead of an NCCL all-reduce. torch DDP owns its bucket
memory (ordinary cudaMalloc blocks that peers cannot map), so the hook stages each bucket through a symmetric buffer:
bucket -> staging, in-place peer all-reduce, staging -> bucket. Those two device copies are what this route costs over
`b200_train.prepare_model`, whose gradients live in symmetric memory to begin with (no copy, fused sum of squares,
sharded optimizer); prefer that one when the wrapper may change.

Contract kept from torch's default hook: the bucket is NOT pre-divided when a hook is registered, the returned tensor
has the shape and dtype of `bucket.buffer()`, and after the hook the bucket holds the mean over ranks. The hook runs on
the autograd engine's thread; everything it enqueues goes to the stream that is current there (the backward stream), so
the returned future is already complete and ordered with what follows on that stream.
"""
import torch
import torch.distributed as dist

from ._lib import B2Error
from .symm import PeerComm

_SUPPORTED = (torch.float32, torch.bfloat16)


class PeerAllreduceState:
    """`state` argument of `peer_allreduce_hook`: the peer-memory handle of this process plus one symmetric staging
    buffer per gradient dtype, grown on demand.

    Growing is a collective (every rank maps every other rank's new block). DDP calls the hook for the same buckets in
    the same order on every rank, so the ranks grow together; call `reserve()` up front to keep it out of the first
    backward. Construct after `init_process_group` (it exchanges memory handles through the group)."""

    def __init__(self, process_group=None, device=None):
        if not dist.is_available() or not dist.is_initialized():
            raise B2Error("PeerAllreduceState needs an initialised torch.distributed process group")
        if device is None:
            if not torch.cuda.is_available():
                raise B2Error("peer_allreduce_hook runs on CUDA devices only (no CPU path)")
            device = torch.device("cuda", torch.cuda.current_device())
        device = torch.device(device)
        if device.type != "cuda":
            raise B2Error("peer_allreduce_hook runs on CUDA devices only (no CPU path)")
        self.device = device
        self.process_group = process_group
        self.comm = PeerComm(device, process_group)
        self._stage = {}                      # dtype -> flat symmetric tensor
        self.buckets_reduced = 0

    def reserve(self, numel, dtype):
        """Collective: makes sure the staging buffer for `dtype` holds at least `numel` elements."""
        if dtype not in _SUPPORTED:
            raise B2Error(f"peer_allreduce_hook: gradient dtype {dtype} is not supported (float32 / bfloat16 only)")
        have = self._stage.get(dtype)
        if have is not None and have.numel() >= numel:
            return have
        per = 16 // torch.empty((), dtype=dtype).element_size()
        want = max(int(numel), 2 * have.numel() if have is not None else 0)
        want = (want + per - 1) // per * per
        # the previous block stays mapped until close(): unmapping is collective too and nothing is gained by doing it
        # in the middle of a backward pass
        self._stage[dtype] = self.comm.alloc(want, dtype, self.device)
        return self._stage[dtype]

    def close(self):
        self._stage.clear()
        self.comm.close()


def peer_allreduce_hook(state, bucket):
    """Communication hook: mean over ranks of one DDP gradient bucket through the peer-memory all-reduce kernel."""
    buf = bucket.buffer()
    if not buf.is_cuda or buf.device != state.device:
        raise B2Error(f"peer_allreduce_hook: bucket lives on {buf.device}, the hook state on {state.device}")
    n = buf.numel()
    fut = torch.futures.Future()
    if n:
        per = 16 // buf.element_size()                     # the kernel moves 16-byte vectors
        npad = (n + per - 1) // per * per
        stage = state.reserve(npad, buf.dtype)
        stage[:n].copy_(buf)
        if npad > n:
            stage[n:npad].zero_()
        state.comm.allreduce_(stage, numel=npad)           # in place, scale 1 / world
        buf.copy_(stage[:n])
        state.buckets_reduced += 1
    fut.set_result(buf)
    return fut
