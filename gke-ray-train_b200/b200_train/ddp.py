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

`parallel_strategy_kwargs={"shard_optimizer": True}` switches to the partitioning SURVEY.md §8(e) describes: each
bucket is only reduce-SCATTERED during backward (rank r keeps slice r of every bucket), AdamW runs on those slices
(optimizer state and optimizer HBM traffic divided by the world size), and the updated slices are all-gathered bucket by
bucket, front of the model first, on the side stream — under the remaining AdamW launches and, with
`"overlap_param_sync": True`, under the next forward pass (each module waits for the bucket that holds its parameters).
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


# set by the trainer shim for its workers: ScalingConfig(use_gpu=False) means CPU workers even on a box that has GPUs
_use_gpu = None


def get_device():
    """`ray.train.torch.get_device()` (ref: ray-jobs/pytorch_llm_ray.py:128): this worker's GPU (the CPU for a worker of
    a `use_gpu=False` trainer)."""
    if _use_gpu is False or not torch.cuda.is_available():
        return torch.device("cpu")
    return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())


def plan_buckets(offsets, total, esize, cap_bytes, front_params=0):
    """Cuts a flat buffer (parameter start offsets `offsets`, `total` elements, `esize` bytes each) into contiguous
    ranges from its END, each closed as soon as it holds at least `cap_bytes`; the last one takes whatever is left at the
    front. `front_params` > 0 makes the first that many parameters a bucket of their own: the embeddings sit at the front
    of the buffer and their gradient is the LAST one backward produces (for a tied LM head: only after the embedding
    lookup's backward), so everything else must not wait for them. Returns [(start, numel, [parameter indices])] in
    launch order (last parameters first)."""
    out, end, members = [], total, []
    floor = front_params if 0 < front_params < len(offsets) else 0
    for i in range(len(offsets) - 1, floor - 1, -1):
        members.append(i)
        start = offsets[i] if i > 0 else 0
        if (end - start) * esize >= cap_bytes or i == floor:
            out.append((start, end - start, members))
            end, members = start, []
    if floor:
        out.append((0, end, list(range(floor - 1, -1, -1))))
    return out


class _Bucket:
    """A contiguous element range of one flat gradient buffer, reduced by one kernel launch."""
    __slots__ = ("buf", "offset", "numel", "nparams", "pending", "shard", "ag_event", "key")

    def __init__(self, buf, offset, numel, nparams, key):
        self.buf, self.offset, self.numel, self.nparams, self.pending = buf, offset, numel, nparams, nparams
        self.shard = None          # (begin, end) of this rank's slice, in buffer elements (sharded optimizer)
        self.ag_event = None
        self.key = key


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, process_group=None, broadcast_buffers=True,
                 bucket_cap_mb=None, overlap=True, shard_optimizer=False, overlap_param_sync=False, **unused):
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
        for buf in self.flat_buffers:
            if getattr(buf.data, "_b2_symm", None) is None or getattr(buf.grad, "_b2_symm", None) is None:
                raise B2Error("prepare_model found parameters that a b200_train.AdamW has already moved into ordinary "
                              "device memory; the peer-memory all-reduce needs them in symmetric memory. Create the "
                              "optimizer AFTER prepare_model(model), as the reference does (pytorch_llm_ray.py:230,236)")
        self.shard_optimizer = bool(shard_optimizer)
        # DDP's _sync_module_states: every rank starts from rank 0's weights and buffers (bootstrap over NCCL)
        for buf in self.flat_buffers:
            dist.broadcast(buf.data, src=0, group=process_group)
            buf.comm = self.comm
            buf.reduced_sumsq = None
            buf._sumsq_store = torch.zeros(self.world_size, dtype=torch.float32, device=device)
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
        self.overlap = bool(int(os.environ.get("B2_DDP_OVERLAP", int(bool(overlap)))))
        self.overlap_param_sync = bool(int(os.environ.get("B2_DDP_PARAM_OVERLAP", int(bool(overlap_param_sync))))) \
            and self.overlap and self.shard_optimizer
        # measured on GPT-2 124M / 2 GPUs (profiles/README.md): 100 MiB buckets on <= 96 CTAs beat torch's 25 MiB default
        bucket_cap_mb = float(os.environ.get("B2_DDP_BUCKET_MB", 100 if bucket_cap_mb is None else bucket_cap_mb))
        self._overlap_blocks = int(os.environ.get("B2_DDP_MAX_BLOCKS", 96))
        cap_bytes = int(bucket_cap_mb * 1024 * 1024) if self.overlap else (1 << 62)      # no overlap: one launch per buffer
        emb = {id(p) for m in module.modules() if isinstance(m, nn.Embedding) for p in m.parameters(recurse=False)}
        self._buckets, self._bucket_of = [], {}
        for buf in self.flat_buffers:
            front = 0
            while self.overlap and front < len(buf.params) and id(buf.params[front]) in emb:
                front += 1
            mine = []
            for start, numel, members in plan_buckets(buf.offsets, buf.numel, buf.data.element_size(), cap_bytes, front):
                for j in members:
                    self._bucket_of[id(buf.params[j])] = len(self._buckets)
                b = _Bucket(buf, start, numel, len(members), len(self._buckets))
                self._buckets.append(b)
                mine.append(b)
            if self.shard_optimizer:
                # rank r reduces, clips and updates slice r of EVERY bucket, then pushes it to the peers; exp_avg /
                # exp_avg_sq exist for those slices only. Listed front of the buffer first: that is the order in which
                # the optimizer updates them and the parameters travel, so the first layers are complete first.
                mine.sort(key=lambda b: b.offset)
                so, shards = 0, []
                for b in mine:
                    lo, hi = self.comm.slice_bounds(b.numel, buf.dtype)
                    b.shard = (b.offset + lo, b.offset + hi)
                    b.ag_event = torch.cuda.Event()
                    shards.append((b.offset + lo, b.offset + hi, so))
                    so += hi - lo
                buf.shards = shards
                buf._shard_buckets = mine
                buf.after_shard_update = self._make_shard_sync(buf)
                buf.param_gate = self._make_param_gate(buf)
        self._next_bucket = 0
        # B2_DDP_TRACE=1: CUDA events around every bucket launch of a backward pass (trace_report() reads them)
        self._trace = [] if os.environ.get("B2_DDP_TRACE") == "1" else None
        # the side stream outranks the compute stream (B2_DDP_STREAM_PRIORITY, default -1 = high): a bucket's CTAs are placed
        # as soon as SM slots free up instead of after the backward kernels queued at that moment, so the exchange starts
        # (and every rank reaches its start barrier) when the bucket is ready, not a few kernels later
        prio = int(os.environ.get("B2_DDP_STREAM_PRIORITY", "-1"))
        self._comm_stream = torch.cuda.Stream(device=device, priority=prio) if self.overlap else None
        self._ag_pending = {}
        for p in params:
            p.register_post_accumulate_grad_hook(self._on_grad)
        if self.overlap_param_sync:
            self._mod_buckets = {}
            for m in module.modules():
                keys = sorted({self._bucket_of[id(p)] for p in m.parameters(recurse=False) if id(p) in self._bucket_of})
                if keys:
                    self._mod_buckets[id(m)] = keys
                    m.register_forward_pre_hook(self._gate_module)

    # ------------------------------------------------------------------ backward-side plumbing
    def _launch_bucket(self, b, tail, last_of_buf):
        """Peer-memory all-reduce (or, with a sharded optimizer, reduce-scatter) of one bucket. With overlap it runs on
        the side stream, ordered after everything the backward stream has enqueued so far, on a capped grid so it shares
        the SMs with the remaining backward kernels. The launch that completes a buffer also delivers the W per-rank
        sums of squares (its last act, after the end barrier) into the buffer's own store."""
        b.buf.flush_landing()             # gradients still waiting in their own allocations move into the buffer first
        first = self._next_bucket == 0 or b.buf is not self._buckets[self._next_bucket - 1].buf
        mode = 1 if first else 2          # sum-of-squares slots: restart on a buffer's first bucket, then keep adding
        kw = dict(want_sumsq=mode, offset=b.offset, numel=b.numel, scatter_only=self.shard_optimizer,
                  sumsq_out=b.buf._sumsq_store if last_of_buf else None)
        tr = None
        if self._trace is not None:
            tr = {"key": b.key, "mb": b.numel * b.buf.data.element_size() / 2 ** 20, "tail": bool(tail),
                  "ready": torch.cuda.Event(enable_timing=True), "start": torch.cuda.Event(enable_timing=True),
                  "end": torch.cuda.Event(enable_timing=True)}
            self._trace.append(tr)
        if self.overlap:
            ev = tr["ready"] if tr else torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                if tr:
                    tr["start"].record()
                # tail: nothing left to overlap with -> whole GPU
                self.comm.allreduce_(b.buf.grad, max_blocks=0 if tail else self._overlap_blocks, **kw)
                if tr:
                    tr["end"].record()
        else:
            if tr:
                tr["ready"].record(); tr["start"].record()
            self.comm.allreduce_(b.buf.grad, **kw)
            if tr:
                tr["end"].record()
        if last_of_buf:
            b.buf.reduced_sumsq = b.buf._sumsq_store

    def _launch_ready(self, force=False):
        while self._next_bucket < len(self._buckets):
            b = self._buckets[self._next_bucket]
            if b.pending > 0 and not force:
                break
            last_of_buf = (self._next_bucket + 1 == len(self._buckets)
                           or self._buckets[self._next_bucket + 1].buf is not b.buf)
            self._launch_bucket(b, tail=force, last_of_buf=last_of_buf)
            self._next_bucket += 1

    def _on_grad(self, p):
        # (the flat buffer's own hook ran first: p.grad already IS the parameter's slice of the flat gradient buffer,
        # whatever the caller did to the gradients between forward and backward)
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
        self.reduce_gradients()

    def reduce_gradients(self):
        """Reduces every bucket that has not been launched yet and joins the side stream. Called at the end of each
        synchronised backward pass; callable by hand after backward passes under `no_sync()` (the buckets, kernels and
        launch order are the ones the overlapped path uses)."""
        for buf in self.flat_buffers:
            buf.install_grads()            # parameters that received no gradient contribute zeros
        if self._trace is not None:
            self._trace_bwd_end = torch.cuda.Event(enable_timing=True)
            self._trace_bwd_end.record()
        self._launch_ready(force=True)     # whatever is left (all of it without overlap; unused parameters' buckets)
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
        if self._trace is not None:
            self._trace_joined = torch.cuda.Event(enable_timing=True)
            self._trace_joined.record()
        self._next_bucket = 0
        for b in self._buckets:
            b.pending = b.nparams

    def trace_report(self):
        """B2_DDP_TRACE=1: timeline of the LAST backward pass's bucket launches, in ms relative to the moment the first
        bucket became ready — when each bucket's last gradient had landed (`ready`), when its kernel started and ended on
        the side stream, whether it ran as tail (after backward, whole GPU), when backward ended and when the main
        stream had joined the side stream (`exposed_ms` = joined - backward end: communication nothing could hide)."""
        if not self._trace:
            return None
        torch.cuda.synchronize()
        n = len(self._buckets)
        last = self._trace[-n:]
        t0 = last[0]["ready"]
        out = {"buckets": [{"bucket": t["key"], "MiB": round(t["mb"], 1), "tail": t["tail"],
                            "ready_ms": round(t0.elapsed_time(t["ready"]), 3), "start_ms": round(t0.elapsed_time(t["start"]), 3),
                            "end_ms": round(t0.elapsed_time(t["end"]), 3)} for t in last],
               "backward_end_ms": round(t0.elapsed_time(self._trace_bwd_end), 3),
               "joined_ms": round(t0.elapsed_time(self._trace_joined), 3)}
        out["exposed_ms"] = round(out["joined_ms"] - out["backward_end_ms"], 3)
        self._trace.clear()
        return out

    # ------------------------------------------------------------------ sharded optimizer: parameter all-gather
    def _make_shard_sync(self, buf):
        def sync(k):
            b = buf._shard_buckets[k]
            if not self.overlap:
                self.comm.allgather_(buf.data, offset=b.offset, numel=b.numel)
                return
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                self.comm.allgather_(buf.data, offset=b.offset, numel=b.numel,
                                     max_blocks=self._overlap_blocks if self.overlap_param_sync else 0)
                b.ag_event.record(self._comm_stream)
            if self.overlap_param_sync:
                self._ag_pending[b.key] = b.ag_event
            elif k + 1 == len(buf._shard_buckets):
                # every later reader of the parameters is ordered behind the gathers (they still ran under the AdamW
                # launches of the buckets that followed them)
                torch.cuda.current_stream().wait_stream(self._comm_stream)
        return sync

    def _make_param_gate(self, buf):
        def gate(offset):
            """Called by code that reads a parameter of this buffer without going through its module's forward."""
            if not self._ag_pending:
                return
            for b in buf._shard_buckets:
                if b.offset <= offset < b.offset + b.numel:
                    ev = self._ag_pending.pop(b.key, None)
                    if ev is not None:
                        torch.cuda.current_stream().wait_event(ev)
                    return
        return gate

    def _gate_module(self, mod, args):
        if not self._ag_pending:
            return
        cur = None
        for k in self._mod_buckets.get(id(mod), ()):
            ev = self._ag_pending.pop(k, None)
            if ev is not None:
                cur = cur or torch.cuda.current_stream()
                cur.wait_event(ev)

    def wait_parameters(self):
        """Orders the current stream behind every parameter all-gather still in flight (sharded optimizer with
        overlap_param_sync). forward() ends with it, so backward passes and anything after them are safe; call it
        yourself before reading parameters right after optimizer.step() (state_dict() does)."""
        if self._ag_pending:
            cur = torch.cuda.current_stream()
            for ev in self._ag_pending.values():
                cur.wait_event(ev)
            self._ag_pending.clear()

    # ------------------------------------------------------------------ nn.Module surface
    def forward(self, *args, **kwargs):
        self.forward_prepare()
        out = self.module(*args, **kwargs)
        self.wait_parameters()
        return out

    def forward_prepare(self):
        """Forward-side bookkeeping (stale reductions forgotten). forward() calls it; call it yourself only when you
        drive `self.module` piecewise (e.g. its feature extractor + a fused loss)."""
        if not torch.is_grad_enabled():
            return                  # an evaluation pass between backward and clip must not forget the reduction's Σg²
        for buf in self.flat_buffers:
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
        self.wait_parameters()
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
