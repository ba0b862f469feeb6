"""The slice of the `ray.train` surface that the reference's two jobs touch, for one multi-GPU box.

ref: ray-jobs/pytorch_llm_ray.py:12-16 imports `ray.train`, `ScalingConfig`, `RunConfig`, `Checkpoint`,
`TorchTrainer`, `TorchConfig`; it calls `train.get_context()` (:125), `train.torch.get_device()` (:128),
`train.torch.prepare_data_loader` (:216), `train.torch.prepare_model` (:230), `train.report` (:310),
`train.CheckpointConfig` (:355) and `TorchTrainer(...).fit()` (:368-376). ref: ray-jobs/fine_tune_llama_ray.py:445-457
does the same with `ScalingConfig(num_workers, use_gpu=True, resources_per_worker={"GPU": 1})`, reads
`get_context().get_world_rank()/get_world_size()` (:201-202) and imports `RayTrainReportCallback`, `prepare_trainer` from
`ray.train.huggingface.transformers` (:8) — here `train.huggingface.transformers`.

A job written against Ray switches with
    from b200_train import train
    from b200_train.train import ScalingConfig, RunConfig, Checkpoint, TorchTrainer, TorchConfig
If real Ray is importable AND `B2_USE_RAY=1`, TorchTrainer delegates to it (workers then call our prepare_model
inside Ray's process group). Otherwise workers are plain processes: one per GPU, rendezvous over 127.0.0.1, the
process group initialised with the backend TorchConfig names (nccl for GPUs, gloo for use_gpu=False) — which is
all the reference relies on Ray for on this path. Cluster provisioning, storage sync and fault tolerance are out
of scope.
"""
import importlib.util
import os
import shutil
import socket
import tempfile
import time as _time
import types
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Optional

import torch as _t
import torch.distributed as dist
import torch.multiprocessing as mp

from . import ddp as _ddp


# ----------------------------------------------------------------------------- configs (same field names as Ray's)
@dataclass
class ScalingConfig:
    num_workers: int = 1
    use_gpu: bool = False
    resources_per_worker: Optional[Dict[str, float]] = None
    trainer_resources: Optional[Dict[str, float]] = None
    placement_strategy: str = "PACK"


@dataclass
class CheckpointConfig:
    num_to_keep: Optional[int] = None
    checkpoint_score_attribute: Optional[str] = None
    checkpoint_score_order: str = "max"


@dataclass
class RunConfig:
    name: Optional[str] = None
    storage_path: Optional[str] = None
    checkpoint_config: Optional[CheckpointConfig] = None
    failure_config: Any = None
    verbose: Any = None


@dataclass
class TorchConfig:
    backend: Optional[str] = None
    init_method: str = "env"
    timeout_s: int = 1800


class Checkpoint:
    """Directory-backed checkpoint handle (ref: pytorch_llm_ray.py:305 `Checkpoint.from_directory`)."""

    def __init__(self, path):
        self.path = path

    @classmethod
    def from_directory(cls, path):
        return cls(str(path))

    def to_directory(self, path=None):
        path = path or tempfile.mkdtemp(prefix="b2_ckpt_")
        if os.path.abspath(path) != os.path.abspath(self.path):
            shutil.copytree(self.path, path, dirs_exist_ok=True)
        return path

    def as_directory(self):
        import contextlib
        return contextlib.nullcontext(self.path)


@dataclass
class Result:
    metrics: Optional[Dict[str, Any]] = None
    checkpoint: Optional[Checkpoint] = None
    error: Optional[BaseException] = None
    path: Optional[str] = None
    metrics_history: list = field(default_factory=list)


# ----------------------------------------------------------------------------- per-worker session
class _Context:
    def __init__(self, rank, world, local_rank, storage_dir, ckpt_cfg):
        self._rank, self._world, self._local = rank, world, local_rank
        self.storage_dir, self.ckpt_cfg = storage_dir, ckpt_cfg
        self.reports = []
        self.kept = []          # (score, path) of persisted checkpoints

    def get_world_rank(self):
        return self._rank

    def get_world_size(self):
        return self._world

    def get_local_rank(self):
        return self._local

    def get_local_world_size(self):
        return self._world

    def get_node_rank(self):
        return 0


_ctx: Optional[_Context] = None


def get_context():
    global _ctx
    if _ctx is None:      # outside a trainer: behave like a single worker (or adopt torchrun's environment)
        _ctx = _Context(int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
                        int(os.environ.get("LOCAL_RANK", 0)), None, None)
    return _ctx


def get_checkpoint():
    return None          # the reference never restores


def report(metrics, checkpoint=None):
    """All workers call this (ref: pytorch_llm_ray.py:310). Rank 0's checkpoint directory is persisted under the
    run's storage path, keeping `num_to_keep` best by `checkpoint_score_attribute`."""
    ctx = get_context()
    entry = dict(metrics)
    if checkpoint is not None and ctx.storage_dir:
        dst = os.path.join(ctx.storage_dir, f"checkpoint_{len(ctx.reports):06d}")
        shutil.copytree(checkpoint.path, dst, dirs_exist_ok=True)
        cfg = ctx.ckpt_cfg
        score = None
        if cfg and cfg.checkpoint_score_attribute:
            score = metrics.get(cfg.checkpoint_score_attribute)
            if score is not None and cfg.checkpoint_score_order == "min":
                score = -score
        ctx.kept.append((score if score is not None else len(ctx.kept), dst))
        if cfg and cfg.num_to_keep:
            ctx.kept.sort(key=lambda t: t[0], reverse=True)
            for _, path in ctx.kept[cfg.num_to_keep:]:
                shutil.rmtree(path, ignore_errors=True)
            ctx.kept = ctx.kept[:cfg.num_to_keep]
        entry["_checkpoint_path"] = dst
    ctx.reports.append(entry)
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def prepare_data_loader(data_loader, add_dist_sampler=True, move_to_device=True, auto_transfer=False):
    """`ray.train.torch.prepare_data_loader` (ref: pytorch_llm_ray.py:216): shard with a DistributedSampler when
    world_size > 1 and move each batch to this worker's device."""
    from torch.utils.data import DataLoader, DistributedSampler, RandomSampler
    world = get_context().get_world_size()
    dl = data_loader
    if add_dist_sampler and world > 1 and not isinstance(dl.sampler, DistributedSampler):
        shuffle = isinstance(dl.sampler, RandomSampler)
        sampler = DistributedSampler(dl.dataset, num_replicas=world, rank=get_context().get_world_rank(), shuffle=shuffle)
        dl = DataLoader(dl.dataset, batch_size=dl.batch_size, sampler=sampler, num_workers=dl.num_workers,
                        collate_fn=dl.collate_fn, pin_memory=dl.pin_memory, drop_last=dl.drop_last,
                        timeout=dl.timeout, worker_init_fn=dl.worker_init_fn)
    if not move_to_device:
        return dl
    device = _ddp.get_device()

    class _DeviceLoader:
        def __init__(self, inner):
            self._inner = inner
            self.sampler = inner.sampler
            self.dataset = inner.dataset
            self.batch_size = inner.batch_size

        def __len__(self):
            return len(self._inner)

        def __iter__(self):
            def move(x):
                if isinstance(x, _t.Tensor):
                    return x.to(device, non_blocking=True)
                if isinstance(x, (list, tuple)):
                    return type(x)(move(v) for v in x)
                if isinstance(x, dict):
                    return {k: move(v) for k, v in x.items()}
                return x
            for batch in self._inner:
                yield move(batch)

    return _DeviceLoader(dl)


# `train.torch.*` namespace, as in `from ray import train; train.torch.prepare_model(model)`
torch = types.SimpleNamespace(prepare_model=_ddp.prepare_model, prepare_data_loader=prepare_data_loader,
                              get_device=_ddp.get_device)


def _prepare_trainer(trainer):
    """`ray.train.huggingface.transformers.prepare_trainer(trainer)` (imported at ref: fine_tune_llama_ray.py:8): Ray's
    version validates the trainer's callbacks and swaps its data loaders for Ray Data ones when Ray Datasets are passed; the
    reference passes HF datasets, so the trainer comes back unchanged."""
    return trainer


def _make_report_callback():
    """`RayTrainReportCallback` (ref: fine_tune_llama_ray.py:8): a transformers TrainerCallback that forwards the
    Trainer's logs, and the checkpoint it just saved, to `train.report`. Built lazily: transformers is only needed by
    jobs that use it."""
    from transformers.trainer_callback import TrainerCallback

    class RayTrainReportCallback(TrainerCallback):
        CHECKPOINT_NAME = "checkpoint"

        def __init__(self):
            self.metrics = {}

        def on_log(self, args, state, control, logs=None, **kwargs):
            self.metrics.update(logs or {})

        def on_save(self, args, state, control, **kwargs):
            ckpt_dir = os.path.join(args.output_dir, f"checkpoint-{state.global_step}")
            ckpt = None
            if os.path.isdir(ckpt_dir):
                tmp = tempfile.mkdtemp()
                shutil.copytree(ckpt_dir, os.path.join(tmp, self.CHECKPOINT_NAME))
                ckpt = Checkpoint.from_directory(tmp)
            report({**self.metrics, "step": state.global_step, "epoch": state.epoch}, checkpoint=ckpt)

    return RayTrainReportCallback


class _HFTransformersNS(types.SimpleNamespace):
    def __getattr__(self, name):
        if name == "RayTrainReportCallback":
            cls = _make_report_callback()
            setattr(self, name, cls)
            return cls
        raise AttributeError(name)


# `train.huggingface.transformers.*`, as in `from ray.train.huggingface.transformers import RayTrainReportCallback, prepare_trainer`
huggingface = types.SimpleNamespace(transformers=_HFTransformersNS(prepare_trainer=_prepare_trainer))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(local_rank, world, port, loop, config, backend, use_gpu, storage_dir, ckpt_cfg, queue):
    global _ctx
    os.environ.update(RANK=str(local_rank), LOCAL_RANK=str(local_rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    _run_worker(local_rank, world, local_rank, loop, config, backend, use_gpu, storage_dir, ckpt_cfg, queue, True)


def _run_worker(rank, world, local_rank, loop, config, backend, use_gpu, storage_dir, ckpt_cfg, queue, own_pg):
    global _ctx
    _ddp._use_gpu = bool(use_gpu)
    if use_gpu:
        _t.cuda.set_device(local_rank % _t.cuda.device_count())
    if own_pg and world > 1 and not dist.is_initialized():
        kw = {}
        if backend == "nccl":
            kw["device_id"] = _t.device("cuda", local_rank % _t.cuda.device_count())
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    _ctx = _Context(rank, world, local_rank, storage_dir if rank == 0 else None, ckpt_cfg)
    err = None
    try:
        if config is None:
            loop()
        else:
            loop(config)
    except BaseException as e:  # noqa: BLE001 - reported to the driver, then re-raised
        err = e
    if queue is not None:
        queue.put((rank, _ctx.reports, repr(err) if err else None))
    if own_pg and dist.is_initialized():
        try:
            dist.destroy_process_group()
        except Exception:
            pass
    if err:
        raise err
    return _ctx.reports


class TorchTrainer:
    def __init__(self, train_loop_per_worker: Callable, *, train_loop_config: Optional[dict] = None,
                 torch_config: Optional[TorchConfig] = None, scaling_config: Optional[ScalingConfig] = None,
                 run_config: Optional[RunConfig] = None, datasets=None, dataset_config=None, metadata=None,
                 resume_from_checkpoint=None):
        self.loop, self.config = train_loop_per_worker, train_loop_config
        self.scaling = scaling_config or ScalingConfig()
        self.run = run_config or RunConfig()
        self.torch_config = torch_config or TorchConfig()

    def fit(self) -> Result:
        if importlib.util.find_spec("ray") is not None and os.environ.get("B2_USE_RAY") == "1":
            return self._fit_with_ray()
        sc = self.scaling
        use_gpu = bool(sc.use_gpu)
        backend = self.torch_config.backend or ("nccl" if use_gpu else "gloo")
        world = int(sc.num_workers)
        if use_gpu and world > torch_cuda_count():
            raise RuntimeError(f"ScalingConfig asks for {world} GPU workers but this box has {torch_cuda_count()} GPUs")
        storage_dir = None
        if self.run.storage_path or self.run.name:
            base = self.run.storage_path or tempfile.gettempdir()
            try:
                os.makedirs(base, exist_ok=True)
            except OSError:
                base = tempfile.gettempdir()
            storage_dir = os.path.join(base, self.run.name or "b2_run")
            os.makedirs(storage_dir, exist_ok=True)
        ckpt_cfg = self.run.checkpoint_config
        if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) == world:
            # already launched one-process-per-GPU (torchrun): this process IS worker RANK
            reports = _run_worker(int(os.environ["RANK"]), world, int(os.environ.get("LOCAL_RANK", 0)), self.loop,
                                  self.config, backend, use_gpu, storage_dir, ckpt_cfg, None, not dist.is_initialized())
            return _result_from(reports, storage_dir)
        if world == 1:
            reports = _run_worker(0, 1, 0, self.loop, self.config, backend, use_gpu, storage_dir, ckpt_cfg, None, False)
            return _result_from(reports, storage_dir)
        ctx = mp.get_context("spawn")
        queue = ctx.SimpleQueue()
        port = _free_port()
        procs = mp.start_processes(_worker, args=(world, port, self.loop, self.config, backend, use_gpu, storage_dir,
                                                  ckpt_cfg, queue), nprocs=world, join=False, start_method="spawn")
        # Collect one report list per worker, without ever blocking on a worker that can no longer post one (killed by
        # the OOM killer, a segfault in native code, a failed import in the spawned interpreter): the queue is polled
        # while the process context is checked; ProcessContext.join raises as soon as any worker has failed.
        results = {}
        deadline = _time.monotonic() + float(self.torch_config.timeout_s or 1800)
        try:
            while len(results) < world:
                if _time.monotonic() > deadline:
                    for pr in procs.processes:
                        if pr.is_alive():
                            pr.terminate()
                    return Result(metrics=None, path=storage_dir,
                                  error=TimeoutError(f"workers did not report within {self.torch_config.timeout_s} s"))
                if not queue.empty():
                    rank, reports, err = queue.get()
                    results[rank] = (reports, err)
                    if err:
                        break
                elif procs.join(timeout=0.2) and queue.empty():
                    break                      # every worker has exited and nothing is left to read
            while not procs.join(timeout=1.0):
                pass
        except Exception as e:  # a worker raised or died: surface it like Ray's Result.error
            return Result(metrics=None, error=e, path=storage_dir)
        missing = [r for r in range(world) if r not in results]
        if missing:
            return Result(metrics=None, error=RuntimeError(f"workers {missing} exited without reporting"), path=storage_dir)
        return _result_from(results.get(0, ([], None))[0], storage_dir)

    def _fit_with_ray(self):
        import ray.train
        import ray.train.torch
        trainer = ray.train.torch.TorchTrainer(
            self.loop, train_loop_config=self.config,
            scaling_config=ray.train.ScalingConfig(num_workers=self.scaling.num_workers, use_gpu=self.scaling.use_gpu,
                                                   resources_per_worker=self.scaling.resources_per_worker),
            torch_config=ray.train.torch.TorchConfig(backend=self.torch_config.backend))
        return trainer.fit()


def torch_cuda_count():
    return _t.cuda.device_count() if _t.cuda.is_available() else 0


def _result_from(reports, storage_dir):
    last = reports[-1] if reports else None
    ckpt = None
    for r in reversed(reports or []):
        if "_checkpoint_path" in r and os.path.isdir(r["_checkpoint_path"]):
            ckpt = Checkpoint(r["_checkpoint_path"])
            break
    metrics = {k: v for k, v in last.items() if not k.startswith("_")} if last else None
    return Result(metrics=metrics, checkpoint=ckpt, path=storage_dir, metrics_history=reports or [])
