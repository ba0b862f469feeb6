"""Host-side logic that needs no GPU: flat storage, optimizer bookkeeping, the ray.train-shaped trainer shim with
a world_size-2 gloo run (the reference's use_gpu=False path), data-loader sharding."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.nn as nn


def test_flatten_keeps_parameters_and_values():
    from b200_train.flat import ALIGN, buffers_of, flatten_params
    torch.manual_seed(0)
    m = nn.Sequential(nn.Linear(5, 7), nn.LayerNorm(7), nn.Linear(7, 3))
    before = {n: p.detach().clone() for n, p in m.named_parameters()}
    ids = [id(p) for p in m.parameters()]
    bufs = flatten_params(list(m.parameters()))
    assert len(bufs) == 1 and bufs[0].numel % ALIGN == 0
    assert [id(p) for p in m.parameters()] == ids                       # same Parameter objects
    for n, p in m.named_parameters():
        assert torch.equal(p, before[n])
        assert p.grad is None                                            # as in torch: no gradient before a backward
        assert p._b2_flat[1] % ALIGN == 0
    # the first backward's gradients LAND in the flat gradient buffer (copied in by the buffer's hook, not added to a
    # zero-filled buffer) ...
    x = torch.randn(4, 5)
    ref = [g.clone() for g in torch.autograd.grad(m(x).sum(), list(m.parameters()))]
    buf = bufs[0]
    buf.grad.fill_(7.0)                                                 # stale contents must not leak into the result
    m(x).sum().backward()
    for i, p in enumerate(buf.params):
        assert p.grad.data_ptr() == buf.grad_view(i).data_ptr()
        assert torch.equal(p.grad, ref[i])
    # ... and a second backward accumulates in place, like any torch gradient
    m(x).sum().backward()
    for i, p in enumerate(buf.params):
        assert p.grad is buf.grad_view(i) and torch.allclose(p.grad, 2 * ref[i])
    # zero_grad(set_to_none=True) drops them (nothing is zero-filled) and the next backward lands again
    buf.zero_grad()
    assert all(p.grad is None for p in buf.params)
    m(x).sum().backward()
    assert all(torch.equal(p.grad, r) for p, r in zip(buf.params, ref))
    # a foreign .grad is adopted by copy; a parameter without a gradient contributes a zeroed slice
    p0, p1 = buf.params[0], buf.params[1]
    p0.grad = torch.ones_like(p0)
    p1.grad = None
    assert buf.install_grads() == 1 and torch.equal(buf.grad_view(0), torch.ones_like(p0))
    assert p1.grad is buf.grad_view(1) and float(p1.grad.abs().sum()) == 0.0
    again = flatten_params(list(m.parameters()))
    assert again == bufs                                                # idempotent
    assert buffers_of(list(m.parameters()))[0] == bufs


def test_gradient_landing_queue_logic(monkeypatch):
    """The bookkeeping around b2_land_copy, on the CPU with the kernel replaced by a plain copy loop: gradients wait in
    their own allocations until the byte threshold, a flush moves them all and re-points .grad, a second micro-batch adds
    into a gradient that still waits (queued once), zero_grad() forgets the queue, install_grads() flushes."""
    from b200_train import flat as F
    from b200_train import ops
    calls = []

    def fake_land_copy(srcs, dst, offs):
        calls.append(len(srcs))
        for t, o in zip(srcs, offs):
            dst[o:o + t.numel()] = t.reshape(-1)

    monkeypatch.setattr(ops, "land_copy", fake_land_copy)
    monkeypatch.setattr(F, "_LAND_FLUSH_BYTES", 1 << 30)          # nothing flushes by size
    torch.manual_seed(0)
    m = nn.Sequential(nn.Linear(5, 7), nn.LayerNorm(7), nn.Linear(7, 3))
    buf = F.flatten_params(list(m.parameters()))[0]
    monkeypatch.setattr(buf, "_can_queue", lambda g, p: True)
    x = torch.randn(4, 5)
    ref = [g.clone() for g in torch.autograd.grad(m(x).sum(), list(m.parameters()))]
    buf.grad.fill_(9.0)
    m(x).sum().backward()
    assert not calls and sorted(buf._stolen) == list(range(len(buf.params)))      # all of them wait, in their own memory
    assert all(p.grad is not buf.grad_view(i) and torch.equal(p.grad, ref[i]) for i, p in enumerate(buf.params))
    m(x).sum().backward()                                         # second micro-batch: in-place adds, queued only once
    assert len(buf._stolen) == len(set(buf._stolen)) == len(buf.params)
    buf.install_grads()                                           # what clip / step / the bucket launch do first
    assert calls == [len(buf.params)] and not buf._stolen
    for i, p in enumerate(buf.params):
        assert p.grad is buf.grad_view(i) and torch.allclose(p.grad, 2 * ref[i])
    # a small threshold flushes while backward is still running
    monkeypatch.setattr(F, "_LAND_FLUSH_BYTES", 100)
    buf.zero_grad()
    del calls[:]
    m(x).sum().backward()
    assert len(calls) >= 2
    buf.install_grads()
    assert all(torch.equal(p.grad, r) for p, r in zip(buf.params, ref))
    # gradients dropped while some still wait: nothing of them may be copied later
    monkeypatch.setattr(F, "_LAND_FLUSH_BYTES", 1 << 30)
    buf.zero_grad()
    m(x).sum().backward()
    buf.zero_grad()
    del calls[:]
    assert not buf._stolen and all(p.grad is None for p in buf.params)
    buf.flush_landing()
    assert not calls


def test_tied_weights_are_flattened_once():
    from b200_train.flat import flatten_params
    emb = nn.Embedding(11, 4)
    head = nn.Linear(4, 11, bias=False)
    head.weight = emb.weight
    bufs = flatten_params(list(emb.parameters()) + list(head.parameters()))
    assert len(bufs[0].params) == 1 and head.weight is emb.weight


def test_adamw_bookkeeping_matches_torch_layout():
    import b200_train
    m = nn.Sequential(nn.Linear(6, 6), nn.LayerNorm(6))
    decay = [m[0].weight]
    no_decay = [m[0].bias, m[1].weight, m[1].bias]
    opt = b200_train.AdamW([{"params": decay, "weight_decay": 0.1}, {"params": no_decay, "weight_decay": 0.0}], lr=1e-3)
    ref = torch.optim.AdamW([{"params": [p.detach().clone().requires_grad_() for p in decay], "weight_decay": 0.1},
                             {"params": [p.detach().clone().requires_grad_() for p in no_decay], "weight_decay": 0.0}],
                            lr=1e-3)
    assert [set(g) - {"params"} for g in opt.param_groups] == [set(g) - {"params"} for g in ref.param_groups] or \
        all(k in opt.param_groups[0] for k in ("lr", "betas", "eps", "weight_decay"))
    bufs = opt.flat_buffers()
    table = opt._chunk_group[id(bufs[0])]
    buf = bufs[0]
    for p, o in zip(buf.params, buf.offsets):
        want = 0 if any(p is q for q in decay) else 1
        assert int(table[o // 64]) == want
    sd = opt.state_dict()
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"}
    assert sd["state"][0]["exp_avg"].shape == m[0].weight.shape
    # LambdaLR drives group["lr"] exactly as with torch's optimizer
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 0.5)
    assert opt.param_groups[0]["lr"] == pytest.approx(5e-4)
    # there is no CPU arithmetic behind step(): it must fail loudly, not fall back
    m(torch.randn(2, 6)).sum().backward()
    with pytest.raises(b200_train.B2Error):
        opt.step()
    with pytest.raises(b200_train.B2Error):
        b200_train.AdamW(m.parameters(), amsgrad=True)
    with pytest.raises(ValueError):
        b200_train.AdamW(m.parameters(), lr=-1.0)
    # once the flat layout exists a late group would never be updated: refused rather than ignored
    with pytest.raises(b200_train.B2Error):
        opt.add_param_group({"params": [nn.Parameter(torch.zeros(3))]})


def test_loss_modules_refuse_cpu_tensors():
    import b200_train
    crit = b200_train.CrossEntropyLoss()
    with pytest.raises(b200_train.B2Error):
        crit(torch.randn(4, 5), torch.tensor([0, 1, 2, 3]))
    with pytest.raises(b200_train.B2Error):
        b200_train.CrossEntropyLoss(label_smoothing=0.1)
    with pytest.raises(b200_train.B2Error):
        b200_train.fused_lm_head_ce(torch.randn(2, 3, 8), torch.randn(10, 8), torch.zeros(2, 3, dtype=torch.long))


def test_prepare_model_single_worker_returns_bare_module():
    import b200_train
    m = nn.Linear(3, 3)
    out = b200_train.prepare_model(m, move_to_device=False)
    assert out is m and not hasattr(out, "module")        # same as Ray with world_size == 1


def _cpu_loop(config):
    """Shape of the reference's train_loop_per_worker (ref: pytorch_llm_ray.py:122-313) on the use_gpu=False path."""
    import torch.distributed as dist
    from b200_train import train
    from oracle import torch_ref as TR
    ctx = train.get_context()
    rank, world = ctx.get_world_rank(), ctx.get_world_size()
    assert train.torch.get_device().type == "cpu"
    torch.set_num_threads(1)
    torch.manual_seed(100 + rank)                          # ranks start DIFFERENT: DDP must sync rank 0's weights
    V = 37
    model = TR.disable_dropout(TR.BasicLLMRestated(V, embed_dim=32, num_heads=2, num_layers=1, hidden_dim=64, max_seq_len=64))
    model = train.torch.prepare_model(model)
    assert hasattr(model, "module")
    opt = torch.optim.AdamW(model.parameters(), lr=config["lr"], weight_decay=0.01)
    crit = nn.CrossEntropyLoss()
    stream = TR.synthetic_char_stream(2000, V)
    ds = torch.utils.data.TensorDataset(*TR.windows(stream, torch.arange(64), 16))
    dl = train.torch.prepare_data_loader(torch.utils.data.DataLoader(ds, batch_size=4, shuffle=True, drop_last=True))
    assert len(dl) == 64 // world // 4
    dl.sampler.set_epoch(0)
    for step, (x, y) in enumerate(dl):
        loss, _ = TR.basic_llm_step(model, opt, None, crit, x, y, V, 1.0)
        if step == 2:
            break
    checksum = torch.stack([p.detach().double().sum() for p in model.parameters()]).sum().reshape(1)
    both = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(both, checksum)
    assert all(torch.equal(both[0], b) for b in both), "ranks diverged"
    ckpt = None
    if rank == 0:
        d = tempfile.mkdtemp()
        torch.save(model.module.state_dict(), os.path.join(d, "model.pth"))
        ckpt = train.Checkpoint.from_directory(d)
    train.report({"loss": float(loss), "checksum": float(checksum)}, checkpoint=ckpt)


def test_trainer_shim_two_gloo_workers(tmp_path):
    from b200_train.train import CheckpointConfig, RunConfig, ScalingConfig, TorchConfig, TorchTrainer
    trainer = TorchTrainer(_cpu_loop, train_loop_config={"lr": 3e-4},
                           scaling_config=ScalingConfig(num_workers=2, use_gpu=False),
                           run_config=RunConfig(name="t", storage_path=str(tmp_path),
                                                checkpoint_config=CheckpointConfig(num_to_keep=1,
                                                                                   checkpoint_score_attribute="loss",
                                                                                   checkpoint_score_order="min")),
                           torch_config=TorchConfig(backend="gloo"))
    result = trainer.fit()
    assert result.error is None
    assert np.isfinite(result.metrics["loss"]) and "checksum" in result.metrics
    assert result.checkpoint is not None and os.path.exists(os.path.join(result.checkpoint.path, "model.pth"))


def _dying_loop(config):
    import os as _os
    from b200_train import train
    if train.get_context().get_world_rank() == 1:
        _os._exit(3)                      # dies without reporting anything (OOM kill, segfault in native code, ...)
    import time as _time
    _time.sleep(600)                      # the survivor would wait in a collective forever


def test_trainer_shim_surfaces_a_worker_that_dies_without_reporting():
    """fit() must neither hang on the queue nor return while workers still run."""
    import time
    from b200_train.train import ScalingConfig, TorchConfig, TorchTrainer
    t0 = time.time()
    res = TorchTrainer(_dying_loop, train_loop_config={}, scaling_config=ScalingConfig(num_workers=2, use_gpu=False),
                       torch_config=TorchConfig(backend="gloo")).fit()
    assert res.error is not None and res.metrics is None
    assert time.time() - t0 < 120


def test_front_parameters_get_their_own_bucket():
    """Embeddings sit at the front of the flat buffer and receive their gradient last: they must not hold back the
    bucket of the first transformer layers."""
    from b200_train.ddp import plan_buckets
    offsets = [0, 1000, 1100, 1200, 1300, 1400]
    plan = plan_buckets(offsets, 1500, 2, cap_bytes=400, front_params=2)
    assert plan[-1] == (0, 1100, [1, 0])                                  # the two leading parameters, launched last
    assert sum(n for _, n, _ in plan) == 1500 and plan[0][0] + plan[0][1] == 1500
    assert sorted(i for _, _, m in plan for i in m) == list(range(6))
    for (s0, n0, _), (s1, n1, _) in zip(plan, plan[1:]):
        assert s1 + n1 == s0                                              # contiguous, back to front
    assert plan_buckets(offsets, 1500, 2, cap_bytes=1 << 40, front_params=0) == [(0, 1500, [5, 4, 3, 2, 1, 0])]


def test_trainer_shim_single_worker_inline():
    from b200_train import train
    seen = {}

    def loop(cfg):
        ctx = train.get_context()
        seen["rank"], seen["world"], seen["cfg"] = ctx.get_world_rank(), ctx.get_world_size(), cfg
        train.report({"loss": 1.5})

    r = train.TorchTrainer(loop, train_loop_config={"a": 1}, scaling_config=train.ScalingConfig(num_workers=1)).fit()
    assert seen == {"rank": 0, "world": 1, "cfg": {"a": 1}} and r.metrics == {"loss": 1.5}


def test_patch_causal_lm_keeps_state_dict_and_parameters():
    import b200_train
    import workloads as WL
    m = WL.make_gpt2(dtype=torch.bfloat16, seed=0, vocab_size=101, n_embd=32, n_layer=1, n_head=2, n_positions=16)
    keys, n_params = list(m.state_dict().keys()), len(list(m.parameters()))
    b200_train.patch_causal_lm(m)
    assert list(m.state_dict().keys()) == keys and len(list(m.parameters())) == n_params
    # without labels the stock forward still answers (logits present)
    out = m(input_ids=torch.zeros(1, 4, dtype=torch.long))
    assert out.logits is not None and out.logits.shape == (1, 4, 101)


def test_bucket_plan_covers_the_buffer_in_reverse_order():
    from b200_train.ddp import plan_buckets
    from b200_train.flat import TOTAL_MULTIPLE
    offsets = [0, 64, 1088, 1152, 9344]                  # 64-aligned parameter starts
    total = TOTAL_MULTIPLE                                # padded buffer length
    plan = plan_buckets(offsets, total, 2, cap_bytes=4096)
    # contiguous, disjoint, covering [0, total), launched from the end towards the front
    assert plan[0][0] + plan[0][1] == total and plan[-1][0] == 0
    for (s0, n0, _), (s1, n1, _) in zip(plan, plan[1:]):
        assert s1 + n1 == s0
    assert sorted(i for _, _, m in plan for i in m) == list(range(len(offsets)))
    assert all(s % 64 == 0 and n % 8 == 0 for s, n, _ in plan)
    # every bucket but the front one reaches the cap
    assert all(n * 2 >= 4096 for _, n, _ in plan[:-1])
    # one bucket when the cap is huge (the no-overlap / sharded path)
    assert plan_buckets(offsets, total, 2, 1 << 62) == [(0, total, [4, 3, 2, 1, 0])]


def test_comm_hook_state_refuses_to_exist_without_group_or_gpu():
    """The communication hook for stock DDP has no CPU path and needs the process group."""
    import inspect
    import b200_train
    with pytest.raises(b200_train.B2Error):
        b200_train.PeerAllreduceState()                       # no process group
    # torch's register_comm_hook checks the hook's annotations; ours carries none, and takes (state, bucket)
    sig = inspect.signature(b200_train.peer_allreduce_hook)
    assert list(sig.parameters) == ["state", "bucket"]
    assert all(p.annotation is inspect.Parameter.empty for p in sig.parameters.values())
    assert sig.return_annotation is inspect.Signature.empty


def test_hf_report_callback_and_prepare_trainer(tmp_path):
    """`from ray.train.huggingface.transformers import RayTrainReportCallback, prepare_trainer`
    (ref: ray-jobs/fine_tune_llama_ray.py:8) has a counterpart under b200_train.train; the callback forwards the HF
    Trainer's logs and the checkpoint it just saved to train.report (driven by hand here: a real Trainer needs accelerate,
    which is not installed)."""
    from types import SimpleNamespace
    from b200_train import train
    hf = train.huggingface.transformers
    sentinel = object()
    assert hf.prepare_trainer(sentinel) is sentinel
    cb = hf.RayTrainReportCallback()
    assert hf.RayTrainReportCallback is type(cb)                    # built once, then cached on the namespace
    out = tmp_path / "out"
    (out / "checkpoint-7").mkdir(parents=True)
    (out / "checkpoint-7" / "weights.bin").write_bytes(b"abc")
    store = tmp_path / "store"
    store.mkdir()
    old = train._ctx
    train._ctx = train._Context(0, 1, 0, str(store), train.CheckpointConfig(num_to_keep=1))
    try:
        args, state = SimpleNamespace(output_dir=str(out)), SimpleNamespace(global_step=7, epoch=0.5)
        cb.on_log(args, state, None, logs={"loss": 1.25})
        cb.on_save(args, state, None)
        rep = train._ctx.reports[-1]
        assert rep["loss"] == 1.25 and rep["step"] == 7
        assert os.path.isfile(os.path.join(rep["_checkpoint_path"], "checkpoint", "weights.bin"))
    finally:
        train._ctx = old


def test_hf_cosine_schedule_drives_our_param_groups_like_torchs():
    """ref: ray-jobs/fine_tune_config.json:18,20 — `cosine` schedule, warm-up ratio 0.03, built by HF on top of the
    optimizer's param_groups. The schedule must see (and mutate) ours exactly as it does torch.optim.AdamW's."""
    from transformers import get_scheduler
    import b200_train
    steps, warm = 200, int(0.03 * 200)
    lrs = {}
    for name, cls in (("ours", b200_train.AdamW), ("torch", torch.optim.AdamW)):
        m = nn.Linear(4, 4)
        opt = cls([{"params": [m.weight], "weight_decay": 1e-3}, {"params": [m.bias], "weight_decay": 0.0}], lr=2e-4)
        sched = get_scheduler("cosine", optimizer=opt, num_warmup_steps=warm, num_training_steps=steps)
        seq = []
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")      # "scheduler before optimizer": no optimizer arithmetic happens in this test
            for _ in range(steps):
                seq.append([g["lr"] for g in opt.param_groups])
                sched.step()
        lrs[name] = seq
    assert lrs["ours"] == lrs["torch"]
    assert lrs["ours"][0] == [0.0, 0.0] and max(v[0] for v in lrs["ours"]) == pytest.approx(2e-4)


def test_reference_arm_scrubs_the_torchrun_environment(monkeypatch):
    """bench.py's CPU legs start their own two gloo workers; under torchrun the parent's RANK / WORLD_SIZE / TORCHELASTIC_*
    variables (TORCHELASTIC_USE_AGENT_STORE makes an env:// rendezvous wait for the agent's store) must not leak into them,
    and must be back afterwards."""
    import bench
    for k, v in {"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": "29999",
                 "TORCHELASTIC_USE_AGENT_STORE": "True", "TORCHELASTIC_RUN_ID": "x", "OMP_NUM_THREADS": "1"}.items():
        monkeypatch.setenv(k, v)
    seen = []

    class FakeTR:
        @staticmethod
        def cpu_gloo_small_gpt_baseline(steps, warmup, world, threads):
            seen.append({k for k in os.environ if k.startswith("TORCHELASTIC_") or k in ("RANK", "WORLD_SIZE", "MASTER_PORT", "OMP_NUM_THREADS")})
            return dict(tokens_per_s=1.0, ms_per_step=1.0, sample="fake")

    out = bench.configs0_gloo_baseline(FakeTR)
    assert seen == [set(), set()]
    assert out["threads_1_per_worker"]["value"] == 1.0 and out["workers"] == 2
    assert os.environ["TORCHELASTIC_USE_AGENT_STORE"] == "True" and os.environ["RANK"] == "0"
