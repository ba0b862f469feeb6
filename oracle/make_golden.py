"""Generates tests/golden/*.npz by running the REFERENCE's own classes and call sequence on stock torch (CPU).

Run in the build container only (it reads /root/reference, which does not exist on the GPU box):
    python oracle/make_golden.py
The reference has no tests or golden vectors of its own, so these files are what pins the oracle:
  - BasicLLM / TextDataset are imported from /root/reference/ray-jobs/pytorch_llm_ray.py (with the import-only `ray`
    stub in oracle/ray_stub) and driven through the loop body ref :272-279;
  - the HF causal-LM loss, clip_grad_norm_, AdamW and DDP(gloo) results come from the installed torch 2.11 /
    transformers 5.5 (the reference pins torch 2.3.0 / transformers 4.50.0 — version skew recorded in DESIGN.md).
Everything is seeded; outputs are small (a few hundred KB in total).
"""
import importlib.util
import os
import socket
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference/ray-jobs/pytorch_llm_ray.py"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(HERE, "ray_stub"))

from oracle import torch_ref as TR  # noqa: E402


def load_reference_module():
    spec = importlib.util.spec_from_file_location("ref_pytorch_llm_ray", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


TINY = dict(embed_dim=32, num_heads=2, num_layers=2, hidden_dim=64, max_seq_len=64)
V_TINY = 37


def npz(name, **arrays):
    np.savez_compressed(os.path.join(OUT, name), **{k: np.asarray(v) for k, v in arrays.items()})
    print("wrote", name, {k: np.asarray(v).shape for k, v in arrays.items()})


def gen_basicllm(ref):
    torch.manual_seed(0)
    model = ref.BasicLLM(vocab_size=V_TINY, **TINY)
    mine = TR.BasicLLMRestated(V_TINY, **TINY)
    mine.load_state_dict(model.state_dict())
    model.eval(); mine.eval()
    stream = TR.synthetic_char_stream(400, V_TINY, seed=1234)
    ds = ref.TextDataset(stream, 16)
    x = torch.stack([ds[i][0] for i in (0, 5, 17, 100)]); y = torch.stack([ds[i][1] for i in (0, 5, 17, 100)])
    xw, yw = TR.windows(stream, torch.tensor([0, 5, 17, 100]), 16)
    assert torch.equal(x, xw) and torch.equal(y, yw), "windows() must reproduce TextDataset"
    with torch.no_grad():
        a, b = model(x), mine(x)
        feats = mine.features(x)
    assert torch.equal(a, b), "restated BasicLLM must be bit-identical to the reference class"
    loss = nn.CrossEntropyLoss()(a.reshape(-1, V_TINY), y.reshape(-1))
    # three training steps of the reference loop body (dropout off for determinism across torch builds)
    TR.disable_dropout(model)
    model.train()
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4, weight_decay=0.01)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, TR.lr_lambda_factory(20))
    crit = nn.CrossEntropyLoss()
    w0 = model.fc_out.weight.detach().clone(); b0 = model.fc_out.bias.detach().clone()
    losses, norms, lrs = [], [], []
    for s in range(3):
        xs, ys = TR.windows(stream, torch.tensor([3 * s, 3 * s + 40, 3 * s + 90, 3 * s + 200]), 16)
        lrs.append(opt.param_groups[0]["lr"])
        l, n = TR.basic_llm_step(model, opt, sched, crit, xs, ys, V_TINY, 1.0)
        losses.append(float(l)); norms.append(float(n))
    st = opt.state[model.fc_out.weight]
    npz("basicllm_tiny.npz", x=x.numpy(), y=y.numpy(), logits=a.numpy(), features=feats.numpy(),
        fc_w=w0.numpy(), fc_b=b0.numpy(), loss0=loss.item(), losses=losses, norms=norms, lrs=lrs,
        fc_w_after=model.fc_out.weight.detach().numpy(), fc_b_after=model.fc_out.bias.detach().numpy(),
        fc_w_exp_avg=st["exp_avg"].numpy(), fc_w_exp_avg_sq=st["exp_avg_sq"].numpy(),
        lr_lambda=[TR.lr_lambda_factory(20)(s) for s in range(25)])


def gen_ce_kats():
    g = torch.Generator().manual_seed(7)
    x = torch.randn(6, 11, generator=g) * 2
    y = torch.tensor([3, -100, 0, 10, -100, 7])
    F = torch.nn.functional
    x.requires_grad_(True)
    lm = F.cross_entropy(x, y); lm.backward()
    gm = x.grad.clone(); x.grad = None
    ls = F.cross_entropy(x, y, reduction="sum")
    # HF shift on 2x5
    from transformers.loss.loss_utils import ForCausalLMLoss
    xl = torch.randn(2, 5, 13, generator=g)
    yl = torch.tensor([[1, 2, 3, 4, 5], [6, 7, -100, -100, 9]])
    hf_mean = ForCausalLMLoss(xl, yl, 13)
    hf_items = ForCausalLMLoss(xl, yl, 13, num_items_in_batch=torch.tensor(9))
    all_ign_mean = F.cross_entropy(x.detach(), torch.full((6,), -100))
    all_ign_sum = F.cross_entropy(x.detach(), torch.full((6,), -100), reduction="sum")
    npz("ce_kat.npz", x=x.detach().numpy(), y=y.numpy(), loss_mean=lm.item(), loss_sum=ls.item(), grad_mean=gm.numpy(),
        hf_logits=xl.numpy(), hf_labels=yl.numpy(), hf_mean=hf_mean.item(), hf_items9=hf_items.item(),
        all_ignored_mean=all_ign_mean.item(), all_ignored_sum=all_ign_sum.item())


def gen_adamw():
    out = {}
    for tag, dt in (("f32", torch.float32), ("bf16", torch.bfloat16)):
        g = torch.Generator().manual_seed(11)
        n = 257
        p = nn.Parameter((torch.randn(n, generator=g)).to(dt))
        out[f"{tag}_p0"] = p.detach().float().numpy().copy()
        opt = torch.optim.AdamW([p], lr=2e-4 if dt == torch.bfloat16 else 3e-4, weight_decay=1e-3 if dt == torch.bfloat16 else 0.01,
                                foreach=False)
        grads = []
        for s in range(10):
            gr = (torch.randn(n, generator=g) * 0.01).to(dt)
            grads.append(gr.float().numpy().copy())
            p.grad = gr.clone()
            opt.step()
        st = opt.state[p]
        out[f"{tag}_grads"] = np.stack(grads)
        out[f"{tag}_p"] = p.detach().float().numpy()
        out[f"{tag}_m"] = st["exp_avg"].float().numpy()
        out[f"{tag}_v"] = st["exp_avg_sq"].float().numpy()
    npz("adamw_traj.npz", **out)


def gen_clip():
    g = torch.Generator().manual_seed(5)
    ps = [nn.Parameter(torch.zeros(s)) for s in ((7, 3), (19,), (4, 4, 2))]
    grads = [torch.randn(p.shape, generator=g) for p in ps]
    res = {}
    for tag, mx in (("clip", 0.3), ("noclip", 100.0)):
        for p, gr in zip(ps, grads):
            p.grad = gr.clone()
        total = torch.nn.utils.clip_grad_norm_(ps, mx)
        res[f"{tag}_total"] = total.item()
        for i, p in enumerate(ps):
            res[f"{tag}_g{i}"] = p.grad.numpy().copy()
    for i, gr in enumerate(grads):
        res[f"g{i}"] = gr.numpy()
    npz("clip_kat.npz", **res)


def _ddp_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, os.path.join(HERE, "ray_stub"))
    ref = load_reference_module()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    base = ref.BasicLLM(vocab_size=V_TINY, **TINY)
    TR.disable_dropout(base)
    stream = TR.synthetic_char_stream(400, V_TINY, seed=1234)
    x, y = TR.windows(stream, torch.tensor([rank * 50, rank * 50 + 7]), 16)
    # local (unsynchronised) gradient of this rank
    crit = nn.CrossEntropyLoss()
    crit(base(x).reshape(-1, V_TINY), y.reshape(-1)).backward()
    local = torch.cat([p.grad.reshape(-1) for p in base.parameters()]).clone()
    for p in base.parameters():
        p.grad = None
    model = nn.parallel.DistributedDataParallel(base)                 # what prepare_model installs
    crit(model(x).reshape(-1, V_TINY), y.reshape(-1)).backward()
    synced = torch.cat([p.grad.reshape(-1) for p in base.parameters()]).clone()
    q.put((rank, local.numpy(), synced.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def gen_ddp():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = mp.start_processes(_ddp_worker, args=(2, port, q), nprocs=2, join=False, start_method="spawn")
    got = {}
    while len(got) < 2:
        r, l, s = q.get(); got[r] = (l, s)
    procs.join()
    assert np.array_equal(got[0][1], got[1][1])
    npz("ddp_gloo2.npz", local0=got[0][0][:4096], local1=got[1][0][:4096], synced=got[0][1][:4096])


def gen_hf_tiny():
    torch.manual_seed(0)
    model = TR.make_gpt2(dtype=torch.bfloat16, seed=0, vocab_size=503, n_embd=64, n_layer=2, n_head=2, n_positions=32,
                         resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
    model.train()
    ids, labels = TR.synthetic_hf_batch(2, 16, 503, seed=1234)
    hidden = model.transformer(input_ids=ids)[0].detach().requires_grad_(True)
    W = model.lm_head.weight.detach().clone().requires_grad_(True)
    logits = torch.nn.functional.linear(hidden, W)
    from transformers.loss.loss_utils import ForCausalLMLoss
    loss = ForCausalLMLoss(logits, labels, 503)
    loss.backward()
    full = model(input_ids=ids, labels=labels).loss
    assert abs(full.item() - loss.item()) < 1e-6, (full.item(), loss.item())
    n_valid = int((labels[:, 1:] != -100).sum())
    npz("hf_tiny_gpt2.npz", ids=ids.numpy(), labels=labels.numpy(), hidden=hidden.detach().float().numpy(),
        W=W.detach().float().numpy(), loss=loss.item(), n_valid=n_valid, dH=hidden.grad.float().numpy(),
        dW=W.grad.float().numpy())


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    torch.use_deterministic_algorithms(True)
    ref = load_reference_module()
    gen_basicllm(ref)
    gen_ce_kats()
    gen_adamw()
    gen_clip()
    gen_ddp()
    gen_hf_tiny()
