"""End-to-end parity on the GPU: the drop-in surface (prepare_model / CrossEntropyLoss / fused LM-head loss /
clip_grad_norm_ / AdamW) against the reference's stock call sequence (oracle/torch_ref.py), the committed golden
fixtures, and size-independent properties at BASELINE.json's full sizes.

Stated tolerances: masking / token counts bit-exact; fp32 path — loss rtol 1e-5, parameters atol 2e-6 after 3 steps
(elements with a mathematically zero gradient excepted); bf16 path — loss atol 5e-3, parameters within 1 bf16 ulp (at
operand magnitude) for 99.5 % of the model's elements, never more than 2 ulp + one Adam sign-flip (2*lr),
and in total no farther from the same step computed in fp32 than the stock bf16 path is.
"""
import copy
import ctypes
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import step_tail as O
from oracle import torch_ref as TR

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def dev():
    return torch.device("cuda", 0)


@pytest.fixture(scope="module")
def b200():
    import b200_train
    return b200_train


# ------------------------------------------------------------------------------------------ golden fixtures
def test_golden_basicllm_head_loss(b200):
    """features/fc_out of the reference's own BasicLLM -> our CrossEntropyLoss == the reference model's loss."""
    g = np.load(os.path.join(G, "basicllm_tiny.npz"))
    H = torch.from_numpy(g["features"]).to(dev()).reshape(-1, 32)
    logits = torch.nn.functional.linear(H, torch.from_numpy(g["fc_w"]).to(dev()), torch.from_numpy(g["fc_b"]).to(dev()))
    y = torch.from_numpy(g["y"]).to(dev()).reshape(-1)
    loss = b200.CrossEntropyLoss()(logits, y)
    np.testing.assert_allclose(loss.item(), g["loss0"], rtol=1e-5)


def test_golden_hf_tiny_gpt2_fused_loss_and_grads(b200):
    g = np.load(os.path.join(G, "hf_tiny_gpt2.npz"))
    H = torch.from_numpy(g["hidden"]).to(dev()).to(torch.bfloat16).requires_grad_(True)
    W = torch.from_numpy(g["W"]).to(dev()).to(torch.bfloat16).requires_grad_(True)
    labels = torch.from_numpy(g["labels"]).to(dev())
    loss = b200.fused_lm_head_ce(H, W, labels)
    loss.backward()
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=0, atol=2e-3)
    from b200_train import ops
    r = ops.lmhead_ce_fwd(H.detach().reshape(-1, 64), W.detach(), labels.reshape(-1), shift_T=labels.shape[1])
    assert int(r.n_valid.item()) == int(g["n_valid"])
    for got, want in ((H.grad.float().cpu().numpy(), g["dH"]), (W.grad.float().cpu().numpy(), g["dW"])):
        err = np.abs(got - want)
        assert (err <= 2 * O.bf16_ulp(want) + 2e-6).mean() > 0.999
        assert err.max() <= 8 * O.bf16_ulp(np.abs(want).max())


# ------------------------------------------------------------------------------------------ BasicLLM loop, fp32
def test_basicllm_fp32_three_steps_match_reference_loop(b200):
    V = 283
    torch.manual_seed(0)
    ref_model = TR.disable_dropout(TR.BasicLLMRestated(V, embed_dim=64, num_heads=4, num_layers=2, hidden_dim=128,
                                                       max_seq_len=128)).to(dev())
    my_model = copy.deepcopy(ref_model)
    my_model = b200.prepare_model(my_model)                         # world 1: bare module, as with Ray
    assert my_model is not None and not hasattr(my_model, "module")
    ref_opt = torch.optim.AdamW(ref_model.parameters(), lr=3e-4, weight_decay=0.01)
    my_opt = b200.AdamW(my_model.parameters(), lr=3e-4, weight_decay=0.01)
    ref_sched = torch.optim.lr_scheduler.LambdaLR(ref_opt, TR.lr_lambda_factory(20))
    my_sched = torch.optim.lr_scheduler.LambdaLR(my_opt, TR.lr_lambda_factory(20))
    ref_crit, my_crit = nn.CrossEntropyLoss(), b200.CrossEntropyLoss()
    stream = TR.synthetic_char_stream(4000, V).to(dev())
    for s in range(3):
        idx = torch.arange(8, device=dev()) * 37 + s
        ar = torch.arange(32, device=dev())
        x, y = stream[idx[:, None] + ar], stream[idx[:, None] + ar + 1]
        l_ref, n_ref = TR.basic_llm_step(ref_model, ref_opt, ref_sched, ref_crit, x, y, V, 1.0)
        # the reference's loop body (ref :272-279) with our drop-ins
        my_opt.zero_grad(set_to_none=True)
        out = my_model(x)
        loss = my_crit(out.reshape(-1, V), y.reshape(-1))
        loss.backward()
        n_my = b200.clip_grad_norm_(my_model.parameters(), max_norm=1.0)
        my_opt.step()
        my_sched.step()
        np.testing.assert_allclose(loss.item(), l_ref.item(), rtol=1e-5)
        np.testing.assert_allclose(n_my.item(), n_ref.item(), rtol=1e-4)
    for (n, a), (_, b) in zip(my_model.named_parameters(), ref_model.named_parameters()):
        # Elements whose gradient is mathematically zero (e.g. the key bias of attention: soft-max is shift-invariant)
        # carry only rounding noise, which Adam normalises to +-lr steps of arbitrary sign in ANY implementation; they
        # are identified by the reference's own second-moment estimate and left out.
        live = (ref_opt.state[b]["exp_avg_sq"] > 1e-14).cpu().numpy()
        got, want = a.detach().cpu().numpy(), b.detach().cpu().numpy()
        np.testing.assert_allclose(got[live], want[live], rtol=0, atol=2e-6, err_msg=n)
        assert np.abs(got - want).max() <= 3 * 3e-4 * 2 + 1e-6, n
    # optimizer.state_dict() keeps torch's layout (ref :302 saves it)
    sd, rd = my_opt.state_dict(), ref_opt.state_dict()
    assert sd["state"].keys() == rd["state"].keys()
    for k in rd["state"]:
        assert float(sd["state"][k]["step"]) == float(rd["state"][k]["step"]) == 3.0
        np.testing.assert_allclose(sd["state"][k]["exp_avg"].cpu().numpy(), rd["state"][k]["exp_avg"].cpu().numpy(), atol=1e-7)


# ------------------------------------------------------------------------------------------ HF causal LM, bf16
def _tiny_gpt2():
    return TR.make_gpt2(dtype=torch.bfloat16, seed=0, vocab_size=2003, n_embd=128, n_layer=2, n_head=4, n_positions=64,
                        resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0).to(dev())


def _bf16_close(got, want, before, name, lr=2e-4, steps=1):
    """Per tensor: never more than 2 bf16 ulp (at operand magnitude) + one Adam sign-flip away. At step 1 Adam's update
    is lr*sign(g), so an element whose gradient is at the level of bf16 rounding noise can legitimately land 2*lr apart
    in any two implementations. Returns (#elements off by > 1 ulp, > 2 ulp, total) for the whole-model fractions."""
    got, want, before = (t.detach().float().cpu().numpy() for t in (got, want, before))
    ulp = O.bf16_ulp(np.maximum(np.abs(want), np.abs(before)))
    err = np.abs(got - want)
    assert (err <= 2 * ulp + 2.2 * lr * steps).all(), f"{name}: max err {err.max()}"
    return int((err > ulp).sum()), int((err > 2 * ulp).sum()), err.size


@pytest.mark.parametrize("micro_batches", [1, 2])
def test_hf_gpt2_bf16_step_matches_stock(b200, micro_batches):
    ref_model = _tiny_gpt2().train()
    my_model = b200.patch_causal_lm(copy.deepcopy(ref_model)).train()
    hi_model = copy.deepcopy(ref_model).float().train()            # same step in fp32: the yardstick for both bf16 paths
    before = {n: p.detach().clone() for n, p in ref_model.named_parameters()}
    ref_opt = torch.optim.AdamW(TR.hf_param_groups(ref_model, 1e-3), lr=2e-4)
    my_opt = b200.AdamW(TR.hf_param_groups(my_model, 1e-3), lr=2e-4)
    hi_opt = torch.optim.AdamW(TR.hf_param_groups(hi_model, 1e-3), lr=2e-4)
    batches = [TR.synthetic_hf_batch(4, 64, 2003, seed=1234 + i, device=dev()) for i in range(micro_batches)]
    l_ref, n_ref = TR.hf_causal_lm_step(ref_model, ref_opt, batches, 0.3)
    TR.hf_causal_lm_step(hi_model, hi_opt, batches, 0.3)
    # same step with the drop-ins
    num_items = sum(int((lab[..., 1:] != -100).sum()) for _, lab in batches)
    my_opt.zero_grad(set_to_none=True)
    total = 0.0
    for ids, lab in batches:
        out = my_model(input_ids=ids, labels=lab, num_items_in_batch=num_items) if micro_batches > 1 \
            else my_model(input_ids=ids, labels=lab)
        assert out.logits is None
        out.loss.backward()
        total += float(out.loss.item())
    n_my = b200.clip_grad_norm_(my_model.parameters(), 0.3)
    my_opt.step()
    np.testing.assert_allclose(total, float(l_ref), rtol=0, atol=5e-3)
    np.testing.assert_allclose(n_my.item(), float(n_ref), rtol=2e-2)
    mine_vs_hi = stock_vs_hi = 0.0
    over1 = over2 = total = s_over1 = s_over2 = 0
    for (n, a), (_, b), (_, h) in zip(my_model.named_parameters(), ref_model.named_parameters(), hi_model.named_parameters()):
        c1, c2, c = _bf16_close(a, b, before[n], n)
        over1, over2, total = over1 + c1, over2 + c2, total + c
        # how far the reference's own bf16 path is from the fp32 step, counted the same way
        ulp = O.bf16_ulp(np.maximum(np.abs(b.detach().float().cpu().numpy()), np.abs(before[n].float().cpu().numpy())))
        e = np.abs(b.detach().float().cpu().numpy() - h.detach().cpu().numpy())
        s_over1 += int((e > ulp).sum()); s_over2 += int((e > 2 * ulp).sum())
        mine_vs_hi += float((a.detach().float() - h.detach()).abs().sum())
        stock_vs_hi += float((b.detach().float() - h.detach()).abs().sum())
    # whole model: within 1 bf16 ulp of the stock path for 99.5 % of the elements (the rest are the sign-flip elements
    # described above), and we disagree with the stock path on fewer elements than the stock path disagrees with fp32
    assert over1 / total < 5e-3 and over2 / total < 3e-3, (over1, over2, total)
    assert over1 <= s_over1 and over2 <= s_over2, (over1, s_over1, over2, s_over2)
    # we are not farther from the fp32 trajectory than the reference's own bf16 path is (SURVEY.md §7 hard part 7)
    assert mine_vs_hi <= 1.02 * stock_vs_hi, (mine_vs_hi, stock_vs_hi)


def test_patch_causal_lm_llama_shape(b200):
    """Llama-shaped (untied head, GQA, RMSNorm) model through the same patch: loss equals the stock HF loss."""
    model = TR.make_llama(dtype=torch.bfloat16, seed=0, vocab_size=4099, hidden_size=128, intermediate_size=256,
                          num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2, head_dim=32,
                          tie_word_embeddings=False, max_position_embeddings=128).to(dev()).train()
    ids, labels = TR.synthetic_hf_batch(2, 96, 4099, seed=5, device=dev())
    with torch.no_grad():
        want = model(input_ids=ids, labels=labels).loss.item()
    b200.patch_causal_lm(model)
    out = model(input_ids=ids, labels=labels)
    np.testing.assert_allclose(out.loss.item(), want, rtol=0, atol=5e-3)
    out.loss.backward()
    assert model.lm_head.weight.grad is not None and torch.isfinite(model.lm_head.weight.grad).all()


# ------------------------------------------------------------------------------------------ full-size properties
def test_full_size_lmhead_properties(b200):
    """BASELINE configs[1] head shape: N = 8*1024 tokens, D = 768, V = 50257 (ragged last tile)."""
    from b200_train import ops
    N, D, V, T = 8192, 768, 50257, 1024
    g = torch.Generator(device="cpu").manual_seed(3)
    H = (torch.randn(N, D, generator=g) * 0.7).to(dev(), torch.bfloat16)
    W = (torch.randn(V, D, generator=g) * 0.05).to(dev(), torch.bfloat16)
    ids, labels = TR.synthetic_hf_batch(N // T, T, V, seed=9, device=dev())
    y = labels.reshape(-1)
    full = ops.lmhead_ce_fwd(H, W, y, shift_T=T)
    # (1) token count is exact
    assert int(full.n_valid.item()) == int((labels[:, 1:] != -100).sum())
    # (2) additivity over row blocks (different work decomposition: other split counts, other CTA assignment)
    parts = [ops.lmhead_ce_fwd(H[a:a + 2048], W, y[a:a + 2048], shift_T=T) for a in range(0, N, 2048)]
    s = sum(float(p.loss_sum.item()) for p in parts)
    np.testing.assert_allclose(float(full.loss_sum.item()), s, rtol=2e-6)
    assert sum(int(p.n_valid.item()) for p in parts) == int(full.n_valid.item())
    np.testing.assert_array_equal(torch.cat([p.lse[:2048] for p in parts]).cpu().numpy(), full.lse[:N].cpu().numpy())
    # (3) a vocabulary permutation leaves every row's lse (to fp32 rounding) and the loss unchanged
    perm = torch.randperm(V, generator=g).to(dev())
    inv = torch.empty_like(perm); inv[perm] = torch.arange(V, device=dev())
    yp = torch.where(y >= 0, inv[y.clamp(min=0)], y)
    permd = ops.lmhead_ce_fwd(H, W[perm].contiguous(), yp, shift_T=T)
    np.testing.assert_allclose(permd.lse[:N].cpu().numpy(), full.lse[:N].cpu().numpy(), rtol=0, atol=2e-5)
    np.testing.assert_allclose(float(permd.loss_mean.item()), float(full.loss_mean.item()), rtol=1e-6)
    # (4) against torch on the same device (bf16 GEMM, fp32 soft-max — the reference's own arithmetic)
    logits = (H @ W.t()).float()
    ys = torch.full_like(labels, -100); ys[:, :-1] = labels[:, 1:]
    want = torch.nn.functional.cross_entropy(logits, ys.reshape(-1), ignore_index=-100)
    np.testing.assert_allclose(float(full.loss_mean.item()), float(want.item()), rtol=0, atol=2e-3)
    del logits


def test_full_size_lmhead_backward_vs_torch(b200):
    N, D, V, T = 4096, 768, 50257, 1024
    g = torch.Generator(device="cpu").manual_seed(4)
    H0 = (torch.randn(N, D, generator=g) * 0.7).to(dev(), torch.bfloat16)
    W0 = (torch.randn(V, D, generator=g) * 0.05).to(dev(), torch.bfloat16)
    _, labels = TR.synthetic_hf_batch(N // T, T, V, seed=10, device=dev())
    H, W = H0.clone().requires_grad_(True), W0.clone().requires_grad_(True)
    b200.fused_lm_head_ce(H.view(N // T, T, D), W, labels, chunk_rows=2048).backward()      # 2 chunks: fp32 dW accumulation
    Hr, Wr = H0.clone().requires_grad_(True), W0.clone().requires_grad_(True)
    ys = torch.full_like(labels, -100); ys[:, :-1] = labels[:, 1:]
    torch.nn.functional.cross_entropy((Hr @ Wr.t()).float(), ys.reshape(-1), ignore_index=-100).backward()
    for got, want, name in ((H.grad, Hr.grad, "dH"), (W.grad, Wr.grad, "dW")):
        got, want = got.float(), want.float()
        rel = (got - want).norm() / want.norm()
        assert rel < 1e-2, f"{name}: relative Frobenius error {rel:.3e}"
        assert torch.isfinite(got).all()
    # rows whose label is ignored contribute exactly nothing to dH
    ign = (ys.reshape(-1) == -100)
    assert float(H.grad[ign].abs().max()) == 0.0


def test_full_size_kept_logits_route_gives_the_recompute_route_bits(b200, monkeypatch):
    """BASELINE configs[1] head shape through the public fused loss: keeping the forward's bf16 logits for the backward
    (one chunk, the default) and recomputing them (B2_KEEP_LOGITS=0) produce the same loss, dH and dW bits."""
    from b200_train import loss as L
    N, D, V, T = 8192, 768, 50257, 1024
    g = torch.Generator(device="cpu").manual_seed(5)
    H0 = (torch.randn(N, D, generator=g) * 0.7).to(dev(), torch.bfloat16)
    W0 = (torch.randn(V, D, generator=g) * 0.05).to(dev(), torch.bfloat16)
    _, labels = TR.synthetic_hf_batch(N // T, T, V, seed=11, device=dev())
    out = {}
    for keep in (True, False):
        monkeypatch.setattr(L, "_KEEP_LOGITS", keep)
        H, W = H0.clone().requires_grad_(True), W0.clone().requires_grad_(True)
        loss = b200.fused_lm_head_ce(H.view(N // T, T, D), W, labels)
        (loss * 0.25).backward()                               # upstream gradient != 1 (gradient accumulation)
        out[keep] = (loss.detach().clone(), H.grad.clone(), W.grad.clone())
    assert torch.equal(out[True][0], out[False][0])
    assert torch.equal(out[True][1].view(torch.int16), out[False][1].view(torch.int16)), "dH differs between the routes"
    assert torch.equal(out[True][2].view(torch.int16), out[False][2].view(torch.int16)), "dW differs between the routes"
    assert float(out[True][1].float().abs().max()) > 0
    # no gradient wanted -> nothing is kept (evaluation must not pay for the logits buffer)
    monkeypatch.setattr(L, "_KEEP_LOGITS", True)
    with torch.no_grad():
        ev = b200.fused_lm_head_ce(H0.view(N // T, T, D), W0, labels)
    assert torch.equal(ev, out[True][0])
    # the kept logits are consumed by the first backward: a second one over the same graph must say so
    H, W = H0[:1024].clone().requires_grad_(True), W0.clone().requires_grad_(True)
    loss = b200.fused_lm_head_ce(H.view(1, T, D), W, labels[:1])
    loss.backward(retain_graph=True)
    with pytest.raises(b200.B2Error):
        loss.backward()


@pytest.mark.parametrize("D", [2048, 4096])
def test_llama_head_shapes_match_torch_and_routes_agree(b200, monkeypatch, D):
    """BASELINE configs[2] / configs[3] head shapes (ref: ray-jobs/fine_tune_llama_ray.py:230-241,333 loads a 128 256-word
    vocabulary): N = 2 x 2048 tokens, D = 2048 (Llama-3.2-1B) and 4096 (Llama-3-8B; 64 K-blocks, the dH GEMM takes the
    split-K schedule). Loss and token count against torch's cross_entropy on (H @ W^T).float(); dH / dW relative Frobenius
    error < 1e-2 against autograd; ignored rows exactly zero; kept-logits route and recompute route give the same bits."""
    from b200_train import loss as L
    from b200_train import ops
    N, V, T = 4096, 128256, 2048
    g = torch.Generator(device="cpu").manual_seed(40 + D)
    H0 = (torch.randn(N, D, generator=g) * 0.7).to(dev(), torch.bfloat16)
    W0 = (torch.randn(V, D, generator=g) * (0.8 / D ** 0.5)).to(dev(), torch.bfloat16)
    _, labels = TR.synthetic_hf_batch(N // T, T, V, seed=12, device=dev())
    ys = torch.full_like(labels, -100); ys[:, :-1] = labels[:, 1:]
    # reference: the stock arithmetic on the same device
    Hr, Wr = H0.clone().requires_grad_(True), W0.clone().requires_grad_(True)
    want = torch.nn.functional.cross_entropy((Hr @ Wr.t()).float(), ys.reshape(-1), ignore_index=-100)
    (want * 0.25).backward()
    r = ops.lmhead_ce_fwd(H0, W0, labels.reshape(-1), shift_T=T)
    assert int(r.n_valid.item()) == int((ys != -100).sum())           # bit-exact token count
    assert int(r.err.item()) == 0
    out = {}
    for keep in (True, False):
        monkeypatch.setattr(L, "_KEEP_LOGITS", keep)
        H, W = H0.clone().requires_grad_(True), W0.clone().requires_grad_(True)
        loss = b200.fused_lm_head_ce(H.view(N // T, T, D), W, labels)
        (loss * 0.25).backward()
        out[keep] = (loss.detach().clone(), H.grad.clone(), W.grad.clone())
        del H, W
    np.testing.assert_allclose(float(out[True][0]), float(want), rtol=0, atol=2e-3)
    assert torch.equal(out[True][0], out[False][0])
    assert torch.equal(out[True][1].view(torch.int16), out[False][1].view(torch.int16)), "dH differs between the routes"
    assert torch.equal(out[True][2].view(torch.int16), out[False][2].view(torch.int16)), "dW differs between the routes"
    for got, ref, name in ((out[True][1], Hr.grad, "dH"), (out[True][2], Wr.grad, "dW")):
        got, ref = got.float(), ref.float()
        rel = (got - ref).norm() / ref.norm()
        assert torch.isfinite(got).all() and rel < 1e-2, f"{name}: relative Frobenius error {rel:.3e}"
    ign = ys.reshape(-1) == -100
    assert ign.any() and float(out[True][1][ign].abs().max()) == 0.0


def test_full_size_adamw_and_sumsq_properties(b200):
    from b200_train import ops
    n = (1 << 27) + 192
    g = (torch.randn(n, device=dev()) * 0.01).to(torch.bfloat16)
    s1 = ops.sumsq(g).item()
    s2 = ops.sumsq(g * 2).item()                       # exact in bf16: (2g)^2 = 4 g^2
    np.testing.assert_allclose(s2, 4 * s1, rtol=1e-6)
    np.testing.assert_allclose(s1, g.float().double().pow(2).sum().item(), rtol=1e-5)
    p = torch.randn(n, device=dev()).to(torch.bfloat16)
    p0 = p.clone()
    m = torch.zeros_like(p); v = torch.zeros_like(p)
    ops.adamw_(p, g, m, v, lr=0.0, weight_decay=0.0, step=1)     # lr = 0: parameters must come back bit-identical
    assert torch.equal(p, p0)
    assert torch.equal(m, (g.float() * np.float32(1.0 - 0.9)).to(torch.bfloat16))


# ------------------------------------------------------------------------------------------ F2 on one GPU
def test_peer_allreduce_single_gpu_emulation(b200):
    """All `world` ranks' buffers live on one GPU and each rank's kernel runs in turn with the barrier flags already
    satisfied (kernels that wait for one another must not share a GPU). Slices are independent, so the result equals
    the real exchange: every buffer ends with round(mean over ranks), plus per-slice sum of squares."""
    from b200_train import _lib
    for dtype, world in ((torch.bfloat16, 2), (torch.bfloat16, 8), (torch.float32, 4), (torch.bfloat16, 3)):
        n = 64 * 1000 + 64
        rng = torch.Generator(device="cpu").manual_seed(world)
        bufs = [(torch.randn(n, generator=rng) * 0.1).to(dev(), dtype) for _ in range(world)]
        orig = [b.clone() for b in bufs]
        flags = [torch.zeros(_lib.lib.b2_peer_flags_bytes() // 4, dtype=torch.int32, device=dev()) for _ in range(world)]
        epoch = 5
        for f in flags:
            f[:16] = epoch
        bt = (ctypes.c_void_p * world)(*[b.data_ptr() for b in bufs])
        ft = (ctypes.c_void_p * world)(*[f.data_ptr() for f in flags])
        for r in range(world):
            _lib.call("b2_peer_allreduce", bt, ft, world, r, 0, n, _lib.dtype_code(dtype), 1.0 / world, epoch, 1, 0, 0,
                      None, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        acc = torch.zeros(n, dtype=torch.float32, device=dev())
        for o in orig:
            acc += o.float()
        want = (acc * np.float32(1.0 / world)).to(dtype)
        for b in bufs:
            if world == 2 or dtype == torch.float32:
                assert torch.equal(b, want)               # fixed rank order: bit-exact against the same fp32 sum
            else:
                assert (b.float() - want.float()).abs().max() <= float(want.float().abs().max()) * 2 ** -7
        assert all(torch.equal(b, bufs[0]) for b in bufs)  # every rank holds identical bits
        total = sum(float(flags[r].view(torch.float32)[32 + r].item()) for r in range(world))
        np.testing.assert_allclose(total, bufs[0].float().double().pow(2).sum().item(), rtol=1e-5)
        # reduce-scatter + all-gather halves give the same result
        bufs2 = [o.clone() for o in orig]
        bt2 = (ctypes.c_void_p * world)(*[b.data_ptr() for b in bufs2])
        for r in range(world):
            _lib.call("b2_peer_allreduce", bt2, ft, world, r, 0, n, _lib.dtype_code(dtype), 1.0 / world, epoch, 0, 1, 0,
                      None, torch.cuda.current_stream().cuda_stream)
        for r in range(world):
            _lib.call("b2_peer_allgather", bt2, ft, world, r, 0, n, _lib.dtype_code(dtype), epoch, 0,
                      torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        for b in bufs2:
            assert torch.equal(b, bufs[0])


# ------------------------------------------------------------------------------------------ checkpoint parity (§8f-1)
def test_optimizer_state_dict_roundtrip_matches_torch(b200):
    """optimizer.state_dict() (ref: pytorch_llm_ray.py:302) is interchangeable with torch.optim.AdamW's: a checkpoint
    written by torch resumes in ours (and vice versa) and the next step lands on the same parameters."""
    torch.manual_seed(0)
    m_ref = nn.Sequential(nn.Linear(64, 128), nn.GELU(), nn.Linear(128, 32)).to(dev())
    m_my = copy.deepcopy(m_ref)
    o_ref = torch.optim.AdamW(m_ref.parameters(), lr=1e-3, weight_decay=0.01)
    o_my = b200.AdamW(m_my.parameters(), lr=1e-3, weight_decay=0.01)
    x = torch.randn(16, 64, device=dev())
    for _ in range(2):
        for m, o in ((m_ref, o_ref), (m_my, o_my)):
            o.zero_grad(set_to_none=True)
            m(x).pow(2).mean().backward()
            o.step()
    # torch -> ours
    m_a = copy.deepcopy(m_ref); o_a = b200.AdamW(m_a.parameters(), lr=1e-3, weight_decay=0.01)
    o_a.load_state_dict(copy.deepcopy(o_ref.state_dict()))
    # ours -> torch
    m_b = copy.deepcopy(m_my); o_b = torch.optim.AdamW(m_b.parameters(), lr=1e-3, weight_decay=0.01)
    o_b.load_state_dict(copy.deepcopy(o_my.state_dict()))     # state_dict() hands out live tensors, as torch's does
    for m, o in ((m_ref, o_ref), (m_a, o_a), (m_my, o_my), (m_b, o_b)):
        o.zero_grad(set_to_none=True)
        m(x).pow(2).mean().backward()
        o.step()
    for pr, pa in zip(m_ref.parameters(), m_a.parameters()):
        np.testing.assert_allclose(pa.detach().cpu().numpy(), pr.detach().cpu().numpy(), rtol=0, atol=1e-7)
    for pm, pb in zip(m_my.parameters(), m_b.parameters()):
        np.testing.assert_allclose(pb.detach().cpu().numpy(), pm.detach().cpu().numpy(), rtol=0, atol=1e-7)
    assert float(o_a.state_dict()["state"][0]["step"]) == 3.0


# ------------------------------------------------------------------------------------------ input side (§8f-3)
def test_resident_window_loader_matches_reference_pipeline(b200):
    """Batch for batch identical (integers: bit-exact) to TextDataset -> Subset -> DataLoader(shuffle, drop_last) ->
    prepare_data_loader's DistributedSampler, for both ranks of a 2-worker job."""
    from torch.utils.data import DataLoader, DistributedSampler, TensorDataset
    stream = TR.synthetic_char_stream(5000, 283)
    T, B, limit = 64, 16, 1000
    for rank in (0, 1):
        n = min(max(0, len(stream) - T - 1), limit)
        xs, ys = TR.windows(stream, torch.arange(n), T)                 # == TextDataset[i] (pinned by make_golden.py)
        ds = TensorDataset(xs, ys)
        sampler = DistributedSampler(ds, num_replicas=2, rank=rank, shuffle=True, seed=0)
        ref = DataLoader(ds, batch_size=B, sampler=sampler, drop_last=True)
        mine = b200.ResidentWindowLoader(stream, T, B, dev(), shuffle=True, num_replicas=2, rank=rank, limit=limit)
        for epoch in (0, 1):
            sampler.set_epoch(epoch); mine.sampler.set_epoch(epoch)
            assert len(mine) == len(ref)
            for (rx, ry), (mx, my) in zip(ref, mine):
                assert torch.equal(mx.cpu(), rx) and torch.equal(my.cpu(), ry)
    # a window that does not fit is flagged, not read out of bounds
    bad = mine.gather(torch.tensor([len(stream) - T], device=dev()))
    assert int(mine._err.item()) == 1 and int(bad[0].abs().sum()) == 0


def test_adamw_skips_parameters_without_gradient_like_torch(b200):
    """torch.optim.AdamW leaves a parameter whose .grad is None alone (no weight decay either); so do we."""
    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.used = nn.Linear(32, 32)
            self.unused = nn.Linear(32, 32)

        def forward(self, x):
            return self.used(x)

    torch.manual_seed(0)
    m_ref = Net().to(dev()); m_my = copy.deepcopy(m_ref)
    o_ref = torch.optim.AdamW(m_ref.parameters(), lr=1e-2, weight_decay=0.1)
    o_my = b200.AdamW(m_my.parameters(), lr=1e-2, weight_decay=0.1)
    x = torch.randn(8, 32, device=dev())
    for _ in range(3):
        for m, o in ((m_ref, o_ref), (m_my, o_my)):
            o.zero_grad(set_to_none=True)
            m(x).pow(2).mean().backward()
            o.step()
    assert torch.equal(m_my.unused.weight, m_ref.unused.weight)          # untouched: bit-identical to the initial value
    np.testing.assert_allclose(m_my.used.weight.detach().cpu().numpy(), m_ref.used.weight.detach().cpu().numpy(), atol=1e-6)


def test_foreign_zero_grad_keeps_flat_gradients(b200):
    """HF Trainer calls model.zero_grad() (gradients -> None) every step; each gradient then lands in its slice of the flat
    buffer as backward produces it (flat.FlatBuffer._land), so the step still matches the stock path."""
    ref_model = _tiny_gpt2().train()
    my_model = b200.patch_causal_lm(copy.deepcopy(ref_model)).train()
    before = {n: p.detach().clone() for n, p in ref_model.named_parameters()}
    ref_opt = torch.optim.AdamW(TR.hf_param_groups(ref_model, 1e-3), lr=2e-4)
    my_opt = b200.AdamW(TR.hf_param_groups(my_model, 1e-3), lr=2e-4)
    ids, lab = TR.synthetic_hf_batch(4, 64, 2003, seed=77, device=dev())
    for _ in range(2):
        ref_model.zero_grad(); my_model.zero_grad()                      # set_to_none=True is torch's default
        assert all(p.grad is None for p in my_model.parameters())
        ref_model(input_ids=ids, labels=lab).loss.backward()
        my_model(input_ids=ids, labels=lab).loss.backward()
        buf = my_opt.flat_buffers()[0]
        assert all(p.grad is not None for p in buf.params)
        torch.nn.utils.clip_grad_norm_(ref_model.parameters(), 0.3); ref_opt.step()
        b200.clip_grad_norm_(my_model.parameters(), 0.3)
        # by the time anything reads the buffer as a whole, every gradient has landed in its slice
        assert all(p.grad.data_ptr() == buf.grad_view(i).data_ptr() for i, p in enumerate(buf.params))
        my_opt.step()
    for (n, a), (_, b) in zip(my_model.named_parameters(), ref_model.named_parameters()):
        _bf16_close(a, b, before[n], n, steps=2)


def test_head_gradient_is_written_into_the_flat_buffer(b200):
    """The LM head's dW goes straight into the parameter's slice of the flat gradient buffer: accumulated there for a tied
    head (GPT-2: the embedding's share is added in place afterwards), adopted without a copy for an untied one (Llama-3-8B
    shape). Values against the stock path; two micro-batches exercise the accumulate branch."""
    from b200_train.flat import flat_of
    tied = _tiny_gpt2().train()
    untied = TR.make_llama(dtype=torch.bfloat16, seed=0, vocab_size=4099, hidden_size=128, intermediate_size=256,
                           num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2, head_dim=32,
                           tie_word_embeddings=False, max_position_embeddings=128).to(dev()).train()
    for ref_model, vocab in ((tied, 2003), (untied, 4099)):
        my_model = b200.patch_causal_lm(copy.deepcopy(ref_model)).train()
        my_opt = b200.AdamW(TR.hf_param_groups(my_model, 1e-3), lr=2e-4)
        head = my_model.get_output_embeddings().weight
        buf, _ = flat_of(head)
        view = buf.grad_view(buf._index[id(head)])
        batches = [TR.synthetic_hf_batch(2, 64, vocab, seed=300 + i, device=dev()) for i in range(2)]
        n_items = sum(int((lab[:, 1:] != -100).sum()) for _, lab in batches)
        my_opt.zero_grad(set_to_none=True)
        buf.grad.fill_(3.0)                                      # stale contents must not survive
        for k, (ids, lab) in enumerate(batches):
            ref_model(input_ids=ids, labels=lab, num_items_in_batch=n_items).loss.backward()
            my_model(input_ids=ids, labels=lab, num_items_in_batch=n_items).loss.backward()
            assert head.grad is view, f"micro-batch {k}: .grad is not the flat slice"
        got, want = head.grad.float(), ref_model.get_output_embeddings().weight.grad.float()
        rel = float((got - want).norm() / want.norm())
        assert rel < 1e-2, rel
        for (n, a), (_, b) in zip(my_model.named_parameters(), ref_model.named_parameters()):
            assert float((a.grad.float() - b.grad.float()).norm()) <= 2e-2 * float(b.grad.float().norm()) + 1e-6, n


def test_clip_semantics_with_attached_optimizer(b200):
    """torch's clip_grad_norm_ scales .grad in place (ref: pytorch_llm_ray.py:277). With b200_train.AdamW attached the scale
    is deferred into the optimizer kernel: (a) defer_to_optimizer=False restores the in-place behaviour; (b) two deferred
    clips before one step compose like two in-place clips; either way the parameters after step() match torch's."""
    torch.manual_seed(0)
    m_ref = nn.Sequential(nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, 8)).to(dev())
    x = torch.randn(16, 64, device=dev())
    for mode in ("inplace", "twice"):
        ref = copy.deepcopy(m_ref); mine = copy.deepcopy(m_ref)
        o_ref = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.1)
        o_my = b200.AdamW(mine.parameters(), lr=1e-2, weight_decay=0.1)
        (ref(x).pow(2).sum() * 50).backward(); (mine(x).pow(2).sum() * 50).backward()
        if mode == "inplace":
            n_ref = torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
            n_my = b200.clip_grad_norm_(mine.parameters(), 1.0, defer_to_optimizer=False)
            for a, b in zip(mine.parameters(), ref.parameters()):
                np.testing.assert_allclose(a.grad.cpu().numpy(), b.grad.cpu().numpy(), rtol=2e-5, atol=1e-7)
        else:
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 2.0)
            n_ref = torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.5)
            b200.clip_grad_norm_(mine.parameters(), 2.0)
            n_my = b200.clip_grad_norm_(mine.parameters(), 0.5)          # sees the norm AFTER the first clip, like torch
        np.testing.assert_allclose(n_my.item(), n_ref.item(), rtol=1e-4)
        o_ref.step(); o_my.step()
        for a, b in zip(mine.parameters(), ref.parameters()):
            np.testing.assert_allclose(a.detach().cpu().numpy(), b.detach().cpu().numpy(), rtol=0, atol=2e-6)


def test_out_of_range_label_raises_when_asked(b200):
    H = torch.zeros(4, 64, dtype=torch.bfloat16, device=dev()); W = torch.zeros(16, 64, dtype=torch.bfloat16, device=dev())
    y = torch.tensor([[0, 1, 16, 2]], device=dev())
    with pytest.raises(IndexError):
        b200.fused_lm_head_ce(H.view(1, 4, 64), W, y, shift=False, check_labels=True)
    with pytest.raises(IndexError):
        b200.CrossEntropyLoss(check_labels=True)(torch.zeros(4, 16, device=dev()), y.reshape(-1))
    assert torch.isnan(b200.fused_lm_head_ce(H.view(1, 4, 64), W, y, shift=False))        # default: poisoned, no sync
