"""Pins the CPU oracle (oracle/step_tail.py) against tests/golden/*.npz — outputs of the reference's own classes
and call sequence on stock torch, produced by oracle/make_golden.py. Runs without a GPU."""
import os

import numpy as np
import pytest

from oracle import step_tail as O

G = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    return np.load(os.path.join(G, name))


def test_ce_kat_mean_sum_grad():
    g = load("ce_kat.npz")
    loss, n_valid, _ = O.cross_entropy(g["x"], g["y"])
    assert n_valid == 4
    np.testing.assert_allclose(loss, g["loss_mean"], rtol=1e-6)
    loss_s, _, _ = O.cross_entropy(g["x"], g["y"], reduction="sum")
    np.testing.assert_allclose(loss_s, g["loss_sum"], rtol=1e-6)
    d = O.cross_entropy_grad(g["x"], g["y"], 1.0 / n_valid)
    np.testing.assert_allclose(d, g["grad_mean"], rtol=1e-5, atol=1e-7)


def test_ce_kat_hf_shift_and_num_items():
    g = load("ce_kat.npz")
    x, y = g["hf_logits"], g["hf_labels"]
    ys = O.shift_labels(y)
    assert (ys != -100).sum() == 6                       # HF shift on 2x5 -> n_valid = 6
    loss, n_valid, _ = O.cross_entropy(x.reshape(-1, 13), ys.reshape(-1))
    assert n_valid == 6
    np.testing.assert_allclose(loss, g["hf_mean"], rtol=1e-6)
    loss9, _, _ = O.cross_entropy(x.reshape(-1, 13), ys.reshape(-1), num_items=9)
    np.testing.assert_allclose(loss9, g["hf_items9"], rtol=1e-6)


def test_ce_edge_cases_match_torch():
    g = load("ce_kat.npz")
    all_ign = np.full(6, -100)
    loss, n_valid, _ = O.cross_entropy(g["x"], all_ign)
    assert n_valid == 0 and np.isnan(loss) and np.isnan(g["all_ignored_mean"])
    loss, _, _ = O.cross_entropy(g["x"], all_ign, reduction="sum")
    assert loss == 0.0 and g["all_ignored_sum"] == 0.0
    with pytest.raises(IndexError):
        O.cross_entropy(g["x"], np.array([0, 1, 2, 3, 4, 11]))
    with pytest.raises(IndexError):
        O.cross_entropy(g["x"], np.array([0, 1, 2, 3, 4, -1]))


def test_basicllm_head_and_loss_from_reference_model():
    g = load("basicllm_tiny.npz")
    H = g["features"].reshape(-1, g["features"].shape[-1])
    r = O.lmhead_ce(H, g["fc_w"], g["y"].reshape(-1), bias=g["fc_b"], shift_T=0, bf16_logits=False)
    np.testing.assert_allclose(r["logits"], g["logits"].reshape(-1, g["logits"].shape[-1]), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(r["loss"], g["loss0"], rtol=1e-6)
    assert r["n_valid"] == g["y"].size                  # BasicLLM path: nothing is ever ignored


def test_lr_lambda_matches_reference_schedule():
    g = load("basicllm_tiny.npz")
    mine = [O.lr_lambda(s, 20) for s in range(25)]
    np.testing.assert_allclose(mine, g["lr_lambda"], rtol=0, atol=0)
    np.testing.assert_allclose(g["lrs"], [3e-4 * O.lr_lambda(s, 20) for s in range(3)], rtol=1e-12)


def test_adamw_trajectory_fp32():
    g = load("adamw_traj.npz")
    p = g["f32_p0"]; m = np.zeros_like(p); v = np.zeros_like(p)
    for s in range(10):
        p, m, v = O.adamw_step(p, g["f32_grads"][s], m, v, lr=3e-4, weight_decay=0.01, step=s + 1)
    for got, want in ((p, g["f32_p"]), (m, g["f32_m"]), (v, g["f32_v"])):
        assert (np.abs(got.astype(np.float64) - want) <= 4 * O.f32_ulp(want)).all()


def test_adamw_trajectory_bf16_bit_exact():
    g = load("adamw_traj.npz")
    p = g["bf16_p0"]; m = np.zeros_like(p); v = np.zeros_like(p)
    for s in range(10):
        p, m, v = O.adamw_step(p, g["bf16_grads"][s], m, v, lr=2e-4, weight_decay=1e-3, step=s + 1, state_bf16=True)
    np.testing.assert_array_equal(p, g["bf16_p"])
    np.testing.assert_array_equal(m, g["bf16_m"])
    np.testing.assert_array_equal(v, g["bf16_v"])


def test_clip_grad_norm():
    g = load("clip_kat.npz")
    grads = [g["g0"], g["g1"], g["g2"]]
    total, coef, clipped = O.clip_grad_norm(grads, 0.3)
    np.testing.assert_allclose(total, g["clip_total"], rtol=1e-6)
    assert coef < 1
    for i, c in enumerate(clipped):
        np.testing.assert_allclose(c, g[f"clip_g{i}"], rtol=2e-6, atol=1e-9)
    total, coef, clipped = O.clip_grad_norm(grads, 100.0)
    assert coef == 1.0
    for i, c in enumerate(clipped):
        np.testing.assert_array_equal(c, g[f"noclip_g{i}"])


def test_ddp_average_matches_gloo_ddp():
    g = load("ddp_gloo2.npz")
    avg = O.ddp_average([g["local0"], g["local1"]])
    np.testing.assert_allclose(avg, g["synced"], rtol=0, atol=1e-9)   # W=2: division exact, one addition


def test_hf_tiny_gpt2_loss_and_grads():
    g = load("hf_tiny_gpt2.npz")
    H = g["hidden"].reshape(-1, g["hidden"].shape[-1]); W = g["W"]
    T = g["labels"].shape[1]
    r = O.lmhead_ce(H, W, g["labels"], shift_T=T, bf16_logits=True)
    assert r["n_valid"] == int(g["n_valid"])            # masking / token count: bit-exact
    np.testing.assert_allclose(r["loss"], g["loss"], rtol=0, atol=2e-3)
    dH, dW, _, _ = O.lmhead_ce_grads(H, W, g["labels"], 1.0 / r["n_valid"], shift_T=T, bf16_logits=True)
    # torch's bf16 autograd rounds dH / dW to bf16
    assert (np.abs(dH - g["dH"].reshape(dH.shape)) <= 2 * O.bf16_ulp(g["dH"].reshape(dH.shape)) + 1e-6).all()
    assert (np.abs(dW - g["dW"]) <= 2 * O.bf16_ulp(g["dW"]) + 1e-6).all()


def test_bf16_round_matches_torch():
    import torch
    x = np.random.default_rng(0).standard_normal(10000).astype(np.float32) * 100
    want = torch.from_numpy(x).to(torch.bfloat16).float().numpy()
    np.testing.assert_array_equal(O.bf16_round(x), want)
