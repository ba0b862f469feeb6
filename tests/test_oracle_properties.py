"""The CPU oracle against stock torch on randomly drawn shapes and hyper-parameters (hypothesis; no GPU).

tests/test_oracle_golden.py pins the oracle on fixed vectors produced by the reference's own classes; this file widens
that to drawn cases, with stock torch (CPU) as the authority — the arithmetic of the reference's step tail lives in
torch / transformers, not in /root/reference. Example counts are small: the whole CPU suite has to stay
within a few minutes.
"""
import math

import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from oracle import step_tail as O

SET = dict(max_examples=25, deadline=None, derandomize=True, database=None)


@settings(**SET)
@given(n=st.integers(1, 40), v=st.integers(2, 300), frac=st.floats(0.0, 0.9), seed=st.integers(0, 2 ** 16),
       reduction=st.sampled_from(["mean", "sum"]))
def test_cross_entropy_and_gradient_match_torch(n, v, frac, seed, reduction):
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal((n, v)) * 3).astype(np.float32)
    y = rng.integers(0, v, size=n)
    y[rng.random(n) < frac] = -100
    xt = torch.tensor(x, requires_grad=True)
    want = torch.nn.functional.cross_entropy(xt, torch.tensor(y), ignore_index=-100, reduction=reduction)
    loss, n_valid, lse = O.cross_entropy(x, y, reduction=reduction)
    assert n_valid == int((y != -100).sum())
    if n_valid == 0 and reduction == "mean":
        assert math.isnan(loss) and math.isnan(float(want.detach()))            # torch: 0 / 0
        return
    np.testing.assert_allclose(loss, float(want.detach()), rtol=2e-6, atol=1e-6)
    np.testing.assert_allclose(lse, torch.logsumexp(xt.detach(), dim=1).numpy(), rtol=2e-6, atol=1e-6)
    want.backward()
    scale = 1.0 if reduction == "sum" else 1.0 / n_valid
    np.testing.assert_allclose(O.cross_entropy_grad(x, y, scale), xt.grad.numpy(), rtol=2e-5, atol=2e-7)


@settings(**SET)
@given(b=st.integers(1, 4), t=st.integers(2, 12), seed=st.integers(0, 2 ** 16))
def test_shift_labels_is_the_hf_causal_shift(b, t, seed):
    """$SP/transformers/loss/loss_utils.py:58-61: labels = pad(labels, (0, 1), value=-100)[..., 1:]."""
    rng = np.random.default_rng(seed)
    y = rng.integers(-1, 50, size=(b, t))
    y[y == -1] = -100
    want = torch.nn.functional.pad(torch.tensor(y), (0, 1), value=-100)[..., 1:].numpy()
    np.testing.assert_array_equal(O.shift_labels(y), want)


@settings(**SET)
@given(seed=st.integers(0, 2 ** 16), steps=st.integers(1, 6), lr=st.floats(1e-5, 1e-2), wd=st.floats(0.0, 0.2),
       b1=st.floats(0.5, 0.95), b2=st.floats(0.9, 0.9995), eps=st.sampled_from([1e-8, 1e-6]))
def test_adamw_fp32_trajectory_matches_torch_single_tensor(seed, steps, lr, wd, b1, b2, eps):
    rng = np.random.default_rng(seed)
    p0 = rng.standard_normal(257).astype(np.float32)
    pt = torch.nn.Parameter(torch.tensor(p0))
    opt = torch.optim.AdamW([pt], lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd, foreach=False)
    p, m, v = p0.copy(), np.zeros_like(p0), np.zeros_like(p0)
    for s in range(1, steps + 1):
        g = (rng.standard_normal(257) * 0.1).astype(np.float32)
        pt.grad = torch.tensor(g)
        opt.step()
        p, m, v = O.adamw_step(p, g, m, v, lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=wd, step=s)
    got = pt.detach().numpy()
    # torch's CPU kernels fuse some of these ops differently from the op-by-op restatement: a few fp32 ulp per step
    tol = 4 * steps * O.f32_ulp(np.maximum(np.abs(got), np.abs(p0))) + 1e-9
    assert (np.abs(p - got) <= tol).all(), float(np.abs(p - got).max())
    st_ = opt.state[pt]
    # (absolute floors: an exp_avg element can cancel to ~0 while its addends are ~0.1; one fp32 ulp of the addends)
    np.testing.assert_allclose(m, st_["exp_avg"].numpy(), rtol=1e-5, atol=2e-8)
    np.testing.assert_allclose(v, st_["exp_avg_sq"].numpy(), rtol=1e-5, atol=1e-10)


@settings(**SET)
@given(seed=st.integers(0, 2 ** 16), k=st.integers(1, 6), max_norm=st.floats(0.05, 5.0))
def test_clip_grad_norm_matches_torch(seed, k, max_norm):
    rng = np.random.default_rng(seed)
    grads = [(rng.standard_normal(rng.integers(1, 50)) * rng.uniform(0.01, 2.0)).astype(np.float32) for _ in range(k)]
    params = [torch.nn.Parameter(torch.zeros(len(g))) for g in grads]
    for p, g in zip(params, grads):
        p.grad = torch.tensor(g)
    total = torch.nn.utils.clip_grad_norm_(params, max_norm)
    t, coef, clipped = O.clip_grad_norm(grads, max_norm)
    np.testing.assert_allclose(t, float(total), rtol=2e-6)
    assert coef <= 1.0
    for p, c in zip(params, clipped):
        np.testing.assert_allclose(c, p.grad.numpy(), rtol=3e-6, atol=1e-9)


@settings(**SET)
@given(seed=st.integers(0, 2 ** 16), n=st.integers(1, 2000), scale=st.sampled_from([1e-30, 1e-3, 1.0, 1e4, 1e30]))
def test_bf16_round_is_round_to_nearest_even_like_torch(seed, n, scale):
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n) * scale).astype(np.float32)
    want = torch.tensor(x).to(torch.bfloat16).float().numpy()
    np.testing.assert_array_equal(O.bf16_round(x), want)
    # ulp helper: the spacing of bf16 numbers at |x| (8 significant bits)
    nz = want != 0
    up = torch.nextafter(torch.tensor(want).to(torch.bfloat16), torch.tensor(float("inf"), dtype=torch.bfloat16)).float().numpy()
    ok = nz & np.isfinite(up) & (np.abs(want) > 1e-30)
    pos = ok & (want > 0)
    np.testing.assert_allclose(O.bf16_ulp(want[pos]), (up - want)[pos], rtol=0, atol=0)


@settings(**SET)
@given(world=st.sampled_from([2, 4, 8]), seed=st.integers(0, 2 ** 16), n=st.integers(1, 300))
def test_ddp_average_is_exact_mean_for_power_of_two_worlds(world, seed, n):
    """dividing by W before summing is exact for W in {2, 4, 8} (no subnormals involved), so the
    oracle's pre-divided sum equals sum / W computed in float64 and rounded once — when the fp32 sum itself is exact,
    i.e. for inputs with few significant bits."""
    rng = np.random.default_rng(seed)
    grads = [np.round(rng.standard_normal(n) * 64).astype(np.float32) / 64 for _ in range(world)]
    want = (np.sum(np.stack(grads).astype(np.float64), axis=0) / world).astype(np.float32)
    np.testing.assert_array_equal(O.ddp_average(grads), want)


@settings(**SET)
@given(total=st.integers(10, 5000), warm=st.floats(0.0, 0.3), floor=st.floats(0.0, 0.5))
def test_lr_lambda_shape(total, warm, floor):
    """ref: ray-jobs/pytorch_llm_ray.py:240-258 — starts at 0 (when there is a warm-up), peaks at 1 at the end of the
    warm-up, never drops below the floor afterwards, is non-increasing after the peak."""
    vals = [O.lr_lambda(s, total, warm, floor) for s in range(total + 5)]
    w = int(total * warm)
    if w > 0:
        assert vals[0] == 0.0 and abs(vals[w] - 1.0) < 1e-12
        assert all(b >= a for a, b in zip(vals[:w], vals[1:w + 1]))
    assert all(v >= floor - 1e-12 for v in vals[w:])
    assert all(b <= a + 1e-12 for a, b in zip(vals[w:], vals[w + 1:]))
    assert abs(vals[-1] - floor) < 1e-12
