"""Parity of each CUDA kernel, called through the C ABI, against the CPU oracle on seeded inputs.

Tolerances (stated per test): integer outputs (n_valid, error flag, label masking) bit-exact; fp32 results to a
few ulp of the oracle's fp32 arithmetic; bf16 results to 1 bf16 ulp.
"""
import numpy as np
import pytest
import torch

from oracle import step_tail as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from b200_train import ops
    return ops


def dev():
    return torch.device("cuda", 0)


def bf16_tensor(rng, shape, scale=1.0):
    x = O.bf16_round((rng.standard_normal(shape) * scale).astype(np.float32))
    return x, torch.from_numpy(x).to(dev()).to(torch.bfloat16)


# ------------------------------------------------------------------------------------------ GEMM core
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (300, 520, 200), (1000, 768, 1032)])
def test_gemm_bf16_all_majors(ops, a_mn, b_mn, M, N, K):
    rng = np.random.default_rng(M * 7 + N * 3 + K + a_mn * 2 + b_mn)
    # storage must have 16-byte-aligned rows: pad leading dims to multiples of 8
    pad8 = lambda n: (n + 7) // 8 * 8
    A, _ = bf16_tensor(rng, (M, K))
    B, _ = bf16_tensor(rng, (N, K))
    if a_mn:
        At = torch.zeros(K, pad8(M), dtype=torch.bfloat16, device=dev()); At[:, :M] = torch.from_numpy(A.T.copy()).to(dev())
    else:
        At = torch.zeros(M, pad8(K), dtype=torch.bfloat16, device=dev()); At[:, :K] = torch.from_numpy(A).to(dev())
    if b_mn:
        Bt = torch.zeros(K, pad8(N), dtype=torch.bfloat16, device=dev()); Bt[:, :N] = torch.from_numpy(B.T.copy()).to(dev())
    else:
        Bt = torch.zeros(N, pad8(K), dtype=torch.bfloat16, device=dev()); Bt[:, :K] = torch.from_numpy(B).to(dev())
    C = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev())
    ops.gemm_bf16(At, Bt, C, M, N, K, a_mn_major=a_mn, b_mn_major=b_mn)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    got = C.cpu().numpy().astype(np.float64)
    # fp32 accumulation of K products of O(1) values: error ~ K * 2^-24 * |terms|
    tol = 4e-6 * K + 1e-5
    assert np.isfinite(got).all()
    assert np.abs(got - ref).max() <= tol * max(1.0, np.abs(ref).max())


def test_gemm_bf16_accumulate_and_bf16_out(ops):
    rng = np.random.default_rng(5)
    M, N, K = 260, 300, 136
    A, At = bf16_tensor(rng, (M, K))
    B, Bt = bf16_tensor(rng, (N, K))
    C0 = rng.standard_normal((M, N)).astype(np.float32)
    C = torch.from_numpy(C0.copy()).to(dev())
    ops.gemm_bf16(At, Bt, C, M, N, K, accumulate=True)
    ref = C0 + A.astype(np.float64) @ B.astype(np.float64).T
    assert np.abs(C.cpu().numpy() - ref).max() <= 1e-3
    Cb = torch.zeros(M, 304, dtype=torch.bfloat16, device=dev())
    ops.gemm_bf16(At, Bt, Cb, M, N, K)
    refb = A.astype(np.float64) @ B.astype(np.float64).T
    got = Cb[:, :N].float().cpu().numpy()
    assert (np.abs(got - refb) <= O.bf16_ulp(refb) + 1e-3).all()


# ------------------------------------------------------------------------------------------ F1 forward
def make_labels(rng, N, V, frac_ignored=0.2):
    y = rng.integers(0, V, size=N, dtype=np.int64)
    y[rng.random(N) < frac_ignored] = -100
    return y


@pytest.mark.parametrize("N,D,V", [(128, 64, 256), (384, 192, 1000), (200, 256, 283), (1024, 768, 5003),
                                   (8, 72, 8), (136, 1032, 257), (264, 64, 40)])     # ragged in every dimension
@pytest.mark.parametrize("shift_T", [0, 8])
@pytest.mark.parametrize("with_bias", [False, True])
def test_lmhead_ce_fwd_matches_oracle(ops, N, D, V, shift_T, with_bias):
    rng = np.random.default_rng(N + D + V + shift_T)
    H, Ht = bf16_tensor(rng, (N, D))
    W, Wt = bf16_tensor(rng, (V, D), scale=2.0 / np.sqrt(D))
    bias = (rng.standard_normal(V) * 0.5).astype(np.float32) if with_bias else None
    bt = torch.from_numpy(bias).to(dev()) if with_bias else None
    y = make_labels(rng, N, V)
    yt = torch.from_numpy(y).to(dev())
    # exact-logit variant: tight tolerance
    r = ops.lmhead_ce_fwd(Ht, Wt, yt, bias=bt, shift_T=shift_T, round_logits_bf16=False)
    o = O.lmhead_ce(H, W, y, bias, shift_T=shift_T, bf16_logits=False)
    assert int(r.n_valid.item()) == o["n_valid"]                     # bit-exact
    assert int(r.err.item()) == 0
    np.testing.assert_allclose(r.lse[:N].cpu().numpy(), o["lse"], rtol=0, atol=5e-5)
    if o["n_valid"]:
        np.testing.assert_allclose(r.loss_mean.item(), o["loss"], rtol=2e-5, atol=2e-5)
    # reference-faithful variant (logits rounded to bf16 as a bf16 nn.Linear would): a tensor-core sum that lands
    # on the other side of a bf16 rounding boundary moves one logit by one bf16 ulp, hence the looser bound
    r2 = ops.lmhead_ce_fwd(Ht, Wt, yt, bias=bt, shift_T=shift_T, round_logits_bf16=True)
    o2 = O.lmhead_ce(H, W, y, bias, shift_T=shift_T, bf16_logits=True)
    assert int(r2.n_valid.item()) == o2["n_valid"]
    np.testing.assert_allclose(r2.lse[:N].cpu().numpy(), o2["lse"], rtol=0, atol=4e-2)
    if o2["n_valid"]:
        np.testing.assert_allclose(r2.loss_mean.item(), o2["loss"], rtol=0, atol=2e-3)


def test_lmhead_ce_rejects_unaligned_shapes(ops):
    """TMA needs 16-byte-aligned rows: a hidden size that is not a multiple of 8 is an error, never a silent fallback."""
    from b200_train import B2Error
    H = torch.zeros(16, 100, dtype=torch.bfloat16, device=dev())[:, :99]
    W = torch.zeros(32, 100, dtype=torch.bfloat16, device=dev())[:, :99]
    y = torch.zeros(16, dtype=torch.int64, device=dev())
    with pytest.raises(B2Error):
        ops.lmhead_ce_fwd(H.contiguous(), W.contiguous(), y)
    with pytest.raises(B2Error):
        ops.lmhead_ce_fwd(torch.zeros(16, 64, device=dev()), torch.zeros(32, 64, device=dev()), y)     # fp32 inputs


def test_lmhead_ce_edge_cases(ops):
    rng = np.random.default_rng(11)
    N, D, V = 130, 64, 300
    H, Ht = bf16_tensor(rng, (N, D))
    W, Wt = bf16_tensor(rng, (V, D))
    # all ignored: mean is NaN, sum is 0, n_valid 0 (torch: mean -> nan, sum -> 0)
    y = torch.full((N,), -100, dtype=torch.int64, device=dev())
    r = ops.lmhead_ce_fwd(Ht, Wt, y)
    assert int(r.n_valid.item()) == 0 and float(r.loss_sum.item()) == 0.0 and np.isnan(r.loss_mean.item())
    # out-of-range label: flagged, loss poisoned (torch raises IndexError / device assert)
    y2 = torch.zeros(N, dtype=torch.int64, device=dev()); y2[5] = V
    r = ops.lmhead_ce_fwd(Ht, Wt, y2)
    assert int(r.err.item()) == 1 and np.isnan(r.loss_mean.item())
    y2[5] = -1
    r = ops.lmhead_ce_fwd(Ht, Wt, y2)
    assert int(r.err.item()) == 1
    # labels at the first and last vocabulary entry
    y3 = torch.zeros(N, dtype=torch.int64, device=dev()); y3[::2] = V - 1
    r = ops.lmhead_ce_fwd(Ht, Wt, y3, round_logits_bf16=False)
    o = O.lmhead_ce(H, W, y3.cpu().numpy(), bf16_logits=False)
    np.testing.assert_allclose(r.loss_mean.item(), o["loss"], rtol=2e-5)


# ------------------------------------------------------------------------------------------ F1 backward
@pytest.mark.parametrize("N,D,V", [(256, 128, 520), (300, 64, 283)])
@pytest.mark.parametrize("shift_T", [0, 10])
def test_lmhead_ce_bwd_dlogits_matches_oracle(ops, N, D, V, shift_T):
    if shift_T:
        N = N // shift_T * shift_T
    rng = np.random.default_rng(N + V)
    H, Ht = bf16_tensor(rng, (N, D))
    W, Wt = bf16_tensor(rng, (V, D), scale=2.0 / np.sqrt(D))
    y = make_labels(rng, N, V)
    yt = torch.from_numpy(y).to(dev())
    r = ops.lmhead_ce_fwd(Ht, Wt, yt, shift_T=shift_T, round_logits_bf16=False)
    gscale = torch.tensor([0.37], dtype=torch.float32, device=dev())
    ldd = (V + 7) // 8 * 8
    # two chunks to exercise row0
    out = torch.full((N, ldd), float("nan"), dtype=torch.bfloat16, device=dev())
    half = (N // 2 + 127) // 128 * 128 if N > 128 else N
    half = min(half, N)
    ops.lmhead_ce_bwd_dlogits(Ht, Wt, yt, r.lse, gscale, 0, half, out[:half], shift_T=shift_T, round_logits_bf16=False)
    if half < N:
        ops.lmhead_ce_bwd_dlogits(Ht, Wt, yt, r.lse, gscale, half, N - half, out[half:], shift_T=shift_T,
                                  round_logits_bf16=False)
    o = O.lmhead_ce(H, W, y, shift_T=shift_T, bf16_logits=False)
    d = O.cross_entropy_grad(o["logits"], o["labels"], 0.37)
    got = out[:, :V].float().cpu().numpy()
    assert np.isfinite(got).all()
    assert (np.abs(got - d) <= O.bf16_ulp(d) + 2e-6).all()
    # ignored rows are exactly zero
    ign = o["labels"] == -100
    assert (got[ign] == 0).all()


@pytest.mark.parametrize("N,D,V", [(256, 128, 520), (300, 64, 283), (1024, 768, 5003)])
@pytest.mark.parametrize("shift_T", [0, 10])
@pytest.mark.parametrize("with_bias", [False, True])
@pytest.mark.parametrize("gs", [0.37, -2.0])
def test_lmhead_keep_logits_route_equals_recompute_route(ops, N, D, V, shift_T, with_bias, gs):
    """Forward that keeps its bf16 logits + in-place transform == forward + recompute kernel, bit for bit; the kept
    logits are the oracle's bf16 logits (to one bf16 ulp: the accumulation order inside the MMA differs)."""
    if shift_T:
        N = N // shift_T * shift_T
    rng = np.random.default_rng(N + V + 7)
    H, Ht = bf16_tensor(rng, (N, D))
    W, Wt = bf16_tensor(rng, (V, D), scale=2.0 / np.sqrt(D))
    b = (rng.standard_normal(V) * 0.1).astype(np.float32) if with_bias else None
    bt = None if b is None else torch.from_numpy(b).to(dev())
    y = make_labels(rng, N, V)
    yt = torch.from_numpy(y).to(dev())
    ldd = (V + 7) // 8 * 8
    r0 = ops.lmhead_ce_fwd(Ht, Wt, yt, bias=bt, shift_T=shift_T)
    kept = torch.full((N, ldd), float("nan"), dtype=torch.bfloat16, device=dev())
    r1 = ops.lmhead_ce_fwd(Ht, Wt, yt, bias=bt, shift_T=shift_T, keep_logits=kept)
    assert torch.equal(r0.loss, r1.loss) and torch.equal(r0.lse, r1.lse) and torch.equal(r0.n_valid, r1.n_valid)
    o = O.lmhead_ce(H, W, y, bias=b, shift_T=shift_T, bf16_logits=True)
    got = kept[:, :V].float().cpu().numpy()
    assert np.isfinite(got).all()
    # one bf16 ulp, plus the fp32 accumulation-order difference that shows when a logit is close to zero
    assert (np.abs(got - o["logits"]) <= O.bf16_ulp(o["logits"]) + 5e-5).all()
    # (the padding columns [V, ldd) are unspecified after the forward: the TMA store clips in 16-byte units)
    gscale = torch.tensor([gs], dtype=torch.float32, device=dev())
    want = torch.full((N, ldd), float("nan"), dtype=torch.bfloat16, device=dev())
    ops.lmhead_ce_bwd_dlogits(Ht, Wt, yt, r0.lse, gscale, 0, N, want, bias=bt, shift_T=shift_T)
    # in place, in two row ranges to exercise row0
    half = min(N, 128)
    ops.ce_dlogits_inplace_(kept[:half], yt, r1.lse, gscale, 0, half, V, shift_T=shift_T)
    if half < N:
        ops.ce_dlogits_inplace_(kept[half:], yt, r1.lse, gscale, half, N - half, V, shift_T=shift_T)
    assert torch.equal(kept[:, :V].view(torch.int16), want[:, :V].view(torch.int16))
    if ldd > V:
        assert (kept[:, V:] == 0).all()                           # ... and zeroed by the transform
    # against the oracle's soft-max gradient of exactly these (kept) logits
    d = O.cross_entropy_grad(got, o["labels"], gs)
    assert (np.abs(kept[:, :V].float().cpu().numpy() - d) <= 2 * O.bf16_ulp(d) + 2e-6).all()


# ------------------------------------------------------------------------------------------ CE over rows
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,V", [(6, 11), (257, 283), (64, 5000)])
def test_ce_rows_matches_oracle(ops, dtype, N, V):
    rng = np.random.default_rng(N * V)
    x = (rng.standard_normal((N, V)) * 3).astype(np.float32)
    if dtype == torch.bfloat16:
        x = O.bf16_round(x)
    y = make_labels(rng, N, V)
    xt = torch.from_numpy(x).to(dev()).to(dtype)
    yt = torch.from_numpy(y).to(dev())
    r = ops.ce_rows_fwd(xt, yt)
    loss, n_valid, lse = O.cross_entropy(x, y)
    assert int(r.n_valid.item()) == n_valid
    np.testing.assert_allclose(r.lse[:N].cpu().numpy(), lse, atol=2e-5, rtol=0)
    if n_valid:
        np.testing.assert_allclose(r.loss_mean.item(), loss, rtol=2e-5)
    gs = torch.tensor([1.0 / max(n_valid, 1)], dtype=torch.float32, device=dev())
    d = ops.ce_rows_bwd(xt, yt, r.lse, gs).float().cpu().numpy()
    dref = O.cross_entropy_grad(x, y, 1.0 / max(n_valid, 1))
    tol = (O.bf16_ulp(dref) + 1e-7) if dtype == torch.bfloat16 else 1e-6
    assert (np.abs(d - dref) <= tol).all()
    # the ABI allows the gradient to overwrite the logits (dlogits == logits): same bits as out of place
    inplace = xt.clone()
    ops.ce_rows_bwd(inplace, yt, r.lse, gs, out=inplace)
    assert np.array_equal(inplace.float().cpu().numpy(), d)


@pytest.mark.parametrize("kept", [False, True])
@pytest.mark.parametrize("N,D,V,shift_T", [(256, 128, 520, 0), (300, 64, 283, 10), (640, 256, 1031, 64)])
def test_lmhead_ce_bwd_one_call_matches_oracle_gradients(ops, N, D, V, shift_T, kept):
    """b2_lmhead_ce_bwd: d(logits), dH, dW and the bias gradient db behind one entry point, from kept logits and by
    recompute, against the oracle's d(logits) pushed through fp64 matrix products; a second call accumulates dW / db."""
    rng = np.random.default_rng(N * 3 + V)
    H, Ht = bf16_tensor(rng, (N, D))
    W, Wt = bf16_tensor(rng, (V, D), scale=2.0 / np.sqrt(D))
    b = (rng.standard_normal(V) * 0.1).astype(np.float32)
    bt = torch.from_numpy(b).to(dev())
    y = make_labels(rng, N, V)
    yt = torch.from_numpy(y).to(dev())
    ldd = (V + 63) // 64 * 64
    dl = torch.full((N, ldd), float("nan"), dtype=torch.bfloat16, device=dev())
    r = ops.lmhead_ce_fwd(Ht, Wt, yt, bias=bt, shift_T=shift_T, keep_logits=dl if kept else None)
    gs = torch.tensor([0.5 / N], dtype=torch.float32, device=dev())
    dH = torch.empty(N, D, dtype=torch.bfloat16, device=dev())
    dW = torch.empty(V, D, dtype=torch.float32, device=dev())
    db = torch.full((V,), float("nan"), dtype=torch.float32, device=dev())
    ops.lmhead_ce_bwd(Ht, Wt, yt, r.lse, gs, 0, N, dl, kept=kept, dH=dH, dW=dW, db=db, bias=bt, shift_T=shift_T)
    o = O.lmhead_ce(H, W, y, bias=b, shift_T=shift_T, bf16_logits=True)
    d = O.bf16_round(O.cross_entropy_grad(o["logits"], o["labels"], 0.5 / N)).astype(np.float64)
    want_dH, want_dW, want_db = d @ W.astype(np.float64), d.T @ H.astype(np.float64), d.sum(axis=0)
    got = dl[:, :V].float().cpu().numpy()
    assert (np.abs(got - d) <= O.bf16_ulp(d.astype(np.float32)) + 2e-6).all()
    for g, w, name in ((dH.float().cpu().numpy(), want_dH, "dH"), (dW.cpu().numpy(), want_dW, "dW"), (db.cpu().numpy(), want_db, "db")):
        rel = np.linalg.norm(g - w) / max(np.linalg.norm(w), 1e-30)
        assert np.isfinite(g).all() and rel < 1e-2, f"{name}: relative error {rel:.3e}"
    np.testing.assert_allclose(db.cpu().numpy(), got.astype(np.float64).sum(axis=0), rtol=1e-5, atol=1e-7)   # exact sums of the stored values
    # accumulate: the same chunk again on top (recompute route, the kept logits are gone)
    dW2, db2 = dW.clone(), db.clone()
    ops.lmhead_ce_bwd(Ht, Wt, yt, r.lse, gs, 0, N, dl, kept=False, dW=dW2, dW_accumulate=True, db=db2, db_accumulate=True,
                      bias=bt, shift_T=shift_T)
    np.testing.assert_allclose(dW2.cpu().numpy(), 2 * dW.cpu().numpy(), rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(db2.cpu().numpy(), 2 * db.cpu().numpy(), rtol=1e-5, atol=1e-8)


# ------------------------------------------------------------------------------------------ gradient landing
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_land_copy_moves_many_ragged_segments_in_one_call(ops, dtype):
    """b2_land_copy: bit-exact copies of many separately allocated tensors into their slices of a flat buffer — sizes
    from one element to several chunks, more segments than one launch holds, untouched gaps in between."""
    g = torch.Generator(device="cpu").manual_seed(5)
    sizes = [1, 7, 8, 9, 63, 64, 1000, 16384, 16385, (1 << 20) + 3, 40000] + [int(x) for x in torch.randint(1, 5000, (140,), generator=g)]
    srcs = [torch.randn(n, generator=g).to(dev(), dtype) for n in sizes]
    offs, off = [], 0
    for n in sizes:
        offs.append(off)
        off += (n + 63) // 64 * 64 + 64            # 64-element aligned slices with a gap after each
    flat = torch.full((off,), 7.0, dtype=dtype, device=dev())
    ops.land_copy(srcs, flat, offs)
    want = torch.full((off,), 7.0, dtype=dtype, device=dev())
    for t, o in zip(srcs, offs):
        want[o:o + t.numel()] = t
    assert torch.equal(flat, want)
    ops.land_copy([], flat, [])                     # nothing to do is legal
    with pytest.raises(Exception):
        ops.land_copy([srcs[6][1:]], flat, [0])     # a source that is not 16-byte aligned is refused, not mis-copied


# ------------------------------------------------------------------------------------------ clip
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [1, 7, 8, 1000, 1 << 20, (1 << 22) + 13])
def test_sumsq_clip_scale(ops, dtype, n):
    rng = np.random.default_rng(n)
    g = (rng.standard_normal(n) * 0.1).astype(np.float32)
    if dtype == torch.bfloat16:
        g = O.bf16_round(g)
    gt = torch.from_numpy(g).to(dev()).to(dtype)
    ss = ops.sumsq(gt)
    ref = (g.astype(np.float64) ** 2).sum()
    np.testing.assert_allclose(ss.item(), ref, rtol=1e-5)
    ss2 = ops.sumsq(gt, out=ss, accumulate=True)          # scratch reuse + accumulate
    np.testing.assert_allclose(ss2.item(), 2 * ref, rtol=1e-5)
    total, coef, clipped = O.clip_grad_norm([g], 0.3)
    norm_t, coef_t = ops.clip_coef(ops.sumsq(gt).reshape(1), 0.3)
    np.testing.assert_allclose(norm_t.item(), total, rtol=1e-5)
    np.testing.assert_allclose(coef_t.item(), coef, rtol=1e-5)
    ops.scale_(gt, coef_t)
    got = gt.float().cpu().numpy()
    want = clipped[0] if dtype == torch.float32 else O.bf16_round(clipped[0])
    tol = 1e-5 * np.abs(want) + (O.bf16_ulp(want) if dtype == torch.bfloat16 else 1e-9)
    assert (np.abs(got - want) <= tol).all()


# ------------------------------------------------------------------------------------------ AdamW
@pytest.mark.parametrize("n", [5, 257, 1 << 16, (1 << 20) + 3])
def test_adamw_fp32_matches_oracle(ops, n):
    rng = np.random.default_rng(n)
    p = rng.standard_normal(n).astype(np.float32)
    m = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
    pt, mt, vt = (torch.from_numpy(a.copy()).to(dev()) for a in (p, m, v))
    hp = dict(lr=3e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01)
    for step in range(1, 6):
        g = (rng.standard_normal(n) * 0.01).astype(np.float32)
        ops.adamw_(pt, torch.from_numpy(g).to(dev()), mt, vt, step=step, **hp)
        p, m, v = O.adamw_step(p, g, m, v, step=step, **hp)
    # fp32: <= 2 ulp per step against the oracle's torch-ordered arithmetic
    for got, want in ((pt, p), (mt, m), (vt, v)):
        err = np.abs(got.cpu().numpy().astype(np.float64) - want)
        assert (err <= 10 * O.f32_ulp(want) + 1e-30).all()


@pytest.mark.parametrize("state_dtype", [torch.bfloat16, torch.float32])
def test_adamw_bf16_params(ops, state_dtype):
    n = 4099
    rng = np.random.default_rng(3)
    p = O.bf16_round(rng.standard_normal(n).astype(np.float32))
    g = O.bf16_round((rng.standard_normal(n) * 0.01).astype(np.float32))
    m0 = np.zeros(n, np.float32); v0 = np.zeros(n, np.float32)
    pt = torch.from_numpy(p).to(dev()).to(torch.bfloat16)
    gt = torch.from_numpy(g).to(dev()).to(torch.bfloat16)
    mt = torch.zeros(n, dtype=state_dtype, device=dev()); vt = torch.zeros(n, dtype=state_dtype, device=dev())
    hp = dict(lr=2e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-3)
    ops.adamw_(pt, gt, mt, vt, step=1, **hp)
    # exact-arithmetic result, rounded once: the kernel must be within 1 bf16 ulp of it (and so is torch's
    # op-by-op bf16 path, which the oracle models with state_bf16=True)
    pe, me, ve = O.adamw_step(p, g, m0, v0, step=1, state_bf16=False, **hp)
    assert (np.abs(pt.float().cpu().numpy() - pe) <= O.bf16_ulp(np.maximum(np.abs(pe), np.abs(p)))).all()
    if state_dtype == torch.bfloat16:
        po, mo, vo = O.adamw_step(p, g, m0, v0, step=1, state_bf16=True, **hp)
        assert (np.abs(mt.float().cpu().numpy() - mo) <= O.bf16_ulp(mo) + 1e-38).all()
        assert (np.abs(vt.float().cpu().numpy() - vo) <= O.bf16_ulp(vo) + 1e-38).all()
        # 1 bf16 ulp at the magnitude of the operands (an update that crosses zero cancels, so the ulp of the tiny
        # result is not the right yardstick) + 2 % of the update size lr*|m^/sqrt(v^)| <= lr: torch's op-by-op bf16
        # rounding of m, v and the denominator perturbs the update itself by about three half-ulps (2^-9 each)
        assert (np.abs(pt.float().cpu().numpy() - po) <= O.bf16_ulp(np.maximum(np.abs(po), np.abs(p))) + 0.02 * hp["lr"]).all()
    else:
        np.testing.assert_allclose(mt.cpu().numpy(), me, rtol=1e-6, atol=1e-12)
        np.testing.assert_allclose(vt.cpu().numpy(), ve, rtol=1e-6, atol=1e-16)


def test_adamw_clip_coef_is_applied(ops):
    n = 1000
    rng = np.random.default_rng(9)
    p = rng.standard_normal(n).astype(np.float32); g = rng.standard_normal(n).astype(np.float32)
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0)
    total, coef, clipped = O.clip_grad_norm([g], 1.0)
    pt = torch.from_numpy(p.copy()).to(dev()); gt = torch.from_numpy(g).to(dev())
    mt = torch.zeros(n, device=dev()); vt = torch.zeros(n, device=dev())
    _, coef_t = ops.clip_coef(ops.sumsq(gt).reshape(1), 1.0)
    ops.adamw_(pt, gt, mt, vt, step=1, gscale=coef_t, **hp)
    pe, me, ve = O.adamw_step(p, clipped[0], np.zeros(n, np.float32), np.zeros(n, np.float32), step=1, **hp)
    np.testing.assert_allclose(mt.cpu().numpy(), me, rtol=2e-6)
    np.testing.assert_allclose(pt.cpu().numpy(), pe, rtol=1e-6, atol=1e-7)


def test_gemm_split_k_matches_unsplit(ops):
    """dH-shaped GEMM (few output tiles, long K): the split-K path must agree with the single-pass path."""
    rng = np.random.default_rng(21)
    M, N, K = 512, 768, 50257
    ldk = (K + 7) // 8 * 8
    A = torch.zeros(M, ldk, dtype=torch.bfloat16, device=dev()); A[:, :K] = torch.from_numpy(O.bf16_round(rng.standard_normal((M, K)).astype(np.float32) * 0.01)).to(dev())
    B = torch.from_numpy(O.bf16_round(rng.standard_normal((K, N)).astype(np.float32))).to(dev()).to(torch.bfloat16)   # MN-major B
    C1 = torch.empty(M, N, dtype=torch.bfloat16, device=dev()); C2 = torch.empty_like(C1)
    ops.gemm_bf16(A, B, C1, M, N, K, b_mn_major=True, split_k=False)
    ops.gemm_bf16(A, B, C2, M, N, K, b_mn_major=True, split_k=True)
    ref = A[:, :K].float() @ B.float()
    for C in (C1, C2):
        assert ((C.float() - ref).abs() <= 2.0 ** -7 * ref.abs() + 2e-3).all()
    # accumulate through the split path
    C3 = torch.ones(M, N, dtype=torch.float32, device=dev())
    ops.gemm_bf16(A, B, C3, M, N, K, b_mn_major=True, accumulate=True)
    assert ((C3 - 1 - ref).abs() <= 1e-3 * ref.abs().max()).all()


def test_empty_inputs(ops):
    """Zero rows / zero elements are legal everywhere: no launch over nothing, torch's conventions for the results."""
    H = torch.zeros(0, 64, dtype=torch.bfloat16, device=dev())
    W = torch.zeros(32, 64, dtype=torch.bfloat16, device=dev())
    y = torch.zeros(0, dtype=torch.int64, device=dev())
    r = ops.lmhead_ce_fwd(H, W, y)
    assert int(r.n_valid.item()) == 0 and float(r.loss_sum.item()) == 0.0 and np.isnan(r.loss_mean.item())
    r = ops.ce_rows_fwd(torch.zeros(0, 7, device=dev()), y)
    assert int(r.n_valid.item()) == 0 and np.isnan(r.loss_mean.item())
    e = torch.zeros(0, device=dev())
    assert float(ops.sumsq(e).item()) == 0.0
    ops.adamw_(e, e, e.clone(), e.clone(), lr=1e-3, step=1)          # no-op, must not fail
    ops.scale_(e, torch.ones(1, device=dev()))
