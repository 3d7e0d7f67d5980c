"""CPU tests of the oracle (oracle/): pins the restatement before anything trusts it.

The reference holds no golden vectors, so the pins are: analytic known
answers (KAT-1..4, 6), central finite differences (KAT-5), an independent numpy restatement,
and PyTorch-autograd fixtures (tests/golden/).  The properties the reference's own tests assert
(tests/test_forward.cpp:25-26,49-50; tests/test_backward.cpp:30-31,46-48; python/test.py:57-79)
are checked on the oracle too, on the same shapes.
"""
import math

import numpy as np
import pytest

from conftest import seeded_z


def test_kat1_analytic(oracle):
    z = np.array([[1, 0], [0, 1], [1, 0], [0, 1]], dtype=np.float64)
    loss, lse, dz = oracle.forward_backward(z, 1.0)
    assert loss == pytest.approx(math.log(math.e + 2) - 1, abs=1e-14)
    g = 1.0 / (math.e + 2)
    np.testing.assert_allclose(dz[0], [0.0, g], atol=1e-14)
    np.testing.assert_allclose(dz[1], [g, 0.0], atol=1e-14)
    np.testing.assert_allclose(dz[2], [0.0, g], atol=1e-14)


def test_kat2_swapped_pairs(oracle):
    z = np.array([[1, 0], [0, 1], [0, 1], [1, 0]], dtype=np.float64)
    assert oracle.forward(z, 1.0)[0] == pytest.approx(math.log(math.e + 2), abs=1e-14)


def test_kat3_identical_rows(oracle):
    z = np.tile(np.array([[0.3, -1.2, 0.5]]), (10, 1))
    loss, lse, dz = oracle.forward_backward(z, 0.3)
    assert loss == pytest.approx(math.log(9), abs=1e-12)
    assert np.abs(dz).max() < 1e-12


def test_kat4_scale_invariance_and_orthogonality(oracle):
    z = seeded_z(32, 16, seed=3, dtype=np.float64)
    l1, _, d1 = oracle.forward_backward(z, 0.2)
    l2, _, d2 = oracle.forward_backward(7.0 * z, 0.2)
    assert l1 == pytest.approx(l2, rel=1e-13)
    np.testing.assert_allclose(d2, d1 / 7.0, rtol=1e-10, atol=1e-15)
    assert np.abs((z * d1).sum(axis=1)).max() < 1e-14


def test_kat5_finite_differences(oracle):
    z = seeded_z(8, 5, seed=5, dtype=np.float64)
    tau, h = 0.5, 1e-6
    _, _, dz = oracle.forward_backward(z, tau)
    num = np.zeros_like(z)
    for i in range(z.shape[0]):
        for k in range(z.shape[1]):
            zp, zm = z.copy(), z.copy()
            zp[i, k] += h
            zm[i, k] -= h
            num[i, k] = (oracle.forward(zp, tau)[0] - oracle.forward(zm, tau)[0]) / (2 * h)
    assert np.abs(num - dz).max() < 1e-8


def test_temperature_gradient_finite_differences(oracle):
    """d loss / d tau of the numpy restatement against central differences, and the identity the
    CUDA epilogue uses: sum_a dU_a . u_a = -2 tau dL/dtau."""
    z = seeded_z(24, 10, seed=11)
    for tau in (0.1, 0.5, 2.0):
        r = oracle.numpy_forward_backward(z, tau)
        h = 1e-6 * tau
        fd = (oracle.numpy_forward_backward(z, tau + h)["loss"] -
              oracle.numpy_forward_backward(z, tau - h)["loss"]) / (2 * h)
        assert abs(r["dtau"] - fd) <= 1e-6 * max(1.0, abs(fd))
        assert abs(-0.5 / tau * (r["dU"] * r["u"]).sum() - r["dtau"]) <= 1e-12 * max(1.0, abs(fd))
        # the C restatement (direct sum over P_ij S_ij) agrees with the numpy one
        _, lse, _ = oracle.forward(z, tau)
        assert abs(oracle.dtau(z, lse, tau) - r["dtau"]) <= 1e-12 * max(1.0, abs(fd))
        assert abs(oracle.dtau(z, lse, tau, grad_out=-2.5) + 2.5 * r["dtau"]) <= 1e-11 * max(1.0, abs(fd))


def test_kat6_pair_permutation(oracle):
    z = seeded_z(24, 12, seed=6, dtype=np.float64)
    N = 12
    perm = np.random.default_rng(0).permutation(N)
    zp = np.concatenate([z[:N][perm], z[N:][perm]])
    assert oracle.forward(z, 0.3)[0] == pytest.approx(oracle.forward(zp, 0.3)[0], rel=1e-13)


def test_grad_out_scales_gradient(oracle):
    z = seeded_z(16, 8, seed=8, dtype=np.float64)
    loss, lse, _ = oracle.forward(z, 0.4)[0], oracle.forward(z, 0.4)[1], None
    d1, _ = oracle.backward(z, lse, 0.4, 1.0)
    d3, _ = oracle.backward(z, lse, 0.4, -2.5)
    np.testing.assert_allclose(d3, -2.5 * d1, rtol=1e-13)


@pytest.mark.parametrize("M,D,tau", [(2, 8, 0.5), (6, 3, 1.0), (64, 128, 0.07), (130, 24, 0.1)])
def test_c_matches_numpy_restatement(oracle, M, D, tau):
    z = seeded_z(M, D, seed=M + D, dtype=np.float64)
    loss, lse, dz = oracle.forward_backward(z, tau, grad_out=1.5)
    r = oracle.numpy_forward_backward(z, tau, grad_out=1.5)
    assert loss == pytest.approx(r["loss"], rel=1e-12, abs=1e-13)
    np.testing.assert_allclose(lse, r["lse"], rtol=1e-12)
    np.testing.assert_allclose(dz, r["dz"], rtol=1e-9, atol=1e-14)
    _, pp = oracle.backward(z, lse, tau)
    np.testing.assert_allclose(pp, r["pos_prob"], rtol=1e-10)


def test_golden_fixtures(oracle, golden):
    """Oracle vs PyTorch-autograd fixtures (tests/golden/make_golden.py)."""
    assert len(golden) >= 9
    for name, c in golden.items():
        tau = float(c["tau"])
        loss, lse, dz = oracle.forward_backward(c["z"], tau)
        assert loss == pytest.approx(float(c["loss"]), rel=1e-11, abs=1e-12), name
        np.testing.assert_allclose(lse, c["lse"], rtol=1e-11, atol=1e-11, err_msg=name)
        scale = np.abs(c["dz"]).max()
        assert np.abs(dz - c["dz"]).max() <= 1e-9 * max(scale, 1e-300), name
        # d loss / d tau: both restatements against autograd's
        ref = float(c["dtau"])
        assert oracle.dtau(c["z"], lse, tau) == pytest.approx(ref, rel=1e-9, abs=1e-12), name
        assert oracle.numpy_forward_backward(c["z"], tau)["dtau"] == pytest.approx(ref, rel=1e-9, abs=1e-12), name


def test_reference_test_properties_hold(oracle):
    """Every property the reference's tests assert, on the reference's shapes."""
    rng = np.random.default_rng(11)
    for B in (16, 32, 64, 128):  # tests/test_forward.cpp:40-52 (rows = B here, stacked views)
        z = rng.standard_normal((B, 128))
        z /= np.linalg.norm(z, axis=1, keepdims=True)
        loss, lse, dz = oracle.forward_backward(z, 0.07)
        assert loss > 0 and math.isfinite(loss)
        assert np.isfinite(dz).all()
        n = np.linalg.norm(dz)
        assert 0 < n < 100  # tests/test_backward.cpp:46-48
    z = rng.standard_normal((128, 256))
    z /= np.linalg.norm(z, axis=1, keepdims=True)
    for scale in (1e-5, 1.0, 1e5):  # python/test.py:64-65
        for tau in (0.01, 0.07, 1.0):
            loss = oracle.forward(z * scale, tau)[0]
            assert math.isfinite(loss)


def test_rowblock_f32_matches_f64(oracle):
    """The timed fp32 row-block form is the same function: full block and a 4-way partition."""
    M, D, tau = 256, 64, 0.3
    z = seeded_z(M, D, seed=21)
    r = oracle.numpy_forward_backward(z.astype(np.float64), tau)
    lse, lp, du_row, du_col = oracle.rowblock_f32(z, tau, 0, M)
    assert lp / M == pytest.approx(r["loss"], rel=1e-5)
    np.testing.assert_allclose(lse, r["lse"], rtol=1e-5)
    dU = (du_row + du_col) / (tau * M)
    assert np.abs(dU - r["dU"]).max() <= 1e-5 * np.abs(r["dU"]).max()
    # partitioned: sum of column terms over ranks + own row term
    R = 4
    col_total = np.zeros((M, D), dtype=np.float64)
    rows = []
    loss_total = 0.0
    for rk in range(R):
        l, lp, dr, dc = oracle.rowblock_f32(z, tau, rk * M // R, M // R)
        rows.append(dr)
        col_total += dc
        loss_total += lp
    dU2 = (np.concatenate(rows) + col_total) / (tau * M)
    assert loss_total / M == pytest.approx(r["loss"], rel=1e-5)
    assert np.abs(dU2 - r["dU"]).max() <= 1e-5 * np.abs(r["dU"]).max()


def test_literal_reference_is_a_different_function(oracle):
    """Documents SURVEY.md section 0.4: what src/ntxent_kernel.cu evaluates is not NT-Xent."""
    rng = np.random.default_rng(0)
    z = rng.standard_normal((32, 128))
    z /= np.linalg.norm(z, axis=1, keepdims=True)
    literal = oracle.literal_reference_forward(z, 0.07)
    canonical = oracle.forward(np.concatenate([z, z]), 0.07)[0]
    assert math.isfinite(literal) and math.isfinite(canonical)
    # same stacked input, an order of magnitude apart (0.00113 vs 9.6e-5 at this seed)
    assert literal > 5 * canonical


def test_bad_arguments(oracle):
    with pytest.raises(ValueError):
        oracle.forward(np.zeros((3, 4)), 0.5)  # odd M
    with pytest.raises(ValueError):
        oracle.forward(np.zeros((4, 4)), 0.0)  # tau <= 0


# ------------------------------------------------ arbitrary positives
def test_arbitrary_positives_c_against_numpy_and_finite_differences(oracle):
    rng = np.random.default_rng(3)
    M, D, tau = 24, 8, 0.3
    z = rng.standard_normal((M, D))
    pos = np.array([(i + 5) % M for i in range(M)])
    pos[3] = pos[9] = 7  # not an involution, and two rows share a positive
    loss, lse, dz = oracle.forward_backward_pos(z, pos, tau, 1.5)
    ref = oracle.numpy_forward_backward(z, tau, 1.5, pos=pos)
    assert loss == pytest.approx(ref["loss"], rel=1e-12)
    np.testing.assert_allclose(dz, ref["dz"], rtol=0, atol=1e-13)
    # the default positives are the canonical NT-Xent
    l2, _, dz2 = oracle.forward_backward_pos(z, (np.arange(M) + M // 2) % M, tau, 1.5)
    l3, _, dz3 = oracle.forward_backward(z, tau, 1.5)
    assert l2 == pytest.approx(l3, rel=1e-14)
    np.testing.assert_allclose(dz2, dz3, rtol=0, atol=1e-14)
    eps = 1e-6
    for i, k in ((0, 0), (3, 2), (7, 5), (9, 1)):
        zp, zm = z.copy(), z.copy()
        zp[i, k] += eps
        zm[i, k] -= eps
        fd = (oracle.forward_backward_pos(zp, pos, tau)[0] - oracle.forward_backward_pos(zm, pos, tau)[0]) / (2 * eps)
        assert 1.5 * fd == pytest.approx(dz[i, k], abs=1e-8)
    with pytest.raises(ValueError):
        oracle.forward_backward_pos(z, np.arange(M), tau)  # a row cannot be its own positive


def test_cross_view_is_the_symmetric_two_tower_infonce(oracle):
    """cross_n: the denominators over the other view only.  Against the CLIP formula written out,
    against the numpy restatement, and by finite differences."""
    rng = np.random.default_rng(5)
    N, D, tau = 10, 7, 0.25
    z = rng.standard_normal((2 * N, D))
    loss, lse, dz = oracle.forward_backward_cross(z, tau, 0.5)
    u = z / np.linalg.norm(z, axis=1, keepdims=True)
    L = u[:N] @ u[N:].T / tau
    ce = lambda A: np.mean(np.log(np.exp(A).sum(1)) - np.diag(A))  # noqa: E731
    assert loss == pytest.approx((ce(L) + ce(L.T)) / 2, rel=1e-12)
    ref = oracle.numpy_forward_backward(z, tau, 0.5, cross_view=True)
    assert loss == pytest.approx(ref["loss"], rel=1e-12)
    np.testing.assert_allclose(dz, ref["dz"], rtol=0, atol=1e-13)
    eps = 1e-6
    for i, k in ((0, 0), (4, 3), (13, 6)):
        zp, zm = z.copy(), z.copy()
        zp[i, k] += eps
        zm[i, k] -= eps
        fd = (oracle.forward_backward_cross(zp, tau)[0] - oracle.forward_backward_cross(zm, tau)[0]) / (2 * eps)
        assert 0.5 * fd == pytest.approx(dz[i, k], abs=1e-8)
