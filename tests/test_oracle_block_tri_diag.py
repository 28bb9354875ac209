"""
Ports of the reference's unit tests that pin the block-tridiagonal layer, run against the oracle.
Reference: /root/reference/tests/unit/test_block_tri_diag.py (cholesky :97-107, solve :114-136,
abs_log_det :82-91, dense_mult :144-180, diagonal_of_inverse :186-202, upper_diagonal_lower :207-225).
All comparisons are against dense NumPy linear algebra, as in the reference.
"""
import numpy as np
import pytest

from oracle.gauss_markov_np import (LowerTriangularBlockTriDiagonal, SymmetricBlockTriDiagonal)

INNER = [1, 3]
OUTER = [1, 4]
BATCH = [(3,), (), (2, 1)]


def _spd_btd(rng, batch, outer, inner, sub=True):
    """A well conditioned SPD block-tridiagonal matrix L L^T from a random lower block-bidiagonal L."""
    Ld = np.tril(rng.normal(size=batch + (outer, inner, inner))) + 3.0 * np.eye(inner)
    Ls = rng.normal(size=batch + (outer - 1, inner, inner)) if (sub and outer > 1) else None
    L = LowerTriangularBlockTriDiagonal(Ld, Ls)
    dense = L.to_dense() @ np.swapaxes(L.to_dense(), -1, -2)
    n = outer
    Kd = np.stack([dense[..., k * inner:(k + 1) * inner, k * inner:(k + 1) * inner] for k in range(n)], axis=-3)
    Ks = None
    if Ls is not None:
        Ks = np.stack([dense[..., (k + 1) * inner:(k + 2) * inner, k * inner:(k + 1) * inner] for k in range(n - 1)],
                      axis=-3)
    return SymmetricBlockTriDiagonal(Kd, Ks), L, dense


@pytest.mark.parametrize("inner", INNER)
@pytest.mark.parametrize("outer", OUTER)
@pytest.mark.parametrize("batch", BATCH)
@pytest.mark.parametrize("sub", [True, False])
def test_cholesky(inner, outer, batch, sub):
    rng = np.random.default_rng(71892305)
    sym, _, dense = _spd_btd(rng, batch, outer, inner, sub)
    np.testing.assert_allclose(sym.to_dense(), dense, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(sym.cholesky.to_dense(), np.linalg.cholesky(dense), rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("inner", INNER)
@pytest.mark.parametrize("outer", OUTER)
@pytest.mark.parametrize("batch", BATCH)
@pytest.mark.parametrize("transpose", [False, True])
def test_solve(inner, outer, batch, transpose):
    rng = np.random.default_rng(1)
    _, L, _ = _spd_btd(rng, batch, outer, inner)
    rhs = rng.normal(size=batch + (outer, inner))
    dense = L.to_dense()
    if transpose:
        dense = np.swapaxes(dense, -1, -2)
    expect = (np.linalg.inv(dense) @ rhs.reshape(batch + (outer * inner, 1))).reshape(batch + (outer, inner))
    np.testing.assert_allclose(L.solve(rhs, transpose_left=transpose), expect, rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("inner", INNER)
@pytest.mark.parametrize("outer", OUTER)
@pytest.mark.parametrize("batch", BATCH)
def test_abs_log_det_and_dense_mult(inner, outer, batch):
    rng = np.random.default_rng(2)
    sym, L, dense = _spd_btd(rng, batch, outer, inner)
    _, ld = np.linalg.slogdet(L.to_dense())
    np.testing.assert_allclose(L.abs_log_det(), ld, rtol=1e-10, atol=1e-12)
    x = rng.normal(size=batch + (outer, inner))
    xf = x.reshape(batch + (outer * inner, 1))
    np.testing.assert_allclose(sym.dense_mult(x), (dense @ xf).reshape(x.shape), rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(L.dense_mult(x), (L.to_dense() @ xf).reshape(x.shape), rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(L.dense_mult(x, transpose_left=True),
                               (np.swapaxes(L.to_dense(), -1, -2) @ xf).reshape(x.shape), rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("inner", INNER)
@pytest.mark.parametrize("outer", OUTER)
@pytest.mark.parametrize("batch", BATCH)
def test_diagonal_of_inverse(inner, outer, batch):
    rng = np.random.default_rng(3)
    _, L, dense = _spd_btd(rng, batch, outer, inner)
    inv = np.linalg.inv(dense)
    expect = np.stack([inv[..., k * inner:(k + 1) * inner, k * inner:(k + 1) * inner] for k in range(outer)], axis=-3)
    np.testing.assert_allclose(L.block_diagonal_of_inverse(), expect, rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("inner", INNER)
@pytest.mark.parametrize("batch", BATCH)
def test_upper_diagonal_lower(inner, batch):
    rng = np.random.default_rng(4)
    sym, _, dense = _spd_btd(rng, batch, 4, inner)
    Ut, cholD = sym.upper_diagonal_lower()
    U = np.swapaxes(Ut.to_dense(), -1, -2)
    cd = cholD.to_dense()
    np.testing.assert_allclose(U @ (cd @ np.swapaxes(cd, -1, -2)) @ np.swapaxes(U, -1, -2), dense, rtol=1e-9,
                               atol=1e-10)


def test_add():
    rng = np.random.default_rng(5)
    a, _, da = _spd_btd(rng, (), 4, 3)
    b, _, db = _spd_btd(rng, (), 4, 3)
    np.testing.assert_allclose((a + b).to_dense(), da + db, rtol=1e-12)
