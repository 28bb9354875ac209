"""
Mirror of ``markovflow/block_tri_diag.py``: block-tridiagonal matrices stored by their diagonal blocks
``[..., n, D, D]`` and sub-diagonal blocks ``[..., n-1, D, D]``.  The banded_matrices custom ops the
reference calls (``cholesky_band``, ``inverse_from_cholesky_band``, ``solve_triang_mat``,
``product_band_mat``; block_tri_diag.py:22-31) are the ``vdp_btd_*`` entries of the C ABI here, in block
layout, so no band packing is needed.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import ops
from ._util import batched, unbatched


class BlockTriDiagonal:
    """block_tri_diag.py:37-289"""

    def __init__(self, diagonal: torch.Tensor, symmetric: bool, sub_diagonal: Optional[torch.Tensor] = None):
        self._diag = diagonal
        self._sub_diag = sub_diagonal
        self._symmetric = symmetric

    @property
    def batch_shape(self):
        return tuple(self._diag.shape[:-3])

    @property
    def inner_dim(self) -> int:
        return self._diag.shape[-1]

    @property
    def outer_dim(self) -> int:
        return self._diag.shape[-3]

    @property
    def bandwidth(self) -> int:
        # block_tri_diag.py:100-108
        return self.inner_dim * 2 - 1 if self._sub_diag is not None else self.inner_dim - 1

    @property
    def block_diagonal(self) -> torch.Tensor:
        return self._diag

    @property
    def block_sub_diagonal(self) -> Optional[torch.Tensor]:
        return self._sub_diag

    def _b(self):
        d, had = batched(self._diag, 3)
        s = None if self._sub_diag is None else batched(self._sub_diag, 3)[0]
        return d, s, had

    def to_dense(self) -> torch.Tensor:
        """block_tri_diag.py:150-173 (small matrices only: this builds the dense [n D, n D] matrix)."""
        d, s, had = self._b()
        B, n, D, _ = d.shape
        cols = []
        for j in range(n * D):
            e = torch.zeros((B, n * D), dtype=d.dtype, device=d.device)
            e[:, j] = 1.0
            cols.append(ops.btd_matvec(d, s, e.reshape(B, n, D), self._symmetric, False).reshape(B, n * D))
        return unbatched(torch.stack(cols, dim=-1), had)

    @property
    def as_band(self) -> torch.Tensor:
        """block_tri_diag.py:110-118, 206-237: lower-band storage [..., R, n*D], band[r, j] = M[j + r, j]."""
        d, s, had = self._b()
        return unbatched(ops.btd_block_to_band(d, s), had)

    def dense_mult(self, right: torch.Tensor, transpose_left: bool = False) -> torch.Tensor:
        """block_tri_diag.py:175-199: ``self @ right`` (or ``self^T @ right``), right [..., n, D]."""
        d, s, had = self._b()
        r, _ = batched(right, 2)
        return unbatched(ops.btd_matvec(d, s, r, self._symmetric, transpose_left), had)


class LowerTriangularBlockTriDiagonal(BlockTriDiagonal):
    """block_tri_diag.py:291-381"""

    def __init__(self, diagonal: torch.Tensor, sub_diagonal: Optional[torch.Tensor] = None) -> None:
        super().__init__(diagonal, False, sub_diagonal)

    def block_diagonal_of_inverse(self) -> torch.Tensor:
        """Block diagonal of (L L^T)^-1   (block_tri_diag.py:318-337)."""
        d, s, had = self._b()
        Sd, _ = ops.btd_inverse_band(d, s, want_sub=False)
        return unbatched(Sd, had)

    def solve(self, right: torch.Tensor, transpose_left: bool = False) -> torch.Tensor:
        """L^-1 right or L^-T right   (block_tri_diag.py:339-351)."""
        d, s, had = self._b()
        r, _ = batched(right, 2)
        return unbatched(ops.btd_solve(d, s, r, transpose_left), had)

    def abs_log_det(self) -> torch.Tensor:
        """block_tri_diag.py:353-366"""
        d, s, had = self._b()
        out = ops.btd_logdet(d)
        return out[0] if had is False else (out if had is True else out.reshape(had))

    def __add__(self, other: "LowerTriangularBlockTriDiagonal") -> "LowerTriangularBlockTriDiagonal":
        sub = _add_sub(self._sub_diag, other._sub_diag)
        return LowerTriangularBlockTriDiagonal(self._diag + other._diag, sub)


class SymmetricBlockTriDiagonal(BlockTriDiagonal):
    """block_tri_diag.py:384-549"""

    def __init__(self, diagonal: torch.Tensor, sub_diagonal: Optional[torch.Tensor] = None) -> None:
        super().__init__(diagonal, True, sub_diagonal)

    def __add__(self, other: "SymmetricBlockTriDiagonal") -> "SymmetricBlockTriDiagonal":
        """block_tri_diag.py:412-425"""
        return SymmetricBlockTriDiagonal(self._diag + other._diag, _add_sub(self._sub_diag, other._sub_diag))

    @property
    def cholesky(self) -> LowerTriangularBlockTriDiagonal:
        """block_tri_diag.py:428-440 (cholesky_band); raises if a block is not positive definite."""
        d, s, had = self._b()
        status = ops.new_status(d.device)
        Ld, Ls = ops.btd_cholesky(d, s, status)
        code, idx = ops.read_status(status)
        if code != 0:
            raise ValueError(f"SymmetricBlockTriDiagonal.cholesky: block {idx} is not positive definite")
        return LowerTriangularBlockTriDiagonal(unbatched(Ld, had), None if Ls is None else unbatched(Ls, had))


def _add_sub(a, b):
    if a is None:
        return b
    if b is None:
        return a
    return a + b
