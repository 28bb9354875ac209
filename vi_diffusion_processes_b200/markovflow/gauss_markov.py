"""Mirror of ``markovflow/gauss_markov.py``: the BTDGaussian container (gauss_markov.py:220-242)."""
from __future__ import annotations

from .block_tri_diag import SymmetricBlockTriDiagonal


class BTDGaussian:
    """Gaussian in natural parameters with a block-tridiagonal second parameter (gauss_markov.py:220-242)."""

    def __init__(self, nat1, nat2_diag, nat2_subdiag):
        self.nat1 = nat1
        self.nat2 = SymmetricBlockTriDiagonal(nat2_diag, nat2_subdiag)
