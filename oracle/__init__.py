"""
CPU oracle for the VDP / CVI-DP site-update hot path.

TEST INFRASTRUCTURE ONLY.  This package is a NumPy float64 restatement of the
reference algorithm (AaltoML/vi-diffusion-processes, a markovflow fork).  It is
imported by ``tests/``, by ``__graft_entry__.smoke()`` and by the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- and by nothing
in the product package ``vi_diffusion_processes_b200``.

Parity status: **unpinned by the reference**.  The reference ships no golden
vectors for the model-level path and its TensorFlow stack (tensorflow, gpflow,
tensorflow_probability, banded_matrices) cannot be installed in the build
container, so the oracle cannot be checked against outputs of the reference
itself.  What pins it instead (see tests/test_oracle_*.py):
  * the reference's own unit tests for the block-tridiagonal / state-space /
    transformation layers, ported one to one (they compare against dense
    NumPy linear algebra, so they apply to the oracle unchanged);
  * closed forms (polynomial drifts are integrated exactly by Gauss-Hermite);
  * reverse-mode gradients checked against torch.autograd on a restated
    forward pass and against central finite differences;
  * an 80-bit long-double re-evaluation for state dimension 1.
"""
