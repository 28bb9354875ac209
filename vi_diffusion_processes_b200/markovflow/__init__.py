"""
Host-side mirror of the reference's Python interface for the VDP / CVI-DP hot path: the same module,
class and function names as ``markovflow`` (AaltoML/vi-diffusion-processes), holding float64 CUDA
``torch.Tensor``s instead of ``tf.Tensor``s and routing every grid-sized operation to libvdp_b200.so.

    markovflow.block_tri_diag                    SymmetricBlockTriDiagonal, LowerTriangularBlockTriDiagonal
    markovflow.state_space_model                 StateSpaceModel
    markovflow.gauss_markov                      BTDGaussian
    markovflow.ssm_gaussian_transformations      ssm_to_expectations, expectations_to_ssm_params,
                                                 ssm_to_naturals, naturals_to_ssm_params
    markovflow.ssm_natgrad                       SSMNaturalGradient
    markovflow.sde                               SDE and its drifts, LinearDrift, sde_utils
    markovflow.likelihoods                       MultivariateGaussian
    markovflow.models.variational_cvi_sde        CVISitesSDE
    markovflow.models.vi_sde                     VariationalMarkovGP

There is no CPU path: every call needs a CUDA device and the built library.
"""
from . import _util  # noqa: F401
