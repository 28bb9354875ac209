"""Mirror of ``markovflow/models`` for the diffusion-process models."""
from .variational_cvi_sde import CVISitesSDE  # noqa: F401
from .vi_sde import VariationalMarkovGP  # noqa: F401
