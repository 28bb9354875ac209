"""Mirror of ``markovflow/sde``."""
from .drift import LinearDrift  # noqa: F401
from .sde import (SDE, BenesSDE, DoubleWellSDE, Lorenz63SDE, MLPDrift, OrnsteinUhlenbeckSDE, OUNdSDE,  # noqa: F401
                  SineDiffusionSDE, SqrtDiffusionSDE, VanderPolOscillatorSDE)
