from .ctgan import (  # noqa: F401
    Generator, Residual, Discriminator, apply_activate, cond_loss, slerp, calc_gradient_penalty,
)
from .arena import ParamArena, GeneratorArena, DiscriminatorArena, round_up  # noqa: F401
