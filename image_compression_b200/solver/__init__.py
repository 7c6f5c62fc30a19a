"""Drop-in for the reference's solver/optim.py on the hot path's device."""
from .optim import FusedAdamW, make_optimizer

__all__ = ["FusedAdamW", "make_optimizer"]
