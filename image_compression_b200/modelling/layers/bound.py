"""LowerBound / UpperBound with the reference's call form `X.apply(tensor, bound)`
(modelling/layers/bound.py:28-59); kernels: hpc_bound_fwd / hpc_bound_bwd."""
from ...autograd_ops import LowerBound, UpperBound  # noqa: F401
