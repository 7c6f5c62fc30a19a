"""Unused by the hot path; kept so `from modelling.layers import ShapeSpec` resolves
(modelling/layers/shape_spec.py:23-37)."""
from collections import namedtuple


class ShapeSpec(namedtuple("_ShapeSpec", ["channels", "height", "width", "stride"])):
    def __new__(cls, *, channels=None, height=None, width=None, stride=None):
        return super().__new__(cls, channels, height, width, stride)
