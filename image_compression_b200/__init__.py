"""image_compression_b200: B200-native (sm_100a) hot path of the scale-hyperprior codec of
hieu1999210/image_compression behind that repository's own module surface.

    from image_compression_b200 import modelling, utils
    model = modelling.build_model(cfg).cuda()          # same cfg keys / state_dict as the reference
    x_tilde, losses = model(images)                    # images: (N,3,H,W) in [0,1] on cuda
"""
from . import _lib
from .hotpath import compute_mode, get_compute_mode, set_compute_mode
from . import modelling, solver, utils
from .modelling import build_model
from .utils import get_cfg_defaults, get_recipe_cfg

__all__ = ["modelling", "solver", "utils", "build_model", "get_cfg_defaults", "get_recipe_cfg", "set_compute_mode",
           "get_compute_mode", "compute_mode"]
