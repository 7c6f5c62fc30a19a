from .meta_arch import build_model
from .loss import SSIMLoss, MS_SSIMLoss
