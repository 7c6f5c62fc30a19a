from .config import CfgNode, get_cfg_defaults, get_recipe_cfg
from .registry import Registry
from .metric import psnr, SSIM, MS_SSIM, image_metrics
