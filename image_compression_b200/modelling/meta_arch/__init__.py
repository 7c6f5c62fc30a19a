from .build import build_model, META_ARCH_REGISTRY
from .bmshl2018 import Compressor2018
