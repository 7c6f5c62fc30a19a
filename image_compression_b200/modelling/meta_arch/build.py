"""build_model(cfg) -> nn.Module, via the META_ARCH registry (modelling/meta_arch/build.py:21-36)."""
from ...utils.registry import Registry

META_ARCH_REGISTRY = Registry("META_ARCH")


def build_model(cfg):
    return META_ARCH_REGISTRY.get(cfg.MODEL.META_ARCHITECTURE)(cfg)
