"""Import the UNMODIFIED reference (hieu1999210/image_compression) in this container.

TEST INFRASTRUCTURE ONLY.  Used by tests/golden/make_golden.py (run in the build
container, where /root/reference is mounted) to pin oracle/hyperprior_oracle.py against
the reference's own modules and to generate the committed golden vectors.  Nothing on the
GPU box may import this: /root/reference does not exist there.

The reference imports `yacs` (utils/config.py:16), which is not installed and cannot be
(no network).  A ~30 line stand-in for `yacs.config.CfgNode` (attribute dict + clone +
merge_from_file/merge_from_list) is injected into sys.modules; no reference source is
copied or modified.
"""
import copy
import sys
import types

REFERENCE_ROOT = "/root/reference"


class CfgNode(dict):
    """Minimal yacs.config.CfgNode stand-in (attribute access over a dict)."""

    def __init__(self, init=None):
        super().__init__()
        for k, v in (init or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) and not isinstance(v, CfgNode) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def clone(self):
        return copy.deepcopy(self)

    def merge_from_other(self, other):
        for k, v in other.items():
            if isinstance(v, dict) and k in self and isinstance(self[k], dict):
                self[k].merge_from_other(CfgNode(v))
            else:
                self[k] = copy.deepcopy(v)

    def merge_from_file(self, path):
        import yaml
        with open(path) as f:
            self.merge_from_other(CfgNode(yaml.safe_load(f) or {}))

    def merge_from_list(self, lst):
        import ast
        for key, val in zip(lst[0::2], lst[1::2]):
            node = self
            parts = key.split(".")
            for p in parts[:-1]:
                node = node[p]
            try:
                val = ast.literal_eval(val) if isinstance(val, str) else val
            except (ValueError, SyntaxError):
                pass
            node[parts[-1]] = val

    def freeze(self):
        pass

    def defrost(self):
        pass


def install():
    """Make `import modelling`, `from utils import psnr, MS_SSIM` resolve to the reference."""
    if "yacs" not in sys.modules:
        yacs = types.ModuleType("yacs")
        yacs_config = types.ModuleType("yacs.config")
        yacs_config.CfgNode = CfgNode
        yacs.config = yacs_config
        sys.modules["yacs"] = yacs
        sys.modules["yacs.config"] = yacs_config
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


def reference_cfg(yaml_name=None, overrides=()):
    """cfg = reference defaults (utils/config.py:20-133) <- configs/<yaml_name> <- overrides."""
    install()
    from utils.config import get_cfg_defaults  # reference module
    cfg = get_cfg_defaults()
    if yaml_name:
        cfg.merge_from_file(f"{REFERENCE_ROOT}/configs/{yaml_name}")
    if overrides:
        cfg.merge_from_list(list(overrides))
    return cfg
