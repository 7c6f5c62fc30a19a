"""Hot-path configuration keys with the reference's defaults (utils/config.py:20-133).

yacs is not required: `CfgNode` here is a small attribute-dict with the methods the reference's
call sites use (clone, merge_from_file, merge_from_list, freeze/defrost).  The modules in
`modelling/` only ever do `cfg.X.Y` attribute access, so a real yacs node, this class, or any
attribute bag (test/test_model2018.py:23-48) all work.
"""
import ast
import copy


class CfgNode(dict):
    def __init__(self, init=None):
        super().__init__()
        for k, v in (init or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) and not isinstance(v, CfgNode) else v

    def __getattr__(self, k):
        if k in self:
            return self[k]
        raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v

    def clone(self):
        return copy.deepcopy(self)

    def _merge(self, other):
        for k, v in other.items():
            if isinstance(v, dict) and isinstance(self.get(k), dict):
                self[k]._merge(v)
            else:
                self[k] = CfgNode(v) if isinstance(v, dict) else copy.deepcopy(v)

    def merge_from_file(self, path):
        import yaml
        with open(path) as f:
            self._merge(yaml.safe_load(f) or {})

    def merge_from_list(self, kv):
        for key, val in zip(kv[0::2], kv[1::2]):
            node, parts = self, key.split(".")
            for p in parts[:-1]:
                node = node[p]
            if isinstance(val, str):
                try:
                    val = ast.literal_eval(val)
                except (ValueError, SyntaxError):
                    pass
            node[parts[-1]] = val

    def freeze(self):
        pass

    def defrost(self):
        pass


def get_cfg_defaults():
    """Defaults of every key the hot path reads (same values as the reference's `_C`)."""
    return CfgNode({
        "SEED": 0,
        "DATA": {"IN_CHANNELS": 3, "SIZE": 256, "MEAN": [0., 0., 0.], "STD": [1., 1., 1.], "NUM_WORKERS": 4},
        "MODEL": {
            "META_ARCHITECTURE": "Compressor2018",
            "STRIDES": [2, 2, 2, 2], "CONV_KERNEL": 5, "INTER_CHANNELS": 192, "LATENT_CHANNELS": 192,
            "HYPER_PRIOR": {"STRIDES": [1, 2, 2], "KERNELS": [3, 5, 5]},
            "ENTROPY_MODEL": {"DIMS": [3, 3, 3], "INIT_SCALE": 10, "BIN": 1.0, "PROB_EPS": 1e-10,
                              "CONDITIONAL_MODEL": "LaplacianConditionalModel"},
            "LOSS": {"DISTORTION_LOSS_NAMES": ["MSE"], "DISTORTION_LOSS_WEIGHT": 1.0, "REDUCTION": "none",
                     "SSIM": {"MAX_VAL": 255.0, "FILTER_SIZE": 11, "FILTER_SIGMA": 1.5, "K1": 0.01, "K2": 0.03,
                              "LOG_SCALE": False, "EPS": 1e-5},
                     "MS_SSIM_WEIGHTS": [0.0448, 0.2856, 0.3001, 0.2363, 0.1333]},
        },
        # utils/config.py:86-127 of the reference (the keys solver/optim.py and engine/trainer.py read)
        "SOLVER": {"GD_STEPS": 1, "IMS_PER_BATCH": 2, "GRAD_CLIP": 0.0, "BASE_LR": 0.001, "OPT_NAME": "sgd",
                   "EPS": 1e-4, "WEIGHT_DECAY": 0.0005, "WEIGHT_DECAY_BIAS": 0.0, "BIAS_LR_FACTOR": 1.0,
                   "MOMENTUM": 0.9, "USE_NESTEROV": False},
        "VAL": {"BATCH_SIZE": 1},
    })


# the overrides the shipped recipes apply on top of the defaults (configs/*.yaml)
RECIPES = {
    **{f"psnr_{l}": {"MODEL": {"LOSS": {"DISTORTION_LOSS_WEIGHT": float(l), "REDUCTION": "mean"}},
                     "DATA": {"SIZE": 192}, "SOLVER": {"IMS_PER_BATCH": 64, "GRAD_CLIP": 5.0, "BASE_LR": 1e-4, "OPT_NAME": "adamw"}}
       for l in (256, 512, 1024, 2048, 4096, 8192)},
    **{f"ssim_{l}": {"MODEL": {"LOSS": {"DISTORTION_LOSS_NAMES": ["MS_SSIMLoss"], "SSIM": {"LOG_SCALE": True},
                                        "DISTORTION_LOSS_WEIGHT": float(l), "REDUCTION": "mean"}},
                     "DATA": {"SIZE": 192}, "SOLVER": {"IMS_PER_BATCH": 64, "GRAD_CLIP": 5.0, "BASE_LR": 1e-4, "OPT_NAME": "adamw"}}
       for l in (8, 16, 32, 64, 128, 256, 512)},
}


def get_recipe_cfg(name):
    """cfg for one of the reference's shipped recipes (configs/psnr_*.yaml, configs/ssim_*.yaml)."""
    cfg = get_cfg_defaults()
    cfg._merge(RECIPES[name])
    return cfg
