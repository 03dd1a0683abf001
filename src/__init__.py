"""Drop-in for the reference's `src` package (src/__init__.py:3-13): the names train.py and the experiment YAMLs import,
served by the B200 implementation. Put this directory (and phantom_vlb_b200/) on PYTHONPATH ahead of the reference's own
`src/` and `python train.py experiment=VLB_vllama2_friends_lora subject=sub-01` resolves every `_target_` here."""
from phantom_vlb_b200.utils import HRFConvolveLayer, LogValAccuracyCallback, RidgeRegressionLayer, get_hrf_weight

__all__ = ["HRFConvolveLayer", "RidgeRegressionLayer", "get_hrf_weight", "LogValAccuracyCallback"]
