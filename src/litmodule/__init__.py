"""`src.litmodule.VLBLitModule` / `VLBLitModuleConfig` (reference: src/litmodule/__init__.py:3-8), the Hydra `_target_`s of
config/experiment/VLB_vllama2_friends_{lora,baseline}.yaml:19-23."""
from phantom_vlb_b200.litmodule import VLBLitModule, VLBLitModuleConfig

__all__ = ["VLBLitModuleConfig", "VLBLitModule"]
