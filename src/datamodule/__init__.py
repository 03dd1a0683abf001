"""`src.datamodule.VLBDataModule` / `VLBDataModuleConfig` (reference: src/datamodule/__init__.py:3-10), the Hydra `_target_`s
of config/experiment/VLB_vllama2_friends_{lora,baseline}.yaml:3-7."""
from phantom_vlb_b200.datamodule import VLB_Dataset, VLBDataModule, VLBDataModuleConfig, VLBDatasets

__all__ = ["VLBDataModule", "VLBDataModuleConfig", "VLB_Dataset", "VLBDatasets"]
