"""(reference: src/llm_training/data/hf_based/hf_based_datamodule_config.py:4-8)"""
from ..base_datamodule_config import BaseDataModuleConfig


class HFBasedDataModuleConfig(BaseDataModuleConfig):
    dataset_kwargs: dict | None = None
    num_proc: int | None = None
    cleanup_cache_files: bool = False
    enable_cache: bool = True
