"""(reference: src/llm_training/data/base_datamodule_config.py:4-13)"""
from pydantic import BaseModel, ConfigDict


class BaseDataModuleConfig(BaseModel):
    pre_processed_data_path: str | None = None
    validation_split: int | float | None = None
    batch_size: int = 1
    num_workers: int = 0
    pin_memory: bool = True        # batches are staged in pinned memory for the async H2D copy
    prepare_data_per_node: bool = False
    prefetch_factor: int | None = None

    model_config = ConfigDict(arbitrary_types_allowed=True)
