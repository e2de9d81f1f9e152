"""Pipeline ``load_data -> pre_process_data -> (save/load pre-processed) -> split -> post_process`` and the
dataloaders (reference: src/llm_training/data/base_datamodule.py:18-132).  The train loader is resumable and
shuffled; under data parallelism it uses a ``DistributedSampler`` over the dp axis only, so every rank of a
tensor-parallel group sees the same batch (reference: fsdp2_strategy.py:149-153)."""
import logging
import os
import shutil
from typing import Any, Mapping, TextIO

import torch
from torch.utils.data import DataLoader, Dataset, DistributedSampler, RandomSampler, SequentialSampler

from ..lightning.module import LightningDataModule
from .base_datacollator import BaseDataCollator
from .base_datamodule_config import BaseDataModuleConfig
from .resumable_dataloader import ResumableDataLoader

logger = logging.getLogger(__name__)

DatasetDict = Mapping[str, Dataset]


class BaseDataModule(LightningDataModule):
    datacollator_class: type[BaseDataCollator] | None = None

    def __init__(self, config: BaseDataModuleConfig) -> None:
        super().__init__()
        self.config = config
        self.datacollator = self.datacollator_class(config) if self.datacollator_class is not None else None
        self.prepare_data_per_node = config.prepare_data_per_node
        self.raw_dataset_dict: DatasetDict | None = None
        self.pre_processed_dataset_dict: DatasetDict | None = None
        self.dataset_dict: DatasetDict | None = None

    # -- pipeline stages -------------------------------------------------------------------------------
    def load_data(self) -> DatasetDict:
        raise NotImplementedError

    def pre_process_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        return dataset_dict

    def split_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        return dataset_dict

    def post_process_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        return self.split_data(dataset_dict)

    def prepare_data(self) -> None:
        if self.config.pre_processed_data_path is None:
            self.pre_process_data(self.load_data())     # warms the on-disk caches once per node

    def save_pre_processed_data(self, path: str | None = None) -> None:
        raise NotImplementedError

    def load_pre_processed_data(self, path: str | None = None) -> None:
        raise NotImplementedError

    def print_dataset_info(self, file: TextIO | None = None) -> None:
        width = shutil.get_terminal_size().columns
        for title, value in (('Raw Dataset', self.raw_dataset_dict),
                             ('Pre-processed Dataset', self.pre_processed_dataset_dict),
                             ('Final Dataset', self.dataset_dict)):
            print(f' {title} '.center(width, '─'), end='\n\n', file=file)
            print(value, end='\n\n', file=file)

    def setup(self, stage: str | None = None) -> None:
        if self.config.pre_processed_data_path is None:
            self.raw_dataset_dict = self.load_data()
            self.pre_processed_dataset_dict = self.pre_process_data(self.raw_dataset_dict)
        else:
            logger.info('Load pre-processed data')
            self.load_pre_processed_data(self.config.pre_processed_data_path)
        self.dataset_dict = self.post_process_data(self.pre_processed_dataset_dict)

    # -- dataloaders -----------------------------------------------------------------------------------
    def _sampler(self, dataset: Dataset, shuffle: bool):
        strategy = None if self.trainer is None else self.trainer.strategy
        seed = int(os.environ.get('PL_GLOBAL_SEED', 0))
        if strategy is not None and strategy.dp_size > 1:
            return DistributedSampler(dataset, shuffle=shuffle, seed=seed, drop_last=False,
                                      **strategy.distributed_sampler_kwargs)
        if shuffle:
            return _EpochSeededSampler(dataset, seed)
        return SequentialSampler(dataset)

    def _get_dataloader(self, split: str) -> DataLoader | None:
        if self.dataset_dict is None or split not in self.dataset_dict:
            return None
        dataset = self.dataset_dict[split]
        train = split == 'train'
        kwargs: dict[str, Any] = dict(
            dataset=dataset, batch_size=self.config.batch_size, num_workers=self.config.num_workers,
            collate_fn=self.datacollator, pin_memory=self.config.pin_memory and torch.cuda.is_available(),
            sampler=self._sampler(dataset, shuffle=train))
        if self.config.num_workers > 0 and self.config.prefetch_factor is not None:
            kwargs['prefetch_factor'] = self.config.prefetch_factor
        if train:
            return ResumableDataLoader(self.trainer, **kwargs)
        return DataLoader(**kwargs)

    def train_dataloader(self) -> DataLoader | None:
        return self._get_dataloader('train')

    def val_dataloader(self) -> DataLoader | None:
        return self._get_dataloader('validation')

    def test_dataloader(self) -> DataLoader | None:
        return self._get_dataloader('test')

    def predict_dataloader(self) -> DataLoader | None:
        return self._get_dataloader('predict')


class _EpochSeededSampler(RandomSampler):
    """Shuffle that is a pure function of (seed, epoch): a resumed run re-creates the same order, which
    is what the skip-ahead batch sampler relies on."""

    def __init__(self, data_source, seed: int) -> None:
        super().__init__(data_source)
        self.seed = seed
        self.epoch = 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __iter__(self):
        g = torch.Generator()
        g.manual_seed(self.seed + self.epoch)
        yield from torch.randperm(len(self.data_source), generator=g).tolist()
