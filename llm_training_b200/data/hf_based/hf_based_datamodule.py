"""Datamodules backed by Hugging Face ``datasets`` (reference:
src/llm_training/data/hf_based/hf_based_datamodule.py:26-248): ``load_dataset(**dataset_kwargs)``,
``validation_split`` with seed 42, cache on/off, a cache fingerprint that is stable across tokenizer object
identity (tokenizers are hashed by the bytes of their saved files), save/load of pre-processed data."""
import hashlib
import logging
import tempfile
from contextlib import contextmanager
from pathlib import Path
from typing import Any, Callable

from ..base_datamodule import BaseDataModule, DatasetDict
from .hf_based_datamodule_config import HFBasedDataModuleConfig

logger = logging.getLogger(__name__)


@contextmanager
def cache_context(enabled: bool):
    import datasets
    previous = datasets.is_caching_enabled()
    (datasets.enable_caching if enabled else datasets.disable_caching)()
    try:
        yield
    finally:
        (datasets.enable_caching if previous else datasets.disable_caching)()


class HFBasedDataModule(BaseDataModule):
    config: HFBasedDataModuleConfig

    def __init__(self, config: HFBasedDataModuleConfig) -> None:
        super().__init__(config)

    def load_data(self) -> DatasetDict:
        import datasets
        assert self.config.dataset_kwargs is not None, '`dataset_kwargs` is required'
        kwargs = dict(self.config.dataset_kwargs)
        kwargs.setdefault('num_proc', self.config.num_proc)
        data = datasets.load_dataset(**kwargs)
        if isinstance(data, datasets.Dataset):
            data = datasets.DatasetDict({'train': data})
        assert self.config.validation_split is None or ('train' in data and 'validation' not in data)
        if self.config.cleanup_cache_files:
            logger.info(f'Cleanup cache files: {data.cleanup_cache_files()}')
        return data

    def split_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        if self.config.validation_split is not None:
            dataset_dict = dataset_dict['train'].train_test_split(self.config.validation_split, seed=42)
            dataset_dict['validation'] = dataset_dict.pop('test')
        return dataset_dict

    def prepare_data(self) -> None:
        if self.config.pre_processed_data_path is None:
            with cache_context(self.config.enable_cache):
                self.pre_process_data(self.load_data())

    def setup(self, stage: str | None = None) -> None:
        with cache_context(self.config.enable_cache):
            super().setup(stage)

    def save_pre_processed_data(self, path: str | None = None) -> None:
        self.pre_processed_dataset_dict.save_to_disk(path or self.config.pre_processed_data_path,
                                                     num_proc=self.config.num_proc)

    def load_pre_processed_data(self, path: str | None = None) -> None:
        import datasets
        self.pre_processed_dataset_dict = datasets.load_from_disk(path or self.config.pre_processed_data_path)

    def cleanup_cache_files(self) -> None:
        if hasattr(self.raw_dataset_dict, 'cleanup_cache_files'):
            self.raw_dataset_dict.cleanup_cache_files()

    # -- stable fingerprints -----------------------------------------------------------------------------
    @classmethod
    def hash_tokenizer(cls, tokenizer: Any) -> str:
        digest = hashlib.sha256()
        with tempfile.TemporaryDirectory() as tmp:
            tokenizer.save_pretrained(tmp)
            for path in sorted(Path(tmp).glob('*')):
                digest.update(path.name.encode())
                digest.update(path.read_bytes())
        return digest.hexdigest()

    @classmethod
    def hash_fn_kwargs(cls, fn_kwargs: dict[str, Any]) -> str:
        from datasets.fingerprint import Hasher
        stable = {k: (cls.hash_tokenizer(v) if hasattr(v, 'save_pretrained') else v) for k, v in fn_kwargs.items()}
        return Hasher.hash(stable)

    @classmethod
    def map_dataset(cls, dataset, function: Callable, fn_kwargs: dict | None = None, new_fingerprint: str | None = None,
                    **map_kwargs: Any):
        if new_fingerprint is None:
            from datasets.fingerprint import Hasher
            described = {k: v for k, v in map_kwargs.items() if k not in ('num_proc', 'desc', 'load_from_cache_file')}
            described['features'] = str(described.get('features'))
            new_fingerprint = Hasher.hash([dataset._fingerprint, getattr(function, '__qualname__', repr(function)),
                                           cls.hash_fn_kwargs(fn_kwargs or {}), described])
        return dataset.map(function=function, fn_kwargs=fn_kwargs, new_fingerprint=new_fingerprint, **map_kwargs)

    @classmethod
    def map_dataset_dict(cls, dataset_dict, function: Callable, remove_columns: Any = None, **kwargs: Any):
        import datasets
        return datasets.DatasetDict({
            split: cls.map_dataset(ds, function,
                                   remove_columns=ds.column_names if remove_columns is True else remove_columns,
                                   **kwargs)
            for split, ds in dataset_dict.items()})
