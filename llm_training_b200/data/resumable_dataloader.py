"""DataLoader whose batch sampler fast-forwards past the batches already consumed in the current epoch
(index-only, no data is read) — reference: src/llm_training/data/resumable_dataloader.py:8-56."""
from typing import Any, Iterable, Iterator

from torch.utils.data import BatchSampler, DataLoader
from torch.utils.data.sampler import Sampler


class ResumableBatchSampler(BatchSampler):
    def __init__(self, sampler: Sampler[int] | Iterable[int], batch_size: int, drop_last: bool,
                 trainer: Any | None = None) -> None:
        super().__init__(sampler, batch_size, drop_last)
        self.trainer = trainer

    def __iter__(self) -> Iterator[list[int]]:
        skip = 0 if self.trainer is None else int(self.trainer.fit_loop.batch_idx)
        for i, indices in enumerate(super().__iter__()):
            if i < skip:
                continue
            yield indices


class ResumableDataLoader(DataLoader):
    def __init__(self, trainer: Any | None = None, *args: Any, **kwargs: Any) -> None:
        self.trainer = trainer
        super().__init__(*args, **kwargs)

    def __setattr__(self, attr: str, val: Any) -> None:
        if attr == 'batch_sampler' and isinstance(val, BatchSampler) and not isinstance(val, ResumableBatchSampler):
            val = ResumableBatchSampler(val.sampler, val.batch_size, val.drop_last, self.trainer)
        super().__setattr__(attr, val)

    def state_dict(self) -> dict[str, Any]:
        return {}

    def load_state_dict(self, state_dict: dict[str, Any]) -> None: ...
