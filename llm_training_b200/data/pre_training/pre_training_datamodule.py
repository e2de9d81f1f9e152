"""Pre-training data: tokenize (+BOS/EOS) -> sliding-window truncate -> naive packing | best-fit-decreasing
bin packing (per source, per map batch) -> per-source sampling (reference:
src/llm_training/data/pre_training/pre_training_datamodule.py:23-360)."""
import logging
import math
import random
from typing import Any

from ..base_datamodule import DatasetDict
from ..hf_based import HFBasedDataModule
from ..packing import best_fit_bin_packing
from .pre_training_datacollator import PreTrainingDataCollator
from .pre_training_datamodule_config import PackingMethod, PreTrainingDataModuleConfig

logger = logging.getLogger(__name__)

Batch = dict[str, list]


def tokenize(batch: Batch, tokenizer: Any) -> Batch:
    keep = [i for i, text in enumerate(batch['text']) if text]
    texts = [batch['text'][i] for i in keep]
    sources = [batch['source'][i] for i in keep]
    ids = tokenizer(texts, add_special_tokens=False, return_token_type_ids=False, return_attention_mask=False,
                    verbose=False)['input_ids'] if texts else []
    out: Batch = {'source': sources, 'input_ids': [], 'length': []}
    for row in ids:
        row = [tokenizer.bos_token_id] + list(row) + [tokenizer.eos_token_id]
        out['input_ids'].append(row)
        out['length'].append(len(row))
    return out


def truncate(batch: Batch, max_length: int, stride: int) -> Batch:
    out: Batch = {'source': [], 'input_ids': [], 'length': []}
    for source, ids in zip(batch['source'], batch['input_ids']):
        for start in range(0, len(ids), stride):
            window = ids[start:start + max_length]
            out['source'].append(source)
            out['input_ids'].append(window)
            out['length'].append(len(window))
    return out


def naive_packing(batch: Batch, max_length: int) -> Batch:
    """Concatenate consecutive documents of the same source and cut every ``max_length`` tokens; segment ids
    restart at 1 in every emitted row."""
    out: Batch = {'source': [], 'input_ids': [], 'attention_mask': [], 'length': []}

    def emit(source, ids: list[int], seg: list[int]) -> None:
        base = seg[0] - 1
        out['source'].append(source)
        out['input_ids'].append(ids)
        out['attention_mask'].append([s - base for s in seg] if base else seg)
        out['length'].append(len(ids))

    if not batch['input_ids']:
        return out
    current_source = batch['source'][0]
    ids_buf: list[int] = []
    seg_buf: list[int] = []
    for source, ids in zip(batch['source'], batch['input_ids']):
        if len(ids) == max_length:
            emit(source, list(ids), [1] * max_length)
            continue
        if source != current_source:
            if ids_buf:
                emit(current_source, ids_buf, seg_buf)
                ids_buf, seg_buf = [], []
            current_source = source
        seg_id = seg_buf[-1] + 1 if seg_buf else 1
        ids_buf = ids_buf + list(ids)
        seg_buf = seg_buf + [seg_id] * len(ids)
        while len(ids_buf) >= max_length:
            emit(current_source, ids_buf[:max_length], seg_buf[:max_length])
            ids_buf, seg_buf = ids_buf[max_length:], seg_buf[max_length:]
    if ids_buf:
        emit(current_source, ids_buf, seg_buf)
    return out


def best_fit_decreasing(batch: Batch, max_length: int) -> Batch:
    out: Batch = {'input_ids': [], 'attention_mask': [], 'source': [], 'length': []}
    by_source: dict[Any, list[int]] = {}
    for i, source in enumerate(batch['source']):
        by_source.setdefault(source, []).append(i)
    for source, members in by_source.items():
        members = sorted(members, key=lambda i: batch['length'][i], reverse=True)   # stable, like the reference
        lengths = [batch['length'][i] for i in members]
        for group in best_fit_bin_packing(max_length, lengths):
            ids: list[int] = []
            seg: list[int] = []
            for doc, k in enumerate(group, start=1):
                row = batch['input_ids'][members[k]]
                ids += row
                seg += [doc] * len(row)
            out['input_ids'].append(ids)
            out['attention_mask'].append(seg)
            out['source'].append(source)
            out['length'].append(len(ids))
    return out


def pre_process_batch(batch: Batch, tokenizer: Any, max_length: int | None, stride: int | None,
                      packing_method: PackingMethod | str) -> Batch:
    batch = tokenize(batch, tokenizer)
    if max_length is not None:
        batch = truncate(batch, max_length, stride or max_length)
    method = PackingMethod(packing_method)
    if method == PackingMethod.NAIVE_PACKING:
        return naive_packing(batch, max_length)
    if method == PackingMethod.BEST_FIT_BIN_PACKING:
        return best_fit_decreasing(batch, max_length)
    batch['attention_mask'] = [[1] * n for n in batch['length']]
    return batch


class PreTrainingDataModule(HFBasedDataModule):
    config: PreTrainingDataModuleConfig
    datacollator_class = PreTrainingDataCollator

    def __init__(self, config: PreTrainingDataModuleConfig) -> None:
        super().__init__(config)

    # static aliases under the reference's method names
    _tokenize = staticmethod(tokenize)
    _truncate = staticmethod(truncate)
    _naive_packing = staticmethod(naive_packing)
    _best_fit_decreasing = staticmethod(best_fit_decreasing)
    _best_fit_bin_packing = staticmethod(best_fit_bin_packing)
    _pre_process_data = staticmethod(pre_process_batch)

    def pre_process_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        from datasets import Features, Sequence, Value
        logger.info('Sorting data by source')
        for split, ds in list(dataset_dict.items()):
            if 'source' in ds.column_names:
                dataset_dict[split] = ds.sort('source')
            else:
                dataset_dict[split] = ds.add_column('source', [None] * len(ds))
        return self.map_dataset_dict(
            dataset_dict, pre_process_batch,
            fn_kwargs=dict(tokenizer=self.config.tokenizer, max_length=self.config.max_length,
                           stride=self.config.stride, packing_method=str(self.config.packing_method)),
            batched=True, batch_size=self.config.pre_processing_batch_size, remove_columns=True,
            num_proc=self.config.num_proc,
            features=Features({'source': Value('string'), 'input_ids': Sequence(Value('int32')),
                               'attention_mask': Sequence(Value('uint16')), 'length': Value('uint32')}),
            desc='Pre-processing data')

    def sample_data(self, dataset):
        """``sample_rate[source]``: the integer part repeats the source, the fraction subsamples it (seed 42)."""
        rates = self.config.sample_rate
        if all(r == 1.0 for r in rates.values()):
            return dataset
        by_source: dict[Any, list[int]] = {}
        for i, source in enumerate(dataset['source']):
            by_source.setdefault(source, []).append(i)
        rng = random.Random(42)
        unused = dict(rates)
        picked: list[int] = []
        for source, indices in by_source.items():
            rate = rates.get(source, 1.0)
            unused.pop(source, None)
            frac, whole = math.modf(rate)
            picked += indices * int(whole)
            if frac > 0.0:
                picked += rng.sample(indices, k=int(len(indices) * frac))
        if unused:
            logger.warning(f'Some sources specified by `sample_rate` are not found in the dataset:\n {unused}')
        return dataset.select(picked)

    def post_process_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        dataset_dict = super().post_process_data(dataset_dict)
        if 'train' in dataset_dict:
            dataset_dict['train'] = self.sample_data(dataset_dict['train'])
        return dataset_dict

    def print_dataset_info(self, file: Any = None) -> None:
        """Dataset summaries plus the token counts before and after per-source sampling (reference :346-360)."""
        import shutil
        super().print_dataset_info(file)
        print('─' * shutil.get_terminal_size().columns, end='\n\n', file=file)
        print('Original Tokens:', end='\n\n', file=file)
        print(self.token_table(self.pre_processed_dataset_dict), end='\n\n', file=file)
        print('Sampled Tokens:', end='\n\n', file=file)
        print(self.token_table(self.dataset_dict), file=file)

    def token_table(self, dataset_dict: Any = None) -> str:
        """Per-split / per-source document and token counts (reference :312-360)."""
        from tabulate import tabulate
        rows = []
        for split, ds in ((dataset_dict if dataset_dict is not None else self.dataset_dict) or {}).items():
            counts: dict[Any, list[int]] = {}
            for source, n in zip(ds['source'], ds['length']):
                c = counts.setdefault(source, [0, 0])
                c[0] += 1
                c[1] += int(n)
            for source, (docs, tokens) in sorted(counts.items(), key=lambda kv: str(kv[0])):
                rows.append([split, source, docs, tokens])
        return tabulate(rows, headers=['split', 'source', 'samples', 'tokens'], tablefmt='github')
