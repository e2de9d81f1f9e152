"""Instruction tuning data: chat template -> tokens + assistant-only labels -> overlong DROP / TRUNCATE ->
optional GROUP_BY_LENGTH packing with per-document segment ids (contamination-free attention) — reference:
src/llm_training/data/instruction_tuning/instruction_tuning_datamodule.py:17-202."""
import random
from typing import Any

from ..base_datamodule import DatasetDict
from ..chat_utils import encode_conversations
from ..hf_based import HFBasedDataModule
from ..packing import group_indices_by_length
from .instruction_tuning_datacollator import InstructionTuningDataCollator
from .instruction_tuning_datamodule_config import (InstructionTuningDataModuleConfig, OverlongHandlingMethod,
                                                   PackingMethod)

Batch = dict[str, list]


def apply_chat_template_and_tokenize(batch: Batch, tokenizer: Any, chat_template: str | None,
                                     default_system_prompt: str | None,
                                     add_default_system_prompt_rate: float | None) -> Batch:
    conversations = []
    for messages in batch['messages']:
        messages = list(messages)
        if (default_system_prompt is not None and add_default_system_prompt_rate is not None
                and not any(m['role'] == 'system' for m in messages)
                and random.random() < add_default_system_prompt_rate):
            messages.insert(0, {'role': 'system', 'content': default_system_prompt})
        conversations.append(messages)
    ids, labels = encode_conversations(tokenizer, conversations, chat_template)
    return {'input_ids': ids, 'attention_mask': [[1] * len(r) for r in ids], 'labels': labels,
            'length': [len(r) for r in ids]}


def drop_overlong_examples(batch: Batch, max_length: int) -> Batch:
    keep = [i for i, n in enumerate(batch['length']) if n <= max_length]
    return {k: [v[i] for i in keep] for k, v in batch.items()}


def truncate_overlong_examples(batch: Batch, max_length: int) -> Batch:
    """Cuts the EXAMPLE (the reference slices the batch list by mistake)."""
    for i, n in enumerate(batch['length']):
        if n > max_length:
            for key in ('input_ids', 'labels', 'attention_mask'):
                batch[key][i] = batch[key][i][:max_length]
            batch['length'][i] = max_length
    return batch


def group_by_length(batch: Batch, max_length: int) -> Batch:
    out: Batch = {'input_ids': [], 'attention_mask': [], 'labels': [], 'length': []}
    for group in group_indices_by_length(batch['length'], max_length):
        ids: list[int] = []
        seg: list[int] = []
        labels: list[int] = []
        for doc, i in enumerate(group, start=1):
            ids += batch['input_ids'][i]
            seg += [doc] * batch['length'][i]
            labels += batch['labels'][i]
        out['input_ids'].append(ids)
        out['attention_mask'].append(seg)
        out['labels'].append(labels)
        out['length'].append(len(ids))
    return out


def pre_process_batch(batch: Batch, tokenizer: Any, chat_template: str | None, default_system_prompt: str | None,
                      add_default_system_prompt_rate: float | None, max_length: int | None,
                      overlong_handling_method: str, packing_method: str) -> Batch:
    batch = apply_chat_template_and_tokenize(batch, tokenizer, chat_template, default_system_prompt,
                                             add_default_system_prompt_rate)
    if max_length is not None:
        if OverlongHandlingMethod(overlong_handling_method) == OverlongHandlingMethod.DROP:
            batch = drop_overlong_examples(batch, max_length)
        else:
            batch = truncate_overlong_examples(batch, max_length)
        if PackingMethod(packing_method) == PackingMethod.GROUP_BY_LENGTH:
            batch = group_by_length(batch, max_length)
    return batch


class InstructionTuningDataModule(HFBasedDataModule):
    config: InstructionTuningDataModuleConfig
    datacollator_class = InstructionTuningDataCollator

    _apply_chat_template_and_tokenize = staticmethod(apply_chat_template_and_tokenize)
    _drop_overlong_examples = staticmethod(drop_overlong_examples)
    _truncate_overlong_examples = staticmethod(truncate_overlong_examples)
    _group_indices_by_length = staticmethod(group_indices_by_length)
    _group_by_length = staticmethod(group_by_length)
    _pre_process_data = staticmethod(pre_process_batch)

    def __init__(self, config: InstructionTuningDataModuleConfig) -> None:
        super().__init__(config)

    def pre_process_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        from datasets import Features, Sequence, Value
        c = self.config
        return self.map_dataset_dict(
            dataset_dict, pre_process_batch,
            fn_kwargs=dict(tokenizer=c.tokenizer, chat_template=c.chat_template,
                           default_system_prompt=c.default_system_prompt,
                           add_default_system_prompt_rate=c.add_default_system_prompt_rate, max_length=c.max_length,
                           overlong_handling_method=str(c.overlong_handling_method),
                           packing_method=str(c.packing_method)),
            batched=True, remove_columns=True, num_proc=c.num_proc,
            features=Features({'input_ids': Sequence(Value('int32')), 'attention_mask': Sequence(Value('uint16')),
                               'labels': Sequence(Value('int32')), 'length': Value('uint32')}),
            desc='Pre-processing data')
