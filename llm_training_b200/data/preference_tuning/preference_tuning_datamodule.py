"""``{prompt, chosen, rejected}`` -> two chat-templated sequences with assistant-only labels; an example is
dropped when either side is too long (reference:
src/llm_training/data/preference_tuning/preference_tuning_datamodule.py:16-150)."""
from typing import Any

from ..base_datamodule import DatasetDict
from ..chat_utils import encode_conversations
from ..hf_based import HFBasedDataModule
from .preference_tuning_datacollator import PreferenceTuningDataCollator
from .preference_tuning_datamodule_config import PreferenceTuningDataModuleConfig

Batch = dict[str, list]


def apply_chat_template_and_tokenize(batch: Batch, tokenizer: Any, chat_template: str | None) -> Batch:
    out: Batch = {}
    for side in ('chosen', 'rejected'):
        conversations = [[{'role': 'user', 'content': prompt}, {'role': 'assistant', 'content': answer}]
                         for prompt, answer in zip(batch['prompt'], batch[side])]
        ids, labels = encode_conversations(tokenizer, conversations, chat_template)
        out[f'{side}_input_ids'] = ids
        out[f'{side}_labels'] = labels
        out[f'{side}_length'] = [len(r) for r in ids]
    return out


def drop_overlong_examples(batch: Batch, max_length: int) -> Batch:
    keep = [i for i in range(len(batch['chosen_length']))
            if max(batch['chosen_length'][i], batch['rejected_length'][i]) <= max_length]
    return {k: [v[i] for i in keep] for k, v in batch.items()}


def pre_process_batch(batch: Batch, tokenizer: Any, chat_template: str | None, max_length: int | None,
                      overlong_handling_method: str = 'drop') -> Batch:
    batch = apply_chat_template_and_tokenize(batch, tokenizer, chat_template)
    if max_length is not None:
        batch = drop_overlong_examples(batch, max_length)
    return batch


class PreferenceTuningDataModule(HFBasedDataModule):
    config: PreferenceTuningDataModuleConfig
    datacollator_class = PreferenceTuningDataCollator

    _apply_chat_template_and_tokenize = staticmethod(apply_chat_template_and_tokenize)
    _drop_overlong_examples = staticmethod(drop_overlong_examples)
    _pre_process_data = staticmethod(pre_process_batch)

    def __init__(self, config: PreferenceTuningDataModuleConfig) -> None:
        super().__init__(config)

    def pre_process_data(self, dataset_dict: DatasetDict) -> DatasetDict:
        from datasets import Features, Sequence, Value
        c = self.config
        cols = {}
        for side in ('chosen', 'rejected'):
            cols[f'{side}_input_ids'] = Sequence(Value('int32'))
            cols[f'{side}_labels'] = Sequence(Value('int32'))
            cols[f'{side}_length'] = Value('uint32')
        return self.map_dataset_dict(
            dataset_dict, pre_process_batch,
            fn_kwargs=dict(tokenizer=c.tokenizer, chat_template=c.chat_template, max_length=c.max_length,
                           overlong_handling_method=str(c.overlong_handling_method)),
            batched=True, remove_columns=True, num_proc=c.num_proc, features=Features(cols),
            desc='Pre-processing data')
