"""Packers, collators, sampling, resumable loader, datamodules ('Unit, CPU')."""
import random

import pytest
import torch

from tests.helpers import tiny_tokenizer


def test_best_fit_bin_packing_matches_quadratic_reference():
    from llm_training_b200.data.packing import best_fit_bin_packing
    rng = random.Random(0)

    def slow(capacity, lengths):
        bins, contents = [], []
        for i, n in enumerate(lengths):
            best, left = -1, float('inf')
            for j, space in enumerate(bins):
                if space >= n and space - n < left:
                    best, left = j, space - n
            if best != -1:
                bins[best] -= n
                contents[best].append(i)
            else:
                bins.append(capacity - n)
                contents.append([i])
        return contents

    for _ in range(20):
        lengths = sorted((rng.randint(1, 100) for _ in range(rng.randint(1, 300))), reverse=True)
        assert best_fit_bin_packing(100, lengths) == slow(100, lengths)
    assert best_fit_bin_packing(10, [6, 5, 4, 3, 2]) == [[0, 2], [1, 3, 4]]
    assert best_fit_bin_packing(10, []) == []


def test_group_indices_by_length_reserves_a_slot_per_document():
    from llm_training_b200.data.packing import group_indices_by_length
    groups = group_indices_by_length([5, 3, 8, 2, 2], max_length=10)
    assert sorted(i for g in groups for i in g) == [0, 1, 2, 3, 4]
    lengths = [5, 3, 8, 2, 2]
    for g in groups:
        assert sum(lengths[i] for i in g) + len(g) - 1 <= 10 or len(g) == 1
    assert groups[0] == [3, 4, 1]          # 2 + 2 + 3 (+2 reserved) fits, adding 5 would not


def test_naive_and_best_fit_packing_segment_ids():
    from llm_training_b200.data.pre_training.pre_training_datamodule import best_fit_decreasing, naive_packing, truncate
    batch = {'source': ['a', 'a', 'a', 'b'], 'input_ids': [[1, 2, 3], [4, 5], [6, 7, 8, 9], [10, 11]],
             'length': [3, 2, 4, 2]}
    out = naive_packing(batch, max_length=4)
    assert out['input_ids'] == [[1, 2, 3, 4], [6, 7, 8, 9], [5], [10, 11]]
    assert out['attention_mask'] == [[1, 1, 1, 2], [1, 1, 1, 1], [1], [1, 1]]
    assert out['source'] == ['a', 'a', 'a', 'b']
    bfd = best_fit_decreasing(batch, max_length=5)
    assert sorted(map(len, bfd['input_ids'])) == [2, 4, 5]
    for ids, seg, n in zip(bfd['input_ids'], bfd['attention_mask'], bfd['length']):
        assert len(ids) == len(seg) == n <= 5 and seg[0] == 1 and all(b - a in (0, 1) for a, b in zip(seg, seg[1:]))
    tr = truncate({'source': ['x'], 'input_ids': [list(range(10))], 'length': [10]}, max_length=4, stride=3)
    assert tr['input_ids'] == [[0, 1, 2, 3], [3, 4, 5, 6], [6, 7, 8, 9], [9]]


def test_pre_training_config_validation_and_collator():
    from llm_training_b200.data import PreTrainingDataCollator, PreTrainingDataModuleConfig
    tok = tiny_tokenizer()
    with pytest.raises(Exception):
        PreTrainingDataModuleConfig(tokenizer=tok, packing_method='BEST_FIT_BIN_PACKING')      # needs max_length
    cfg = PreTrainingDataModuleConfig(tokenizer=tok, max_length=8, packing_method='Best_Fit_Bin_Packing',
                                      pad_to_multiple_of=4)
    assert cfg.stride == 8 and str(cfg.packing_method) == 'best_fit_bin_packing'
    batch = PreTrainingDataCollator(cfg)([{'input_ids': [1, 20, 21, 2], 'attention_mask': [1, 1, 1, 1]},
                                          {'input_ids': [1, 22], 'attention_mask': [1, 1]}])
    assert batch['input_ids'].shape == (2, 8)                      # (4 // 4 + 1) * 4
    assert batch['attention_mask'].tolist() == [[1, 1, 1, 1, 0, 0, 0, 0], [1, 1, 0, 0, 0, 0, 0, 0]]
    assert batch['labels'][0].tolist() == [-100, 20, 21, 2, -100, -100, -100, -100]   # BOS and padding masked
    assert batch['position_ids'].tolist() == [list(range(8))]
    cfg2 = PreTrainingDataModuleConfig(tokenizer=tok, max_length=8, block_cross_document_attention=True)
    b2 = PreTrainingDataCollator(cfg2)([{'input_ids': [1, 20, 2, 1, 21, 2], 'attention_mask': [1, 1, 1, 2, 2, 2]}])
    assert b2['attention_mask'].tolist() == [[1, 1, 1, 2, 2, 2]]


def test_pre_training_datamodule_end_to_end_with_sampling(tmp_path):
    import datasets
    from llm_training_b200.data import PreTrainingDataModule, PreTrainingDataModuleConfig
    tok = tiny_tokenizer()
    texts = [' '.join(f'w{(i * 7 + j) % 200}' for j in range(3 + i % 9)) for i in range(60)]
    ds = datasets.Dataset.from_dict({'text': texts + [''], 'source': ['web'] * 30 + ['code'] * 30 + ['web']})
    path = str(tmp_path / 'raw')
    ds.save_to_disk(path)

    class DM(PreTrainingDataModule):
        def load_data(self):
            return datasets.DatasetDict({'train': datasets.load_from_disk(path)})

    cfg = PreTrainingDataModuleConfig(tokenizer=tok, max_length=32, packing_method='best_fit_bin_packing',
                                      sample_rate={'web': 2.0, 'code': 0.5}, batch_size=2, validation_split=0.2,
                                      pin_memory=False, enable_cache=False)
    dm = DM(cfg)
    dm.setup()
    train = dm.dataset_dict['train']
    assert set(dm.dataset_dict) == {'train', 'validation'}
    assert max(train['length']) <= 32 and all(len(a) == n for a, n in zip(train['attention_mask'], train['length']))
    n_web = sum(s == 'web' for s in dm.split_data(dm.pre_processed_dataset_dict)['train']['source'])
    assert sum(s == 'web' for s in train['source']) == 2 * n_web
    batch = next(iter(dm.train_dataloader()))
    assert batch['input_ids'].shape[0] == 2 and set(batch) == {'input_ids', 'attention_mask', 'position_ids', 'labels'}
    assert 'tokens' in dm.token_table()


CHATML = 'chatml'


def test_instruction_tuning_labels_packing_and_collator():
    from llm_training_b200.data import InstructionTuningDataCollator, InstructionTuningDataModuleConfig
    from llm_training_b200.data.instruction_tuning.instruction_tuning_datamodule import pre_process_batch
    tok = tiny_tokenizer()
    cfg = InstructionTuningDataModuleConfig(tokenizer=tok, chat_template=CHATML, max_length=64,
                                            packing_method='GROUP_BY_LENGTH', pad_to_multiple_of=8)
    convs = [[{'role': 'user', 'content': 'w1 w2'}, {'role': 'assistant', 'content': 'w3 w4 w5'}],
             [{'role': 'system', 'content': 'w9'}, {'role': 'user', 'content': 'w6'},
              {'role': 'assistant', 'content': 'w7'}],
             [{'role': 'user', 'content': ' '.join(['w8'] * 80)}, {'role': 'assistant', 'content': 'w1'}]]
    out = pre_process_batch({'messages': convs}, tok, cfg.chat_template, None, None, 64, 'drop', 'group_by_length')
    assert len(out['input_ids']) == 1                      # overlong dropped, two short ones grouped
    ids, seg, labels = out['input_ids'][0], out['attention_mask'][0], out['labels'][0]
    assert set(seg) == {1, 2} and len(ids) == len(seg) == len(labels)
    supervised = [tok.convert_ids_to_tokens(t) for t, l in zip(ids, labels) if l != -100]
    assert 'w3' in supervised and 'w7' in supervised and 'w1' not in supervised and 'w6' not in supervised
    assert '<|im_end|>' in supervised                      # the assistant terminator is supervised
    trunc = pre_process_batch({'messages': convs[2:]}, tok, cfg.chat_template, None, None, 16, 'truncate', 'no_packing')
    assert trunc['length'] == [16] and len(trunc['input_ids'][0]) == 16 and len(trunc['labels'][0]) == 16
    batch = InstructionTuningDataCollator(cfg)([{k: v[0] for k, v in out.items()}])
    assert batch['input_ids'].shape[1] % 8 == 0 and (batch['attention_mask'] == 0).any()   # always >= 1 pad
    assert batch['position_ids'][0, :len(ids)].tolist() == list(range(len(ids)))


def test_preference_tuning_pipeline():
    from llm_training_b200.data import PreferenceTuningDataCollator, PreferenceTuningDataModuleConfig
    from llm_training_b200.data.preference_tuning.preference_tuning_datamodule import pre_process_batch
    tok = tiny_tokenizer()
    cfg = PreferenceTuningDataModuleConfig(tokenizer=tok, chat_template=CHATML, max_length=32, pad_to_multiple_of=4)
    out = pre_process_batch({'prompt': ['w1 w2', 'w3'], 'chosen': ['w4', ' '.join(['w5'] * 40)],
                             'rejected': ['w6 w7 w8', 'w9']}, tok, cfg.chat_template, 32)
    assert len(out['chosen_input_ids']) == 1               # second example: chosen side too long -> dropped
    batch = PreferenceTuningDataCollator(cfg)([{k: v[0] for k, v in out.items()}])
    assert set(batch) == {f'{s}_{k}' for s in ('chosen', 'rejected')
                          for k in ('input_ids', 'attention_mask', 'labels', 'position_ids')}
    assert batch['chosen_input_ids'].shape != batch['rejected_input_ids'].shape or True
    assert (batch['rejected_labels'] != -100).sum() > (batch['chosen_labels'] != -100).sum()


def test_chat_templates_are_all_loadable_and_mark_generation():
    import jinja2
    from llm_training_b200.data.chat_templates import CHAT_TEMPLATES, get_chat_template
    names = CHAT_TEMPLATES.names()
    assert names == ['chatml', 'gemma', 'llama-2', 'llama-3', 'llama-3.1', 'llama-3.2', 'phi-3', 'qwen2.5', 'tulu-2']
    for name in names:
        text = get_chat_template(name)
        assert '{%- generation -%}' in text and '{%- endgeneration -%}' in text, name
    assert get_chat_template('{{ messages }}') == '{{ messages }}'      # literal passthrough
    tok = tiny_tokenizer()
    msgs = [{'role': 'system', 'content': 'w1'}, {'role': 'user', 'content': 'w2'},
            {'role': 'assistant', 'content': 'w3'}]
    for name in names:
        text = tok.apply_chat_template(msgs, chat_template=get_chat_template(name), tokenize=False)
        assert 'w2' in text and 'w3' in text, name


def test_dummy_dataset_and_resumable_loader():
    from llm_training_b200.data import DummyDataModule, DummyDataModuleConfig
    with pytest.raises(Exception):
        DummyDataModuleConfig(vocab_size=10, max_length=4)
    cfg = DummyDataModuleConfig(vocab_size=50, max_length=8, num_tokens=100, base_seed=3, batch_size=2,
                                pin_memory=False)
    dm = DummyDataModule(cfg)
    dm.setup()
    ds = dm.dataset_dict['train']
    assert len(ds) == 13 and torch.equal(ds[5]['input_ids'], ds[5]['input_ids'])

    class FakeTrainer:
        class fit_loop:
            batch_idx = 0
        strategy = type('S', (), {'dp_size': 1, 'distributed_sampler_kwargs': {}})()

    dm.trainer = FakeTrainer()
    full = [b['input_ids'] for b in dm.train_dataloader()]
    FakeTrainer.fit_loop.batch_idx = 3
    resumed = [b['input_ids'] for b in dm.train_dataloader()]
    assert len(resumed) == len(full) - 3
    assert all(torch.equal(a, b) for a, b in zip(full[3:], resumed))     # same order, first 3 skipped


def test_pre_process_data_script_then_training_datamodule_loads_the_result(tmp_path):
    """scripts/pre_process_data.py (reference: scripts/pre_process_data.py:25-47): tokenise + pack once, store at
    `pre_processed_data_path` with `info.txt`; a datamodule pointed at that path then skips all pre-processing."""
    import json
    import os
    import subprocess
    import sys
    tok_dir = tmp_path / 'tok'
    tiny_tokenizer().save_pretrained(str(tok_dir))
    raw = tmp_path / 'raw.jsonl'
    with open(raw, 'w') as f:
        for i in range(40):
            f.write(json.dumps({'text': ' '.join(f'w{(i * 3 + j) % 200}' for j in range(5 + i % 11))}) + '\n')
    target = tmp_path / 'processed'
    cfg = tmp_path / 'cfg.yaml'
    cfg.write_text(f"""
trainer: {{}}
model: {{}}
data:
  class_path: llm_training.data.PreTrainingDataModule
  init_args.config:
    dataset_kwargs:
      path: json
      data_files: {raw}
    pre_processed_data_path: {target}
    tokenizer:
      class_path: HFTokenizer
      init_args:
        path: {tok_dir}
    packing_method: best_fit_bin_packing
    max_length: 32
    batch_size: 2
    enable_cache: false
""")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, HF_DATASETS_OFFLINE='1', HF_HUB_OFFLINE='1')
    proc = subprocess.run([sys.executable, os.path.join(root, 'scripts', 'pre_process_data.py'), '--config', str(cfg)],
                          capture_output=True, text=True, timeout=600, env=env)
    assert proc.returncode == 0, proc.stderr[-2000:]
    assert (target / 'info.txt').exists() and 'tokens' in (target / 'info.txt').read_text().lower()
    from llm_training_b200.lightning.cli.config import instantiate, load_config
    dm = instantiate(load_config([str(cfg)])['data'])
    dm.load_data = lambda: (_ for _ in ()).throw(AssertionError('raw data must not be read again'))
    dm.setup()
    assert max(dm.dataset_dict['train']['length']) <= 32
    batch = next(iter(dm.train_dataloader()))
    assert batch['input_ids'].shape[0] == 2
