#!/usr/bin/env python
"""Export a training checkpoint to Hugging Face format.

    python scripts/convert_to_hf.py <checkpoint> [output_dir] [--config_path cfg.yaml]
                                    [--eos_token_id ID[,ID...]] [--dtype bfloat16]

``<checkpoint>`` is an ``epoch=E-step=S.ckpt`` DCP directory (``*.distcp`` + ``.metadata`` + ``meta.pt``) or a
single-file checkpoint.  The LM and datamodule are rebuilt from the config stored INSIDE the checkpoint (or
``--config_path``), keys missing from the checkpoint (frozen modules) are filled from the pre-trained weights,
and the model (+ tokenizer, chat template, ``model_max_length``) is written with ``save_pretrained``.
Same behaviour as the reference script (scripts/convert_to_hf.py:18-177).
"""
import argparse
import os
import sys
from pathlib import Path

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import yaml

from llm_training_b200.lightning.checkpoint_io import read_checkpoint_state_dict
from llm_training_b200.lightning.cli.config import expand_dotted, instantiate
from llm_training_b200.lms import BaseLightningModule
from llm_training_b200.models import HFCompatModel
from llm_training_b200.models.utils import init_empty_weights

_PRECISION_DTYPE = {'32-true': torch.float32, '32': torch.float32, '16-true': torch.float16, '16-mixed': torch.float32,
                    'bf16-true': torch.bfloat16, 'bf16-mixed': torch.float32, 'bf16': torch.float32}


def dtype_from_config(config: dict) -> torch.dtype:
    return _PRECISION_DTYPE.get(str((config.get('trainer') or {}).get('precision', '32-true')), torch.float32)


def convert(checkpoint_path: str, output_dir: str | None = None, config_path: str | None = None,
            eos_token_id: list[int] | None = None, dtype: torch.dtype | None = None) -> Path:
    checkpoint_path = Path(checkpoint_path)
    out = Path(output_dir) if output_dir else checkpoint_path.parent / 'hf' / checkpoint_path.stem
    print('Reading checkpoint')
    state_dict, meta = read_checkpoint_state_dict(str(checkpoint_path))
    if config_path is not None:
        with open(config_path) as f:
            config = expand_dotted(yaml.safe_load(f))
    else:
        config = meta.get('config')
        if config is None:
            raise SystemExit('the checkpoint does not embed its config (SaveConfigCallback); pass --config_path')
    dtype = dtype or dtype_from_config(config)
    lm = instantiate(config['model'])
    datamodule = instantiate(config['data']) if config.get('data') else None
    assert isinstance(lm, BaseLightningModule)
    lm.config.load_weights = False
    lm.config.init_weights = False
    with init_empty_weights(include_buffers=False):
        lm.configure_model()
    required = lm.required_keys
    missing = required - state_dict.keys()
    if missing:
        print(f'{len(missing)} keys are missing from the checkpoint, taking them from the pre-trained weights')
        original = lm.get_pre_trained_weights()
        state_dict |= {k: original[k] for k in missing}
    result = lm.load_state_dict(state_dict, strict=False, assign=True)
    still_missing = set(result.missing_keys) & required
    assert not still_missing, f'Missing keys: {sorted(still_missing)}'
    model = lm.get_model()
    assert isinstance(model, HFCompatModel), f'{type(model).__name__} cannot be converted to a Hugging Face model'
    model.config.torch_dtype = dtype
    hf_model = model.get_hf_model()
    if eos_token_id:
        hf_model.config.eos_token_id = eos_token_id[0]
        if getattr(hf_model, 'generation_config', None) is not None:
            hf_model.generation_config.eos_token_id = eos_token_id if len(eos_token_id) > 1 else eos_token_id[0]
    print(f'Saving model to {out}')
    hf_model.to(dtype).save_pretrained(out)
    tokenizer = getattr(getattr(datamodule, 'config', None), 'tokenizer', None)
    if tokenizer is not None:
        print('Saving tokenizer')
        max_length = getattr(datamodule.config, 'max_length', None)
        if max_length:
            tokenizer.model_max_length = max(tokenizer.model_max_length, max_length)
        chat_template = getattr(datamodule.config, 'chat_template', None)
        if chat_template is not None:
            tokenizer.chat_template = chat_template
        tokenizer.save_pretrained(out)
    return out


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument('checkpoint_path')
    ap.add_argument('output_dir', nargs='?')
    ap.add_argument('--config_path')
    ap.add_argument('--eos_token_id', type=lambda s: [int(x) for x in s.split(',')])
    ap.add_argument('--dtype', type=lambda s: getattr(torch, s.removeprefix('torch.')))
    a = ap.parse_args()
    convert(a.checkpoint_path, a.output_dir, a.config_path, a.eos_token_id, a.dtype)


if __name__ == '__main__':
    main()
