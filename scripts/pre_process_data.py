#!/usr/bin/env python
"""Run a datamodule's pre-processing once and store the result at ``pre_processed_data_path`` (+ ``info.txt``).

    python scripts/pre_process_data.py --config cfg.yaml

Same contract as the reference script (scripts/pre_process_data.py:25-47): the config is parsed exactly like
``llm-training fit`` would (nothing is trained), an existing non-empty target directory is left alone.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from llm_training_b200.data import HFBasedDataModule
from llm_training_b200.lightning.cli.config import instantiate, load_config, parse_overrides, set_by_path


class _Tee:
    def __init__(self, *streams) -> None:
        self.streams = streams

    def write(self, data: str) -> None:
        for s in self.streams:
            s.write(data)

    def flush(self) -> None:
        for s in self.streams:
            s.flush()


def main() -> None:
    configs, overrides, _ = parse_overrides(sys.argv[1:])
    if not configs:
        raise SystemExit('usage: pre_process_data.py --config CONFIG [--data.init_args.config.key value ...]')
    config = load_config(configs)
    for key, value in overrides.items():
        set_by_path(config, key, value)
    datamodule = instantiate(config['data'])
    assert isinstance(datamodule, HFBasedDataModule), 'only datasets-backed datamodules can be pre-processed'
    cfg = datamodule.config
    target = cfg.pre_processed_data_path
    assert target is not None, '`pre_processed_data_path` should not be `None`.'
    if not os.path.exists(target) or not os.listdir(target):
        cfg.pre_processed_data_path = None
        datamodule.setup()
        datamodule.save_pre_processed_data(target)
    else:
        print(f'`pre_processed_data_path="{target}"` is not empty, skipping.')
        datamodule.setup()
    with open(os.path.join(target, 'info.txt'), 'w') as f:
        datamodule.print_dataset_info(file=_Tee(sys.stdout, f))


if __name__ == '__main__':
    main()
