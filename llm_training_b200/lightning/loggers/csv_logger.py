"""Dependency-free logger: ``metrics.csv`` + ``hparams.yaml`` under ``<save_dir>/<name>/version_N``."""
import csv
import os
from typing import Any

import yaml


class CSVLogger:
    def __init__(self, save_dir: str = 'logs', name: str = 'llm_training', version: int | str | None = None) -> None:
        self._save_dir, self.name, self._version = save_dir, name, version
        self._rows: list[dict[str, Any]] = []
        self._keys: list[str] = ['step']

    @property
    def version(self) -> int | str:
        if self._version is None:
            root = os.path.join(self._save_dir, self.name)
            existing = [d for d in (os.listdir(root) if os.path.isdir(root) else []) if d.startswith('version_')]
            self._version = len(existing)
        return self._version

    @property
    def save_dir(self) -> str:
        return self._save_dir

    @property
    def log_dir(self) -> str:
        v = self.version
        return os.path.join(self._save_dir, self.name, v if isinstance(v, str) else f'version_{v}')

    def log_hyperparams(self, params: dict[str, Any]) -> None:
        os.makedirs(self.log_dir, exist_ok=True)
        with open(os.path.join(self.log_dir, 'hparams.yaml'), 'w') as f:
            yaml.safe_dump(params, f, sort_keys=False)

    def log_metrics(self, metrics: dict[str, float], step: int | None = None) -> None:
        row = {'step': step, **metrics}
        for k in row:
            if k not in self._keys:
                self._keys.append(k)
        self._rows.append(row)
        if len(self._rows) % 20 == 0:
            self.save()

    def save(self) -> None:
        os.makedirs(self.log_dir, exist_ok=True)
        with open(os.path.join(self.log_dir, 'metrics.csv'), 'w', newline='') as f:
            w = csv.DictWriter(f, fieldnames=self._keys)
            w.writeheader()
            w.writerows(self._rows)

    def finalize(self, status: str) -> None:
        if self._rows:
            self.save()
