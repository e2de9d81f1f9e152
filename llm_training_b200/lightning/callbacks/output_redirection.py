"""Tee stdout / stderr / python loggers into ``<log_dir>/<index>-<version>.log``
(reference: src/llm_training/lightning/callbacks/output_redirection.py:13-101)."""
import logging
import os
import sys
from typing import TextIO

from .base import Callback


class _Tee:
    def __init__(self, stream: TextIO, sink: TextIO) -> None:
        self.stream, self.sink = stream, sink

    def write(self, data: str) -> int:
        self.sink.write(data)
        self.sink.flush()
        return self.stream.write(data)

    def flush(self) -> None:
        self.stream.flush()
        self.sink.flush()

    def __getattr__(self, name):
        return getattr(self.stream, name)


class OutputRedirection(Callback):
    LOG_FILE_SUFFIX = '.log'

    def __init__(self, log_file_name: str | None = None, redirect_stdout: bool = True, redirect_stderr: bool = True,
                 redirect_loggers: bool = True, rank_zero_only: bool = True, enabled: bool = True) -> None:
        """``log_file_name`` may contain ``{index}`` (number of logs already in the directory) and ``{version}``
        (logger version), like the reference's default ``'{index}-{version}'``; ``enabled=False`` turns the callback
        into a no-op (reference: output_redirection.py:16-33)."""
        self.enabled = enabled
        self.log_file_name = log_file_name
        self.redirect_stdout, self.redirect_stderr = redirect_stdout, redirect_stderr
        self.redirect_loggers = redirect_loggers
        self.rank_zero_only = rank_zero_only
        self._file: TextIO | None = None
        self._saved: tuple | None = None
        self._handler: logging.Handler | None = None

    def _path(self, trainer) -> str:
        log_dir = trainer.log_dir
        os.makedirs(log_dir, exist_ok=True)
        existing = [f for f in os.listdir(log_dir) if f.endswith(self.LOG_FILE_SUFFIX)]
        version = getattr(trainer.logger, 'version', None) or 'run'
        if self.log_file_name:
            name = self.log_file_name.format(index=len(existing), version=version) \
                if '{' in self.log_file_name else self.log_file_name
            if not name.endswith(self.LOG_FILE_SUFFIX) and '{' in self.log_file_name:
                name += self.LOG_FILE_SUFFIX
            return os.path.join(log_dir, name)
        return os.path.join(log_dir, f'{len(existing)}-{version}{self.LOG_FILE_SUFFIX}')

    def setup(self, trainer, pl_module, stage: str) -> None:
        if not self.enabled or self._file is not None or (self.rank_zero_only and not trainer.is_global_zero):
            return
        path = trainer.strategy.broadcast(self._path(trainer)) if not self.rank_zero_only else self._path(trainer)
        if not trainer.is_global_zero:
            root, ext = os.path.splitext(path)
            path = f'{root}.rank{trainer.global_rank}{ext}'
        self._file = open(path, 'a', buffering=1)
        self._saved = (sys.stdout, sys.stderr)
        if self.redirect_stdout:
            sys.stdout = _Tee(sys.stdout, self._file)
        if self.redirect_stderr:
            sys.stderr = _Tee(sys.stderr, self._file)
        if self.redirect_loggers:
            self._redirect_loggers()

    def _redirect_loggers(self) -> None:
        """Attach a handler writing to the log file to the package loggers (reference: ``redirect_loggers``,
        output_redirection.py:60-73)."""
        self._handler = logging.StreamHandler(self._file)
        self._handler.setFormatter(logging.Formatter('[%(asctime)s][%(levelname)s][%(name)s] %(message)s'))
        for name in ('llm_training_b200', 'llm_training'):
            logging.getLogger(name).addHandler(self._handler)

    def teardown(self, trainer, pl_module, stage: str) -> None:
        if self._file is None:
            return
        sys.stdout, sys.stderr = self._saved
        if self._handler is not None:
            for name in ('llm_training_b200', 'llm_training'):
                logging.getLogger(name).removeHandler(self._handler)
        self._file.close()
        self._file = None
