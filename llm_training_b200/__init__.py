"""llm_training_b200 — a Blackwell (sm_100a) native LLM training framework.

Public surface mirrors cchou0519/LLM-Training (reference: src/llm_training/__init__.py:1-6):
importing the package only configures the package logger.
"""
import logging as _logging

__version__ = '0.1.0'

_handler = _logging.StreamHandler()
_handler.setFormatter(_logging.Formatter('[%(asctime)s][%(levelname)s][%(name)s] %(message)s',
                                         datefmt='%Y-%m-%d %H:%M:%S'))
logger = _logging.getLogger(__name__)
if not logger.handlers:
    logger.addHandler(_handler)
logger.propagate = False
