"""Third-party version shims for running the UNMODIFIED reference (baseline/_ref) on this image's stack.

Nothing here touches the reference's own code or llm_training_b200.  One library is missing and two library symbols
moved since the versions the reference pins (`pyproject.toml:4-23` of the reference):

* ``accelerate`` is not on this image and cannot be installed offline; the reference imports one symbol from it
  (``init_empty_weights``, `/root/reference/src/llm_training/models/hf_compat_model/hf_compat_model.py:5`).  A module
  object with that one function is registered AFTER transformers has been imported, so transformers itself keeps
  seeing "accelerate not installed";
* ``transformers.modeling_utils.no_init_weights`` (imported at
  `/root/reference/src/llm_training/models/hf_causal_lm/hf_causal_lm.py:10`) is gone in transformers 5.x;
* ``LigerCrossEntropyFunction.apply`` took ``(logits, target, ignore_index)`` in liger-kernel 0.4.2 (the call at
  `/root/reference/src/llm_training/ops/liger_kernel/cross_entropy_op.py:29-33`) and takes
  ``(logits, target, weight, ignore_index, ...)`` in 0.8.0.
"""
from contextlib import contextmanager


def install() -> list[str]:
    import sys
    import types
    notes = []
    import transformers.modeling_utils as mu
    if 'accelerate' not in sys.modules:
        try:
            import accelerate  # noqa: F401
        except ImportError:
            import torch

            @contextmanager
            def init_empty_weights(include_buffers: bool = False):
                with torch.device('meta'):
                    yield
            stub = types.ModuleType('accelerate')
            stub.init_empty_weights = init_empty_weights
            sys.modules['accelerate'] = stub
            notes.append('accelerate.init_empty_weights stubbed (accelerate not installed)')
    if not hasattr(mu, 'no_init_weights'):
        @contextmanager
        def no_init_weights(_enable: bool = True):
            yield
        mu.no_init_weights = no_init_weights
        notes.append('transformers.modeling_utils.no_init_weights stubbed (removed in transformers 5)')
    try:
        import inspect

        from liger_kernel.ops import cross_entropy as lce
        params = list(inspect.signature(lce.LigerCrossEntropyFunction.forward).parameters)
        if len(params) > 3 and params[3] == 'weight':
            new_apply = lce.LigerCrossEntropyFunction.apply

            class _OldSignature:
                """`apply(logits, target, ignore_index)` of liger-kernel 0.4.2 on top of the 0.8 kernel."""
                @staticmethod
                def apply(logits, target, ignore_index=-100, *rest):
                    out = new_apply(logits, target, None, ignore_index, *rest)
                    return out if isinstance(out, tuple) else (out, None)
            lce.LigerCrossEntropyFunction = _OldSignature
            notes.append('LigerCrossEntropyFunction.apply adapted from the 0.4.2 to the 0.8 positional signature')
    except ImportError:
        notes.append('liger_kernel not importable')
    return notes
