"""Value-by-value parity with the UNMODIFIED reference's own torch code ("the reference's own torch code is
the oracle where it exists").  The reference (installed in ``baseline/_ref``) is imported in a
subprocess so that its ``llm_training`` package and this repository's alias of the same name never meet."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, 'baseline', '_ref')


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, 'llm_training')), reason='reference not installed in baseline/_ref')
def test_ops_match_reference_torch_code(tmp_path):
    dump = tmp_path / 'ref.pt'
    env = dict(os.environ, PYTHONPATH=REF)
    cases = os.path.join(ROOT, 'tests', 'ref_oracle_cases.py')
    proc = subprocess.run([sys.executable, cases, str(dump)], cwd=str(tmp_path), env=env, capture_output=True,
                          text=True, timeout=600)
    assert proc.returncode == 0, proc.stderr[-2000:]
    want = torch.load(dump)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import llm_training  # noqa: F401  -- this repository's alias
    assert 'llm_training_b200' in sys.modules
    from ref_oracle_cases import build_cases
    got = build_cases()
    assert set(got) == set(want), sorted(set(got) ^ set(want))
    known_deviations = set()
    for name in sorted(want):
        a, b = want[name], got[name]
        if isinstance(a, str):
            assert a == b, (name, a, b)
            continue
        assert a.shape == b.shape, (name, a.shape, b.shape)
        if name in known_deviations:
            continue
        rtol, atol = (2e-2, 4e-2) if 'bf16' in name else (1e-5, 1e-5)    # bf16: one rounding instead of two
        assert torch.allclose(a.float(), b.float(), rtol=rtol, atol=atol), (name, (a.float() - b.float()).abs().max())
