"""The driver's contract for ``bench.py`` (one JSON line, fixed keys) checked on CPU with the tiny model."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args: str) -> dict:
    proc = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), *args], capture_output=True, text=True,
                          timeout=600, cwd=ROOT, env=dict(os.environ, CUDA_VISIBLE_DEVICES=''))
    assert proc.returncode == 0, proc.stderr[-2000:]
    lines = [l for l in proc.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines                       # exactly ONE line on stdout
    return json.loads(lines[0])


def test_bench_prints_one_json_line_with_the_contract_keys():
    out = _run('--model', 'llama-tiny', '--steps', '2', '--warmup', '3', '--seq-len', '64')
    for key in ('metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling',
                'vs_baseline', 'dtype', 'data', 'config', 'clocks', 'e2e', 'gpu_launches'):
        assert key in out, key
    assert out['metric'] == 'tokens_per_sec' and out['unit'] == 'tokens/s' and out['higher_is_better'] is True
    assert out['n_gpus'] == 1 and out['steps'] == 2 and out['warmup'] >= 3 and out['scaling'] == 'weak'
    assert out['data'] == 'synthetic' and out['value'] > 0 and out['ms_per_step'] > 0
    assert {'model', 'global_batch', 'seq_len', 'parallelism'} <= set(out['config'])
    assert {'value', 'unit', 'h2d_bytes_per_step', 'd2h_bytes_per_step'} <= set(out['e2e'])
    assert out['e2e']['h2d_bytes_per_step'] > 0 and out['e2e']['d2h_bytes_per_step'] > 0
    assert {'sm_mhz', 'sm_max_mhz', 'reasons'} <= set(out['clocks'])


def test_bench_reference_arm_reports_json():
    out = _run('--impl', 'reference')
    assert out['impl'] == 'reference' and ('unavailable' in out or 'value' in out)
