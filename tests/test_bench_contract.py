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


def test_bench_honours_gpus_and_records_the_data_plane():
    """`--gpus N` is the number of ranks that run (the round-1 bench ignored it and ran on every visible GPU)."""
    out = _run('--model', 'llama-tiny', '--steps', '2', '--warmup', '3', '--seq-len', '64', '--gpus', '2')
    assert out['n_gpus'] == 2 and out['config']['parallelism'] == 'dp2' and out['config']['global_batch'] == 2
    assert {'comm', 'reshard', 'multicast'} <= set(out['config'])
    assert out['config']['reshard'] is True and 'ZeRO-3' in out['config']['sharding']     # the default is FSDP / ZeRO-3
    out = _run('--model', 'llama-tiny', '--steps', '2', '--warmup', '3', '--seq-len', '64', '--gpus', '2', '--resident')
    assert out['n_gpus'] == 2 and out['config']['reshard'] is False


def test_bench_rejects_a_world_size_that_differs_from_gpus():
    proc = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--gpus', '1', '--model', 'llama-tiny'],
                          capture_output=True, text=True, timeout=120, cwd=ROOT,
                          env=dict(os.environ, CUDA_VISIBLE_DEVICES='', WORLD_SIZE='4', RANK='0'))
    assert proc.returncode == 2 and 'must agree' in proc.stderr


def test_bench_reference_arm_runs_the_reference_modules():
    """The reference arm drives the unmodified tree in baseline/_ref (tiny config on CPU here) and never imports
    this package; without baseline/_ref it must say `unavailable` and exit 0."""
    out = _run('--impl', 'reference', '--model', 'llama-tiny', '--steps', '2', '--warmup', '3', '--seq-len', '64')
    assert out['impl'] == 'reference'
    if 'unavailable' in out:
        assert not os.path.isdir(os.path.join(ROOT, 'baseline', '_ref', 'llm_training')), out
        return
    assert out['n_gpus'] == 1 and out['value'] > 0 and out['ours_loaded'] is False
    assert out['e2e']['h2d_bytes_per_step'] > 0 and out['e2e']['d2h_bytes_per_step'] > 0
    assert out['reference_class'] == 'reference modules, bare loop'


def test_bench_host_profile_is_written_next_to_the_one_json_line(tmp_path):
    """`--host-profile` (development aid) writes rank 0's enqueue-path profile to a file and leaves stdout alone."""
    target = tmp_path / 'prof' / 'host.txt'
    out = _run('--model', 'llama-tiny', '--steps', '1', '--warmup', '3', '--seq-len', '64', '--phases',
               '--host-profile', str(target))
    assert out['steps'] == 1
    text = target.read_text()
    assert text.startswith('host wall') and 'tottime' in text
