"""bench.py's contract, as far as it can be checked without a GPU: the reference arm (the CPU
stand-in: the oracle port on the host cores) prints the agreed JSON line, and the product arm
refuses to run on the host."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py')] + list(args), cwd=ROOT,
                          capture_output=True, text=True, timeout=600)


def test_reference_arm_prints_the_contract_line():
    out = _run('--impl', 'reference', '--steps', '1', '--warmup', '0')
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line['impl'] == 'reference'
    assert line['metric'] == 'particle-timesteps/s' and line['unit'] == 'particle-timesteps/s'
    assert line['higher_is_better'] is True and line['vs_baseline'] is None
    assert line['value'] > 0 and line['steps'] == 1 and line['n_gpus'] == 1
    assert 'N=1200000' in line['config']['workload']
    cpu = line['cpu_baseline']
    assert cpu['kind'] == 'port' and cpu['cores'] >= 1 and cpu['value'] == line['value'] and cpu['sample']
    e2e = line['e2e']
    assert e2e['value'] == line['value'] and e2e['h2d_bytes_per_step'] == 0 and e2e['d2h_bytes_per_step'] == 0


def test_product_arm_has_no_cpu_path():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip('a GPU is present')
    out = _run('--steps', '1', '--warmup', '1')
    assert out.returncode != 0
    assert 'no CUDA device' in out.stderr + out.stdout
    assert not any(l.startswith('{') for l in out.stdout.splitlines())
