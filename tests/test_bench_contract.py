"""bench.py's contract on the CPU: both arms print the same `config` (built from the workload alone), and the reference arm
prints one JSON line with the keys the driver reads."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_config_is_a_function_of_the_workload():
    sys.path.insert(0, ROOT)
    import bench
    for name, w in bench.WORKLOADS.items():
        a = bench.make_config(name, dict(w), 1)
        b = bench.make_config(name, dict(w), 1)
        assert a == b and a['name'] == name and 'cpu_reference_sample' in a and 'workload' in a
        assert 'model' not in a
    assert set(bench.SECONDARY) <= set(bench.WORKLOADS)


def test_reference_arm_prints_one_json_line():
    env = dict(os.environ, OMP_NUM_THREADS='2')
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '2', '--warmup', '3'],
                         capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['metric'] == 'DG DOF-updates/sec per RK stage' and d['unit'] == 'DOF-updates/s'
    assert d['higher_is_better'] is True and d['value'] > 0 and d['steps'] == 2
    assert d['cpu_baseline']['kind'] == 'port' and d['cpu_baseline']['cores'] >= 1 and d['cpu_baseline']['value'] == d['value']
    assert d['e2e'] == {'value': d['value'], 'unit': d['unit'], 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    sys.path.insert(0, ROOT)
    import bench
    assert d['config'] == bench.make_config('c4', dict(bench.WORKLOADS['c4']), 1)
