"""bench.py contract checks that need no GPU: the reference arm (the compiled reference on the host cores) prints ONE JSON
line with the keys the driver reads and loads no product code; the product arm fails loudly without a CUDA device."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=600):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def test_reference_arm_prints_the_contract_line(ref):
    p = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--cpu-edges-per-core", "1500"])
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines  # exactly one JSON line on stdout
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("C3") and "mode" in d["config"]
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["value"] > 0


def test_reference_arm_loads_no_product_code(ref):
    """The arm's process must not dlopen libppx.so (its record has to show that no product code ran)."""
    code = ("import sys, runpy; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '1', '--cpu-edges-per-core', '500'];\n"
            "import os\n"
            "try:\n    runpy.run_path('bench.py', run_name='__main__')\nexcept SystemExit:\n    pass\n"
            "maps = open('/proc/self/maps').read()\n"
            "sys.stderr.write('LOADED_PPX=%d\\n' % ('libppx.so' in maps))\n"
            "sys.stderr.write('MODULE_PPX=%d\\n' % ('pathplanning_b200' in sys.modules))\n")
    p = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert "LOADED_PPX=0" in p.stderr and "MODULE_PPX=0" in p.stderr, p.stderr[-2000:]


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    p = _run(["--steps", "1", "--warmup", "1", "--no-cpu-baseline", "--no-hybrid", "--no-extras"], timeout=300)
    assert p.returncode != 0
    assert not [l for l in p.stdout.splitlines() if l.strip().startswith("{")]  # no result line from a CPU fallback
