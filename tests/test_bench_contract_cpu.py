"""bench.py's reference arm (`--impl reference`) runs on host cores only, so its contract can be held on the CPU box:
one JSON line with the keys the driver reads, `impl: reference`, the product library never loaded."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                        "--batch", "256"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "fddp_ocp_solves_per_sec_T100" and d["unit"] == "solves/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["dtype"] == "f64"
    assert d["config"]["problems_per_gpu"] == 256 and "workload" in d["config"]
    cb_ = d["cpu_baseline"]
    assert cb_["kind"] == "port" and cb_["cores"] >= 1 and cb_["value"] == d["value"] and cb_["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_does_not_load_the_product_library():
    code = ("import sys, runpy; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0', '--batch', '64'];"
            "runpy.run_path(%r, run_name='__main__');"
            "print('PRODUCT_LOADED' if any('libcrocoddyl_b200' in l for l in open('/proc/self/maps')) else 'PRODUCT_NOT_LOADED')"
            % os.path.join(ROOT, "bench.py"))
    r = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "PRODUCT_NOT_LOADED" in r.stdout
