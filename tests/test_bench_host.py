"""bench.py host logic that needs no GPU: the algorithmic work per unit equals SURVEY.md 8(d)'s figures,
the reference arm describes the same config as the GPU arm, and never imports the product library."""
import json
import os
import subprocess
import sys

from conftest import ROOT

sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_algorithmic_figures_match_the_survey():
    assert bench.alg_flops_per_unit("cml_gcn") == 6_747_520
    assert sum(bench.lstm_flops("cml_gcn")) == 6_722_816
    assert bench.alg_flops_per_unit("cml_baseline") == 6_376_832
    assert bench.alg_flops_per_unit("soilnet_gcn") == 12_728_832
    assert bench.alg_flops_per_unit("ig", 50) == 51 * 2 * 6_747_520
    assert bench.alg_flops_per_unit("train_b128") == 3 * 6_747_520
    assert bench.alg_bytes_per_unit("cml_gcn") == 34_848
    assert bench.alg_bytes_per_unit("cml_baseline") == 1_452
    assert bench.alg_bytes_per_unit("soilnet_gcn") == 4_044 + 4
    # dead-step elimination: floor(floor(floor(T/3)/3)/3) * 27 ...
    for T, live in bench.LIVE.items():
        assert live[0] == 27 * (((T // 3) // 3) // 3) and live[6] == ((T // 3) // 3) // 3


def test_reference_arm_same_config_and_no_product_library():
    code = ("import sys, json; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0', '--workload', 'cml_baseline'];"
            "import bench; bench.REF_SAMPLE['cml_baseline'] = 256; bench.main();"
            "maps = open('/proc/self/maps').read(); assert 'libgnnqc' not in maps, 'reference arm mapped the product library';"
            "assert not any(m == 'gnnqc' or m.startswith('gnnqc.') for m in sys.modules)")
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["config"] == bench.make_config("cml_baseline", bench.WORKLOADS["cml_baseline"][3])
    assert line["metric"] == bench.WORKLOADS["cml_baseline"][0]
