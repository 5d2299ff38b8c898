"""bench.py's contract pieces that can be checked without a GPU: the algorithmic-byte model of
SURVEY.md 8(d), the reference arm's JSON line, and the refusal to run the product path on a CPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_stage_bytes_sum_to_the_survey_formula():
    sys.path.insert(0, ROOT)
    import bench
    n, k, i, p, t = 100_000, 16, 236_810, 512 * 512, 1024
    per = bench.stage_bytes_per_view(n, k, i, p, t)
    total = sum(per.values())
    assert total == (24 * k + 436) * n + 52 * i + 44 * p + 8 * t        # B_alg, SURVEY.md 8(d)
    fwd = sum(per[s] for s in ("project_sh_fwd", "cumsum", "map_isect", "sort", "bin_edges", "raster_fwd"))
    assert fwd == (12 * k + 188) * n + 48 * i + 20 * p + 8 * t           # B_fwd


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--ref-views", "1"], capture_output=True, text=True, timeout=600,
                         env={**os.environ, "CUDA_VISIBLE_DEVICES": ""})
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Mpix/s" and line["higher_is_better"] is True
    assert line["config"]["workload"] == "objaverse_100k_sh3_512"
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["value"] == line["value"] > 0


def test_product_arm_refuses_to_run_without_a_gpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"], capture_output=True,
                         text=True, timeout=300, env={**os.environ, "CUDA_VISIBLE_DEVICES": ""})
    assert out.returncode != 0 and "no CPU path" in (out.stderr + out.stdout)
