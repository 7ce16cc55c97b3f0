"""bench.py contract checks that need no GPU: the reference arm prints one JSON line with the agreed keys,
non-zero ranks of a torchrun launch stay silent, and the algorithmic-work figures match SURVEY 8(d)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_reference_arm_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, cwd=ROOT, timeout=560)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["impl"] == "reference" and d["metric"] == "MMD train-steps/sec (n=1500)" and d["unit"] == "train-steps/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"]
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, cwd=ROOT, env=env, timeout=60)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_algorithmic_work_matches_survey():
    sys.path.insert(0, ROOT)
    import bench
    exps, flops = bench.algorithmic_work(1, 1500, 1, 0)
    P = 3374250
    assert exps == 7 * P + 7 * (1500 * 1501 // 2)                     # one train step + the once-per-instance TT block
    assert abs(flops - 132.7e6) < 0.1e6                               # 132.7 MFLOP per train step
    _, f_eval = bench.algorithmic_work(1, 1500, 0, 1)
    assert abs(f_eval - 67.5e6) < 0.1e6


def test_executed_roofline_is_recomputable_from_the_tracked_profile():
    """roofline.frac = max(executed ex2/s / MUFU peak, FMA-pipe lane-slots/s / (FP32 peak / 2)) with the per-evaluation counts
    of profiles/r2_roofline_*.json; the profile carries the hash of the kernel sources and a mismatch is flagged."""
    sys.path.insert(0, ROOT)
    import bench
    from cgnn_b200.build import source_hash
    prof = bench.load_profile("pairwise")
    assert prof is not None and prof["per_eval"]["mufu_lane_ops"] > 0
    assert prof["stale"] == (prof["source_hash"] != source_hash("pairwise"))
    assert not prof["stale"], "cgnn_b200/csrc changed after profiles/r2_roofline_pairwise.json was captured: re-run the ncu capture " \
                              "(tools/profile_target.py under ncu --set full, then tools/ncu_summary.py --json)"
    evals_per_s, mufu_peak, fp32_peak = 1.4e12, 4.63e12, 73.5e12
    ex = bench.executed_roofline(prof, evals_per_s, mufu_peak, fp32_peak)
    assert abs(ex["mufu_frac"] - prof["per_eval"]["mufu_lane_ops"] * evals_per_s / mufu_peak) < 1e-12
    assert abs(ex["fma_pipe_frac"] - prof["per_eval"]["fma_pipe_lane_slots"] * evals_per_s / (fp32_peak / 2)) < 1e-12
    assert 0 < ex["mufu_frac"] < 1 and 0 < ex["fma_pipe_frac"] < 1
    w = prof["warp_inst"]                      # slots = (scalar + 2 x packed) x 32 / evaluations, as the note in the file says
    evals = prof["workload"]["evals_per_launch"]
    assert abs(prof["per_eval"]["fma_pipe_lane_slots"] - (w["fma_pipe_scalar"] + 2 * w["fma_pipe_packed_f32x2"]) * 32.0 / evals) < 1e-9
    for which in ("graph_c2", "graph_c5"):
        g = bench.load_profile(which)
        assert g is not None and g["stale"] == (g["source_hash"] != source_hash("graph"))
        assert not g["stale"], "graph kernel sources changed after profiles/%s was captured" % bench.PROFILE_FILES[which]
