"""Acceptance of the drop-in by the reference's OWN callers, run unedited (north_star:
"GaussianPointTrainer, gaussian_point_render.py and the tests run unchanged").

Each case runs tests/ref_callers.py in a fresh process: the reference package installed under
baseline/_ref (pip --target by __graft_entry__.build(), git-ignored, shipped to the GPU box with
the snapshot), its caller files staged beside it, `install()` of the drop-in, and stand-ins only
for the third-party imports the box lacks (tests/golden/ti_shim).  Cases:

  * the operator cases of tests/GaussianPointCloudRasterisation_test.py (T_R.py:14-51, 111-351);
  * gaussian_point_render.py as __main__ (gaussian_point_render.py:100-176);
  * gaussian_point_train.py as __main__ -> GaussianPointCloudTrainer.train() on a small synthetic
    capture (GaussianPointTrainer.py:123-275, validation :342-423), with the trainer imported
    BEFORE install() (late install).
"""
import json
import os
import subprocess
import sys

import pytest

import ref_callers

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def _run(args, timeout):
    if not ref_callers.available():
        pytest.skip("baseline/_ref not staged: __graft_entry__.build() installs the reference there "
                    "when /root/reference exists")
    proc = subprocess.run([sys.executable, os.path.join(HERE, "ref_callers.py"), *args],
                          capture_output=True, text=True, timeout=timeout)
    assert proc.returncode == 0, proc.stderr[-6000:]
    line = [ln for ln in proc.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    print(json.dumps({k: v for k, v in out.items() if k != "report"}))
    return out


def test_reference_operator_tests_pass_unedited():
    out = _run(["tests"], timeout=1500)
    assert out["ran"] == len(ref_callers.DEFAULT_TESTS), out["report"]
    assert out["failures"] == 0 and out["errors"] == 0 and out["skipped"] == 0, out["report"]
    # test_rasterisation_basic: 100 fwd+bwd; two_points: 1 fwd; hook: 1+1; coverage: 10000 + 10000
    assert out["forward"] == 100 + 1 + 1 + 10000 and out["backward"] == 100 + 1 + 10000
    assert out["hook"] == 1


def test_reference_renderer_script_runs_unedited(tmp_path):
    out = _run(["render", str(tmp_path)], timeout=600)
    assert out["frames"] == 4 and out["shape"] == [544, 976, 3]
    assert out["rasteriser_is_drop_in"] and out["frame0_mean"] > 1.0
    assert out["frame0_max_level_diff"] <= 1 and out["frame0_differing_fraction"] < 1e-3
    assert out["forward"] == 4 + 1 and out["backward"] == 0


def test_reference_trainer_runs_unedited_with_late_install(tmp_path):
    iterations = 30
    out = _run(["train", str(tmp_path), str(iterations)], timeout=900)
    # every iteration went through the drop-in, with the densification hook; validation at 10 and 20
    assert out["backward"] == iterations and out["hook_calls"] == iterations
    assert out["forward"] == iterations + 2
    # the YAML values reached the operator through the reference's own nested config class
    r = out["rasteriser"]
    assert r["hook"] and (r["near"], r["far"], r["scale"]) == (0.4, 2000.0, 10.0)
    assert r["config_type"].endswith("GaussianPointCloudRasterisationConfig")
    assert out["last_loss"] < out["first_loss"], out["losses"]
    assert "best_scene.parquet" in out["checkpoints"] and "scene_20.parquet" in out["checkpoints"]
