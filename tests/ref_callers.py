#!/usr/bin/env python
"""Runs the REFERENCE'S OWN callers of the operator, unedited, against the drop-in.

    python tests/ref_callers.py tests  [TestClass.test_name ...]
    python tests/ref_callers.py render <workdir>
    python tests/ref_callers.py train  <workdir> [iterations]

north_star: "GaussianPointTrainer, gaussian_point_render.py and the tests run unchanged".  The files
executed here are the reference's, byte for byte: the package installed by `pip --target` under
baseline/_ref and the three caller files staged beside it by `__graft_entry__.build()`
(baseline/_ref/_callers; git-ignored, never part of this repository's history).  What this script
adds is only what the GPU box lacks: `tests/golden/ti_shim` on sys.path (stand-ins for `taichi`,
`dataclass_wizard`, `pytorch_msssim`, `plyfile`, `matplotlib`, `pytorch3d`, none of which is
installable there) and ONE call, `install()`, of the drop-in.  Each mode prints one JSON line;
tests/test_reference_callers_gpu.py runs every mode in its own process and checks that line.

Test infrastructure only: nothing in the product package imports this file.
"""
from __future__ import annotations

import importlib.util
import io
import json
import os
import runpy
import sys
import time
import unittest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SHIM = os.path.join(HERE, "golden", "ti_shim")
REF_INSTALL = os.path.join(ROOT, "baseline", "_ref")
REF_CALLERS = os.path.join(REF_INSTALL, "_callers")

# the cases of tests/GaussianPointCloudRasterisation_test.py that go through the operator boundary
# (the other classes of that file exercise the reference's own Taichi device functions, which are not
# replaced: GaussianPoint3D, the torch single-point forward)
DEFAULT_TESTS = (
    "TestFindTileStartAndEnd.test_find_tile_start_and_end",
    "TestRasterisation.test_rasterisation_basic",
    "TestRasterisation.test_rasterisation_two_points",
    "TestRasterisation.test_backward_hook",
    "TestRasterisation.test_backward_coverage",
)


def available() -> bool:
    return (os.path.exists(os.path.join(REF_INSTALL, "taichi_3d_gaussian_splatting", "__init__.py"))
            and os.path.exists(os.path.join(REF_CALLERS, "GaussianPointCloudRasterisation_test.py")))


def _paths() -> None:
    for p in (REF_INSTALL, SHIM, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)


def _drop_in():
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as drop_in
    return drop_in


def _counts(drop_in) -> dict:
    return dict(drop_in.CALL_COUNTS)


def run_tests(names) -> dict:
    """install() FIRST, then import the reference's test module (its from-import binds the drop-in)."""
    _paths()
    drop_in = _drop_in()
    drop_in.install()
    spec = importlib.util.spec_from_file_location(
        "reference_rasterisation_test", os.path.join(REF_CALLERS, "GaussianPointCloudRasterisation_test.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    assert mod.GaussianPointCloudRasterisation is drop_in.GaussianPointCloudRasterisation
    assert mod.find_tile_start_and_end is drop_in.find_tile_start_and_end
    suite = unittest.TestSuite()
    for name in names:
        cls, meth = name.split(".")
        suite.addTest(getattr(mod, cls)(meth))
    stream = io.StringIO()
    t0 = time.time()
    # the reference's loops print a loss per iteration: keep stdout for the JSON line
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        result = unittest.TextTestRunner(stream=stream, verbosity=2).run(suite)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
    return dict(mode="tests", ran=result.testsRun, failures=len(result.failures), errors=len(result.errors),
                skipped=len(result.skipped), names=list(names), seconds=round(time.time() - t0, 1),
                report=stream.getvalue()[-4000:], **_counts(drop_in))


def _synthetic_scene_parquet(path, num_points, width, height, seed, with_features=True):
    import numpy as np
    import pandas as pd
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    scene = make_scene(num_points, width, height, seed=seed, scale_factor=0.01, invalid_fraction=0.0)
    cols = {"x": scene.point_cloud[:, 0], "y": scene.point_cloud[:, 1], "z": scene.point_cloud[:, 2]}
    if with_features:
        names = [f"cov_q{i}" for i in range(4)] + [f"cov_s{i}" for i in range(3)] + ["alpha0"] + \
                [f"{c}_sh{i}" for c in "rgb" for i in range(16)]
        for j, name in enumerate(names):
            cols[name] = scene.point_cloud_features[:, j]
    else:
        rng = np.random.default_rng(seed)
        for c in "rgb":
            cols[c] = rng.integers(30, 220, num_points).astype(np.float64)
    pd.DataFrame(cols).to_parquet(path)
    return scene


def _poses(n, seed=11):
    """[n, 4, 4] T_pointcloud_camera close to the identity (camera looks down +z at the cloud)."""
    import numpy as np
    import torch
    rng = np.random.default_rng(seed)
    T = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    for i in range(1, n):
        a = rng.normal(0, 0.03, 3)
        K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        R = np.eye(3) + K + 0.5 * K @ K
        u, _, vt = np.linalg.svd(R)
        T[i, :3, :3] = (u @ vt).astype(np.float32)
        T[i, :3, 3] = rng.normal(0, 0.1, 3)
    return torch.from_numpy(T)


def run_render(workdir) -> dict:
    """gaussian_point_render.py as __main__ (its own argparse, scene loader, PNG writer)."""
    import numpy as np
    import torch
    from PIL import Image
    _paths()
    os.makedirs(workdir, exist_ok=True)
    drop_in = _drop_in()
    drop_in.install()
    parquet = os.path.join(workdir, "scene.parquet")
    width, height = 976, 544                     # the script's default resolution and intrinsics
    _synthetic_scene_parquet(parquet, 60000, width, height, seed=31)
    poses = _poses(4)
    torch.save(poses, os.path.join(workdir, "poses.pt"))
    out = os.path.join(workdir, "frames")
    argv = sys.argv
    sys.argv = ["gaussian_point_render.py", "--parquet_path", parquet, "--poses",
                os.path.join(workdir, "poses.pt"), "--output_prefix", out]
    t0 = time.time()
    try:
        glb = runpy.run_path(os.path.join(REF_CALLERS, "gaussian_point_render.py"), run_name="__main__")
    finally:
        sys.argv = argv
    seconds = time.time() - t0
    frames = sorted(f for f in os.listdir(out) if f.endswith(".png"))
    # the frame the script wrote equals a direct call of the operator on the same inputs
    renderer = glb["renderer"]
    Ras = drop_in.GaussianPointCloudRasterisation
    q, t = glb["SE3_to_quaternion_and_translation_torch"](renderer.cameras[0:1])
    with torch.no_grad():
        image, _, _ = Ras(Ras.GaussianPointCloudRasterisationConfig(near_plane=0.8, far_plane=1000.,
                                                                    depth_to_sort_key_scale=100.))(
            Ras.GaussianPointCloudRasterisationInput(
                point_cloud=renderer.scene.point_cloud, point_cloud_features=renderer.scene.point_cloud_features,
                point_invalid_mask=renderer.scene.point_invalid_mask,
                point_object_id=renderer.scene.point_object_id, camera_info=renderer.camera_info,
                q_pointcloud_camera=q, t_pointcloud_camera=t, color_max_sh_band=3))
    want = torch.clamp(image * 255, 0, 255).byte().cpu().numpy()
    got = np.asarray(Image.open(os.path.join(out, frames[0])).convert("RGB"))
    # (the script's first forward normalised the quaternions in place, R.py:264-266, so this second
    # render starts from features that differ in their last bits: compare at one grey level)
    diff = np.abs(want.astype(np.int16) - got.astype(np.int16))
    return dict(mode="render", frames=len(frames), shape=list(got.shape), seconds=round(seconds, 1),
                frame0_max_level_diff=int(diff.max()), frame0_differing_fraction=float((diff > 0).mean()),
                frame0_mean=float(got.mean()),
                rasteriser_is_drop_in=type(renderer.rasteriser) is Ras, **_counts(drop_in))


def run_train(workdir, iterations=30) -> dict:
    """gaussian_point_train.py as __main__: YAML config -> GaussianPointCloudTrainer.train().

    The trainer module is imported BEFORE install() on purpose (late install: its
    `from .GaussianPointCloudRasterisation import GaussianPointCloudRasterisation` and the
    `default_factory` of TrainConfig.rasterisation_config are already bound)."""
    import numpy as np
    import torch
    import yaml
    from PIL import Image
    _paths()
    os.makedirs(workdir, exist_ok=True)
    import taichi_3d_gaussian_splatting.GaussianPointTrainer as trainer_mod   # the reference's, first
    drop_in = _drop_in()
    drop_in.install()
    Ras = drop_in.GaussianPointCloudRasterisation
    assert trainer_mod.GaussianPointCloudRasterisation is Ras

    # a small synthetic capture: ground-truth images rendered from a known cloud, initial cloud = the
    # same positions with colours only (features initialised by the reference's scene loader)
    width, height, n = 256, 160, 6000
    truth = _synthetic_scene_parquet(os.path.join(workdir, "truth.parquet"), n, width, height, seed=41)
    _synthetic_scene_parquet(os.path.join(workdir, "init.parquet"), n, width, height, seed=41,
                             with_features=False)
    dev = torch.device("cuda:0")
    poses = _poses(3, seed=12)
    from taichi_3d_gaussian_splatting.utils import SE3_to_quaternion_and_translation_torch
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    K = torch.from_numpy(truth.camera_intrinsics)
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
    records = []
    for i in range(poses.shape[0]):
        q, t = SE3_to_quaternion_and_translation_torch(poses[i:i + 1])
        with torch.no_grad():
            image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                point_cloud=torch.from_numpy(truth.point_cloud).to(dev),
                point_cloud_features=torch.from_numpy(truth.point_cloud_features.copy()).to(dev),
                point_object_id=torch.zeros(n, dtype=torch.int32, device=dev),
                point_invalid_mask=torch.zeros(n, dtype=torch.int8, device=dev),
                camera_info=CameraInfo(K.to(dev), height, width, 0),
                q_pointcloud_camera=q.to(dev), t_pointcloud_camera=t.to(dev), color_max_sh_band=3))
        path = os.path.join(workdir, f"image_{i}.png")
        Image.fromarray(torch.clamp(image * 255, 0, 255).byte().cpu().numpy(), "RGB").save(path)
        records.append(dict(image_path=path, T_pointcloud_camera=poses[i].tolist(),
                            camera_intrinsics=K.tolist(), camera_height=height, camera_width=width,
                            camera_id=i))
    for name, recs in (("train.json", records[:2]), ("val.json", records[2:])):
        with open(os.path.join(workdir, name), "w") as f:
            json.dump(recs, f)
    before = _counts(drop_in)
    cfg = {
        "train-dataset-json-path": os.path.join(workdir, "train.json"),
        "val-dataset-json-path": os.path.join(workdir, "val.json"),
        "pointcloud-parquet-path": os.path.join(workdir, "init.parquet"),
        "num-iterations": int(iterations), "val-interval": 10, "log-loss-interval": 5,
        "log-metrics-interval": 10, "log-image-interval": 20, "initial-downsample-factor": 1,
        "increase-color-max-sh-band-interval": 10.0,
        "summary-writer-log-dir": os.path.join(workdir, "logs"),
        "output-model-dir": os.path.join(workdir, "logs"),
        # non-default rasteriser values, to see them arrive through the YAML -> nested dataclass path
        "rasterisation-config": {"near-plane": 0.4, "far-plane": 2000.0, "depth-to-sort-key-scale": 10.0},
        # densification POLICY is out of scope (AC.py:168-347); its statistics hook runs every step
        "adaptive-controller-config": {"num-iterations-warm-up": 100000},
        "gaussian-point-cloud-scene-config": {"max-num-points-ratio": 2.0},
    }
    cfg_path = os.path.join(workdir, "train.yaml")
    with open(cfg_path, "w") as f:
        yaml.safe_dump(cfg, f)

    losses = []
    orig_add_scalar = trainer_mod.SummaryWriter.add_scalar

    def add_scalar(self, tag, value, step=None, *a, **k):
        if tag == "train/loss":
            losses.append((int(step), float(value)))
        return orig_add_scalar(self, tag, value, step, *a, **k)

    trainer_mod.SummaryWriter.add_scalar = add_scalar
    seen = {}
    orig_init = Ras.__init__

    def spy_init(self, config, backward_valid_point_hook=None):
        seen.update(config_type=type(config).__qualname__, near=config.near_plane, far=config.far_plane,
                    scale=config.depth_to_sort_key_scale, hook=backward_valid_point_hook is not None)
        orig_init(self, config, backward_valid_point_hook)

    Ras.__init__ = spy_init
    argv = sys.argv
    sys.argv = ["gaussian_point_train.py", "--train_config", cfg_path]
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    t0 = time.time()
    try:
        runpy.run_path(os.path.join(REF_CALLERS, "gaussian_point_train.py"), run_name="__main__")
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        sys.argv = argv
        Ras.__init__ = orig_init
        trainer_mod.SummaryWriter.add_scalar = orig_add_scalar
    after = _counts(drop_in)
    logs = os.listdir(os.path.join(workdir, "logs"))
    return dict(mode="train", iterations=int(iterations), seconds=round(time.time() - t0, 1),
                losses=losses, first_loss=losses[0][1], last_loss=losses[-1][1],
                checkpoints=sorted(f for f in logs if f.endswith(".parquet")), rasteriser=seen,
                forward=after["forward"] - before["forward"], backward=after["backward"] - before["backward"],
                hook_calls=after["hook"] - before["hook"])


def main(argv):
    mode = argv[1] if len(argv) > 1 else "tests"
    if not available():
        print(json.dumps(dict(mode=mode, unavailable="baseline/_ref is not staged (run __graft_entry__.build() "
                                                       "where /root/reference exists)")))
        return 0
    if mode == "tests":
        out = run_tests(argv[2:] or DEFAULT_TESTS)
    elif mode == "render":
        out = run_render(argv[2])
    elif mode == "train":
        out = run_train(argv[2], int(argv[3]) if len(argv) > 3 else 30)
    else:
        raise SystemExit(f"unknown mode {mode}")
    print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
