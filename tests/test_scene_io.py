"""Scene wire formats: parquet and INRIA .ply, byte-exact.

Reference: GaussianPointCloudScene.py:132-210 (to_parquet / to_ply / from_parquet),
benchmark/inference_benchmark.py:21-81 (.ply import)."""
import os
import struct

import numpy as np
import pandas as pd
import pytest
import torch

from taichi_3d_gaussian_splatting_b200 import GaussianPointCloudScene as S

Scene = S.GaussianPointCloudScene


def _scene(n=37, ratio=None, seed=0):
    rng = np.random.default_rng(seed)
    xyz = rng.normal(0, 1, (n, 3)).astype(np.float32)
    feat = rng.normal(0, 1, (n, 56)).astype(np.float32)
    feat[:, 0:4] /= np.linalg.norm(feat[:, 0:4], axis=1, keepdims=True)
    return Scene(xyz, Scene.PointCloudSceneConfig(max_num_points_ratio=ratio), torch.from_numpy(feat)), xyz, feat


def test_parquet_columns_and_round_trip(tmp_path):
    scene, xyz, feat = _scene(ratio=1.5)
    assert scene.point_cloud.shape[0] == int(37 * 1.5) and int(scene.point_invalid_mask.sum()) == 55 - 37
    path = str(tmp_path / "scene.parquet")
    scene.to_parquet(path)
    df = pd.read_parquet(path)
    want = ["x", "y", "z"] + [f"cov_q{i}" for i in range(4)] + [f"cov_s{i}" for i in range(3)] + ["alpha0"] + \
        [f"r_sh{i}" for i in range(16)] + [f"g_sh{i}" for i in range(16)] + [f"b_sh{i}" for i in range(16)]
    assert list(df.columns) == want and len(df) == 37 and all(df.dtypes == np.float32)
    back = Scene.from_parquet(path)
    assert np.array_equal(back.point_cloud.detach().numpy(), xyz)
    assert np.array_equal(back.point_cloud_features.detach().numpy(), feat)
    assert back.point_invalid_mask.dtype == torch.int8 and back.point_object_id.dtype == torch.int32


def test_parquet_written_the_reference_way_loads(tmp_path):
    """A file produced by the statements of Sc.py:132-147 (pandas only) is read back bit-exactly."""
    _, xyz, feat = _scene(seed=3)
    cols = S.FEATURE_COLUMNS
    df = pd.concat([pd.DataFrame(xyz, columns=["x", "y", "z"]), pd.DataFrame(feat, columns=cols)], axis=1)
    path = str(tmp_path / "ref.parquet")
    df.to_parquet(path)
    scene = Scene.from_parquet(path)
    assert np.array_equal(scene.point_cloud_features.detach().numpy(), feat)


def test_raw_parquet_is_initialised(tmp_path):
    rng = np.random.default_rng(5)
    df = pd.DataFrame(rng.normal(0, 1, (200, 3)), columns=["x", "y", "z"])
    for c in "rgb":
        df[c] = rng.integers(1, 256, 200)   # 0 gives logit(0) = -inf, as in the reference
    path = str(tmp_path / "raw.parquet")
    df.to_parquet(path)
    scene = Scene.from_parquet(path)
    f = scene.point_cloud_features.detach().numpy()
    assert np.allclose(np.linalg.norm(f[:, 0:4], axis=1), 1, atol=1e-6)
    assert np.all(f[:, 7] == -2.0) and np.all(f[:, 9:24] == 0) and np.all(np.isfinite(f))
    assert np.all(f[:, 4] == f[:, 5]) and np.all(f[:, 5] == f[:, 6])       # isotropic
    rgb = np.clip(df[["r", "g", "b"]].to_numpy().astype(np.float32) / 255.0, 0, 0.99)
    assert np.allclose(f[:, 8], np.log(rgb[:, 0] / (1 - rgb[:, 0])) / 0.28209479177387814, rtol=1e-5, atol=1e-6)


def test_ply_export_layout(tmp_path):
    scene, xyz, feat = _scene()
    path = str(tmp_path / "scene.ply")
    scene.to_ply(path)
    raw = open(path, "rb").read()
    header, body = raw.split(b"end_header\n", 1)
    lines = header.decode().strip().split("\n")
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 37"]
    props = [ln.split() for ln in lines[3:]]
    assert all(p[:2] == ["property", "float"] for p in props)
    assert [p[2] for p in props] == S.PLY_PROPERTIES and len(props) == 62
    assert len(body) == 37 * 62 * 4
    row0 = struct.unpack("<62f", body[:62 * 4])
    assert row0[0:3] == tuple(xyz[0]) and row0[3:6] == (0.0, 0.0, 0.0)
    assert row0[6:9] == (feat[0, 8], feat[0, 24], feat[0, 40])            # f_dc: DC of r, g, b
    assert row0[9:24] == tuple(feat[0, 9:24]) and row0[24:39] == tuple(feat[0, 25:40])
    assert row0[54] == feat[0, 7] and row0[55:58] == tuple(feat[0, 4:7])
    assert row0[58:62] == (feat[0, 3], feat[0, 0], feat[0, 1], feat[0, 2])  # rotation stored wxyz


def test_ply_import_matches_the_benchmark_conversion(tmp_path):
    """A hand-packed INRIA file (unnormalised wxyz rotations) against the conversion steps of
    benchmark/inference_benchmark.py:21-81 written out in numpy."""
    rng = np.random.default_rng(9)
    n = 23
    table = rng.normal(0, 1, (n, 62)).astype("<f4")
    path = str(tmp_path / "inria.ply")
    with open(path, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n).encode())
        f.write("".join(f"property float {p}\n" for p in S.PLY_PROPERTIES).encode())
        f.write(b"end_header\n")
        f.write(table.tobytes())
    col = {p: table[:, i].astype(np.float64) for i, p in enumerate(S.PLY_PROPERTIES)}
    dc = np.stack([col[f"f_dc_{i}"] for i in range(3)], axis=1)
    extra = np.stack([col[f"f_rest_{i}"] for i in range(45)], axis=1).reshape(n, 3, 15)
    rots = np.roll(np.stack([col[f"rot_{i}"] for i in range(4)], axis=1), shift=-1, axis=1)
    rots = rots / np.linalg.norm(rots, axis=1, keepdims=True)
    scales = np.stack([col[f"scale_{i}"] for i in range(3)], axis=1)
    want = np.concatenate([rots, scales, col["opacity"][:, None], dc[:, 0:1], extra[:, 0], dc[:, 1:2], extra[:, 1],
                           dc[:, 2:3], extra[:, 2]], axis=1).astype(np.float32)
    xyz, feat = S.read_ply(path)
    assert np.array_equal(xyz, table[:, 0:3]) and np.array_equal(feat, want)
    scene = Scene.from_ply(path)
    assert scene.point_cloud_features.shape == (n, 56)


def test_ply_round_trip_and_errors(tmp_path):
    scene, xyz, feat = _scene(seed=4)
    path = str(tmp_path / "rt.ply")
    scene.to_ply(path)
    x2, f2 = S.read_ply(path)
    assert np.array_equal(x2, xyz)
    assert np.array_equal(f2[:, 4:], feat[:, 4:]) and np.allclose(f2[:, :4], feat[:, :4], atol=1e-7)
    bad = str(tmp_path / "bad.ply")
    open(bad, "wb").write(b"ply\nformat ascii 1.0\nelement vertex 0\nend_header\n")
    with pytest.raises(ValueError):
        S.read_ply(bad)
    open(bad, "wb").write(b"not a ply")
    with pytest.raises(ValueError):
        S.read_ply(bad)


@pytest.mark.gpu
def test_to_device_feeds_the_rasteriser(tmp_path):
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as R
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    sc = make_scene(2000, 64, 48, seed=41, scale_factor=0.05)
    scene = Scene(sc.point_cloud, Scene.PointCloudSceneConfig(), torch.from_numpy(sc.point_cloud_features))
    path = str(tmp_path / "s.parquet")
    scene.to_parquet(path)
    dev = torch.device("cuda")
    loaded = Scene.from_parquet(path).to_device(dev)
    assert loaded.point_cloud.is_cuda and loaded.point_invalid_mask.is_cuda and loaded.point_cloud.requires_grad
    t = lambda a: torch.from_numpy(a).to(dev)
    ras = R(R.GaussianPointCloudRasterisationConfig())
    image, _, _ = ras(R.GaussianPointCloudRasterisationInput(
        point_cloud=loaded.point_cloud, point_cloud_features=loaded.point_cloud_features,
        point_object_id=loaded.point_object_id, point_invalid_mask=loaded.point_invalid_mask,
        camera_info=CameraInfo(t(sc.camera_intrinsics), sc.height, sc.width, 0),
        q_pointcloud_camera=t(sc.q_pointcloud_camera), t_pointcloud_camera=t(sc.t_pointcloud_camera),
        color_max_sh_band=3))
    image.sum().backward()
    assert float(image.abs().sum()) > 0 and loaded.point_cloud.grad is not None
