"""Scene container and wire formats: the data either side of the hot
path.  Mirrors the parts of the reference `GaussianPointCloudScene`
(taichi_3d_gaussian_splatting/GaussianPointCloudScene.py, "Sc.py" below) that define bytes on disk
and tensors on the device:

  * parquet: columns x y z cov_q0..3 cov_s0..2 alpha0 r_sh0..15 g_sh0..15 b_sh0..15, one row per
    VALID point (Sc.py:132-147, 181-210);
  * INRIA .ply export (Sc.py:149-179): binary little-endian float32 vertex elements
    x y z nx ny nz f_dc_0..2 f_rest_0..44 opacity scale_0..2 rot_0..3, rotation stored wxyz;
  * INRIA .ply import (benchmark/inference_benchmark.py:21-81): rotation wxyz -> xyzw and
    normalised, SH regrouped to [DC, 15 rest] per colour channel;
  * the N x 3 / N x 56 / int8 mask / int32 object-id tensors the rasteriser consumes
    (Sc.py:24-66), uploaded through pinned memory by `to_device`.

plyfile is not a dependency: the PLY header is read and written here (the binary layout above is
fixed by the format, not by plyfile).  The k-nearest-neighbour initialisation of a raw point cloud
(Sc.py:71-127) is restated with scipy's cKDTree, as in the reference.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple, Union

import numpy as np
import torch
import torch.nn as nn

try:
    from dataclass_wizard import YAMLWizard
except Exception:  # pragma: no cover
    class YAMLWizard:  # type: ignore
        pass

FEATURE_COLUMNS = ([f"cov_q{i}" for i in range(4)] + [f"cov_s{i}" for i in range(3)] + ["alpha0"]
                   + [f"r_sh{i}" for i in range(16)] + [f"g_sh{i}" for i in range(16)]
                   + [f"b_sh{i}" for i in range(16)])  # Sc.py:137-142
PLY_PROPERTIES = (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)]
                  + [f"f_rest_{i}" for i in range(45)] + ["opacity"] + [f"scale_{i}" for i in range(3)]
                  + [f"rot_{i}" for i in range(4)])  # Sc.py:161-173

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


# ------------------------------------------------------------------------------------------------
# PLY
# ------------------------------------------------------------------------------------------------
def write_ply(path: str, point_cloud: np.ndarray, point_cloud_features: np.ndarray) -> None:
    """Valid points only, in the INRIA layout (Sc.py:149-179)."""
    xyz = np.asarray(point_cloud, dtype=np.float32)
    feat = np.asarray(point_cloud_features, dtype=np.float32)
    n = xyz.shape[0]
    f_sh = feat[:, 8:].reshape(n, 3, 16)
    rows = np.concatenate([xyz, np.zeros_like(xyz), f_sh[:, :, 0], f_sh[:, :, 1:].reshape(n, 45), feat[:, 7:8],
                           feat[:, 4:7], feat[:, [3, 0, 1, 2]]], axis=1).astype("<f4")
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n
    header += "".join(f"property float {name}\n" for name in PLY_PROPERTIES) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(rows).tobytes())


def _read_ply_vertices(path: str) -> np.ndarray:
    """Structured array of the `vertex` element of a binary little-endian PLY file."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.find(b"end_header\n")
    if not data.startswith(b"ply") or end < 0:
        raise ValueError(f"{path}: not a PLY file")
    lines = data[:end].decode("ascii").split("\n")
    offset = end + len(b"end_header\n")
    fmt = [ln for ln in lines if ln.startswith("format ")]
    if not fmt or fmt[0].split()[1] != "binary_little_endian":
        raise ValueError(f"{path}: only binary_little_endian PLY is supported")
    elements, current = [], None
    for ln in lines:
        tok = ln.split()
        if not tok:
            continue
        if tok[0] == "element":
            current = {"name": tok[1], "count": int(tok[2]), "props": []}
            elements.append(current)
        elif tok[0] == "property":
            if tok[1] == "list":
                raise ValueError(f"{path}: list properties are not supported")
            current["props"].append((tok[2], "<" + _PLY_TYPES[tok[1]]))
    for el in elements:
        dtype = np.dtype(el["props"])
        if el["name"] == "vertex":
            return np.frombuffer(data, dtype=dtype, count=el["count"], offset=offset)
        offset += dtype.itemsize * el["count"]
    raise ValueError(f"{path}: no vertex element")


def read_ply(path: str) -> Tuple[np.ndarray, np.ndarray]:
    """INRIA gaussian-splatting .ply -> (point_cloud [N,3] f32, point_cloud_features [N,56] f32),
    the conversion of benchmark/inference_benchmark.py:21-81."""
    v = _read_ply_vertices(path)
    names = v.dtype.names
    n = v.shape[0]
    xyz = np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32)
    rest_names = [p for p in names if p.startswith("f_rest_")]
    if len(rest_names) != 45:
        raise ValueError(f"{path}: expected 45 f_rest_* properties (SH degree 3), found {len(rest_names)}")
    dc = np.stack([v[f"f_dc_{i}"] for i in range(3)], axis=1).astype(np.float64)          # [N,3]
    rest = np.stack([v[p] for p in rest_names], axis=1).astype(np.float64).reshape(n, 3, 15)
    scales = np.stack([v[p] for p in names if p.startswith("scale_")], axis=1).astype(np.float64)
    rots = np.stack([v[p] for p in names if p.startswith("rot")], axis=1).astype(np.float64)
    rots = np.roll(rots, shift=-1, axis=1)                                                # wxyz -> xyzw
    rots = rots / np.linalg.norm(rots, axis=1, keepdims=True)
    opac = np.asarray(v["opacity"], dtype=np.float64)[:, None]
    feat = np.concatenate([rots, scales, opac, dc[:, 0:1], rest[:, 0], dc[:, 1:2], rest[:, 1], dc[:, 2:3],
                           rest[:, 2]], axis=1)
    return xyz, feat.astype(np.float32)


# ------------------------------------------------------------------------------------------------
# parquet
# ------------------------------------------------------------------------------------------------
def write_parquet(path: str, point_cloud: np.ndarray, point_cloud_features: np.ndarray) -> None:
    """Valid points only; column names and order of Sc.py:132-147."""
    import pandas as pd
    df = pd.concat([pd.DataFrame(np.asarray(point_cloud), columns=["x", "y", "z"]),
                    pd.DataFrame(np.asarray(point_cloud_features), columns=FEATURE_COLUMNS)], axis=1)
    df.to_parquet(path)


def read_parquet(path: str):
    """-> (point_cloud [N,3], features [N,56] or None, rgb [N,3] or None)  (Sc.py:181-210)."""
    import pandas as pd
    df = pd.read_parquet(path)
    xyz = df[["x", "y", "z"]].to_numpy()
    feat = df[FEATURE_COLUMNS].to_numpy() if set(FEATURE_COLUMNS).issubset(df.columns) else None
    rgb = df[["r", "g", "b"]].to_numpy() if {"r", "g", "b"}.issubset(df.columns) else None
    return xyz, feat, rgb


# ------------------------------------------------------------------------------------------------
# container
# ------------------------------------------------------------------------------------------------
class GaussianPointCloudScene(nn.Module):
    @dataclass
    class PointCloudSceneConfig(YAMLWizard):
        num_of_features: int = 56
        max_num_points_ratio: Optional[float] = None
        add_sphere: bool = False
        sphere_radius_factor: float = 4.0
        num_points_sphere: int = 10000
        max_initial_covariance: Optional[float] = None
        initial_alpha: float = -2.0
        initial_covariance_ratio: float = 1.0

    def __init__(self, point_cloud: Union[np.ndarray, torch.Tensor], config: "GaussianPointCloudScene.PointCloudSceneConfig",
                 point_cloud_features: Optional[torch.Tensor] = None, point_object_id: Optional[torch.Tensor] = None):
        super().__init__()
        if point_cloud.ndim != 2 or point_cloud.shape[1] != 3:
            raise ValueError("point_cloud must be [N, 3]")
        if isinstance(point_cloud, np.ndarray):
            point_cloud = torch.from_numpy(np.array(point_cloud, dtype=np.float32))  # owned, writable copy
        point_cloud = point_cloud.to(torch.float32)
        if point_cloud_features is not None:
            point_cloud_features = torch.as_tensor(point_cloud_features, dtype=torch.float32)
        num_points = point_cloud.shape[0]
        if config.max_num_points_ratio is not None:  # spare capacity for densification (Sc.py:39-47)
            capacity = int(num_points * config.max_num_points_ratio)
            if capacity <= num_points:
                raise ValueError("max_num_points_ratio should be greater than 1.0")
            point_cloud = torch.cat([point_cloud, torch.zeros((capacity - num_points, 3))], dim=0)
            if point_cloud_features is not None:
                point_cloud_features = torch.cat(
                    [point_cloud_features, torch.zeros((capacity - num_points, config.num_of_features))], dim=0)
        self.config = config
        self.point_cloud = nn.Parameter(point_cloud.contiguous())
        if point_cloud_features is None:
            point_cloud_features = torch.zeros(point_cloud.shape[0], config.num_of_features)
        self.point_cloud_features = nn.Parameter(point_cloud_features.contiguous())
        mask = torch.zeros(point_cloud.shape[0], dtype=torch.int8)
        mask[num_points:] = 1
        self.register_buffer("point_invalid_mask", mask)
        if point_object_id is None:
            point_object_id = torch.zeros(point_cloud.shape[0], dtype=torch.int32)
        self.register_buffer("point_object_id", point_object_id)

    def forward(self):
        return self.point_cloud, self.point_cloud_features

    def to_device(self, device) -> "GaussianPointCloudScene":
        """Uploads parameters and buffers through pinned host memory (one async copy each)."""
        device = torch.device(device)
        with torch.no_grad():
            for holder in (self._parameters, self._buffers):
                for name, t in list(holder.items()):
                    if t is None or t.device == device:
                        continue
                    src = t.data.pin_memory() if (device.type == "cuda" and t.device.type == "cpu") else t.data
                    moved = src.to(device, non_blocking=True)
                    holder[name] = nn.Parameter(moved, requires_grad=t.requires_grad) if isinstance(t, nn.Parameter) \
                        else moved
        if device.type == "cuda":
            torch.cuda.current_stream(device).synchronize()
        return self

    def _valid(self):
        keep = (self.point_invalid_mask == 0).cpu()
        return (self.point_cloud.detach().cpu()[keep].numpy(), self.point_cloud_features.detach().cpu()[keep].numpy())

    def to_parquet(self, path: str) -> None:
        write_parquet(path, *self._valid())

    def to_ply(self, path: str) -> None:
        write_ply(path, *self._valid())

    @staticmethod
    def from_ply(path: str, config: Optional["GaussianPointCloudScene.PointCloudSceneConfig"] = None):
        xyz, feat = read_ply(path)
        config = config or GaussianPointCloudScene.PointCloudSceneConfig()
        return GaussianPointCloudScene(xyz, config, point_cloud_features=torch.from_numpy(feat))

    @staticmethod
    def from_parquet(path: str, config: Optional["GaussianPointCloudScene.PointCloudSceneConfig"] = None):
        config = config or GaussianPointCloudScene.PointCloudSceneConfig()
        xyz, feat, rgb = read_parquet(path)
        if config.add_sphere:
            xyz, feat, rgb = _add_sphere(xyz, feat, rgb, config.sphere_radius_factor, config.num_points_sphere)
        if feat is None:
            scene = GaussianPointCloudScene(xyz, config)
            scene.initialize(point_cloud_rgb=rgb)
            return scene
        return GaussianPointCloudScene(xyz, config, point_cloud_features=torch.from_numpy(np.array(feat, dtype=np.float32)))

    def initialize(self, point_cloud_rgb=None) -> None:
        """Raw point cloud -> isotropic Gaussians (Sc.py:71-127): log-scale = log of the mean distance
        to the 3 nearest neighbours, random unit quaternion, opacity logit initial_alpha, DC colour
        from rgb (logit(rgb/255)/c0) or 1."""
        from scipy.spatial import cKDTree
        cfg = self.config
        with torch.no_grad():
            valid = self.point_invalid_mask == 0
            pts = self.point_cloud[valid].detach().cpu().numpy()
            dist, _ = cKDTree(pts).query(pts, k=4)
            sigma = np.clip(dist[:, 1:].mean(axis=1) * cfg.initial_covariance_ratio, 1e-6, cfg.max_initial_covariance)
            f = self.point_cloud_features
            f[valid, 4:7] = torch.tensor(np.log(sigma), dtype=torch.float32, device=f.device).unsqueeze(1)
            q = torch.rand_like(f[:, 0:4])
            f[:, 0:4] = q / torch.norm(q, dim=1, keepdim=True)
            f[:, 7] = cfg.initial_alpha
            f[:, 8:56] = 0.0
            f[:, 8] = f[:, 24] = f[:, 40] = 1.0
            if point_cloud_rgb is not None:
                rgb = torch.as_tensor(np.asarray(point_cloud_rgb), dtype=torch.float32, device=f.device) / 255.0
                rgb = torch.clamp(rgb, 0.0, 0.99)
                c0 = 0.28209479177387814
                for ch, col in enumerate((8, 24, 40)):
                    f[valid, col] = torch.log(rgb[:, ch] / (1.0 - rgb[:, ch])) / c0


def _add_sphere(xyz, feat, rgb, radius_factor: float, num_points: int):
    """Background shell of points (Sc.py:212-245); only meaningful for raw (feature-less) clouds."""
    extent = (xyz.max(axis=0) - xyz.min(axis=0)).max() / 2.0
    radius = extent * radius_factor
    phi = 2.0 * np.pi * np.random.rand(num_points)
    theta = np.arccos(2.0 * np.random.rand(num_points) - 1.0)
    shell = radius * np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta)], axis=1)
    xyz = np.concatenate([xyz, shell], axis=0)
    if rgb is not None:
        rgb = np.concatenate([rgb, np.full((num_points, 3), 255 // 2, dtype=rgb.dtype)], axis=0)
    if feat is not None:  # the reference would produce NaN rows here (concat of frames); keep rows aligned instead
        feat = np.concatenate([feat, np.full((num_points, feat.shape[1]), np.nan, dtype=feat.dtype)], axis=0)
    return xyz, feat, rgb
