"""Seeded synthetic scenes for the parity tests and bench.py.

Everything is generated on the CPU with a seeded numpy Generator so the oracle and
the CUDA path see identical bits.  No file or network access.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class SyntheticScene:
    point_cloud: np.ndarray           # (N,3) f32
    point_cloud_features: np.ndarray  # (N,56) f32, layout of reference R.py:208-236
    point_invalid_mask: np.ndarray    # (N,) i8
    point_object_id: np.ndarray       # (N,) i32
    camera_intrinsics: np.ndarray     # (3,3) f32
    q_pointcloud_camera: np.ndarray   # (K,4) f32 xyzw
    t_pointcloud_camera: np.ndarray   # (K,3) f32
    width: int
    height: int


def make_scene(num_points: int, width: int, height: int, seed: int = 1000,
               scale_factor: float = 0.004, num_objects: int = 1,
               invalid_fraction: float = 0.02, z_range=(1.0, 50.0)) -> SyntheticScene:
    """Random Gaussians seen by a pinhole camera at the origin looking down +z.

    Pixel position uniform over the image plus a 5 % margin, depth log-uniform in
    `z_range`, log-scales ~ N(log(scale_factor*z), 0.5), un-normalised quaternions,
    opacity logit ~ N(0,1.5), SH DC ~ N(0,1), higher bands ~ N(0,0.1).
    With `num_objects` > 1 each object gets its own small rigid pose.
    """
    assert width % 16 == 0 and height % 16 == 0
    rng = np.random.default_rng(seed)
    n = num_points
    f = 0.6 * width
    K = np.array([[f, 0, width / 2], [0, f, height / 2], [0, 0, 1]], np.float32)
    u = rng.uniform(-0.05 * width, 1.05 * width, n)
    v = rng.uniform(-0.05 * height, 1.05 * height, n)
    z = np.exp(rng.uniform(np.log(z_range[0]), np.log(z_range[1]), n))
    xyz = np.stack([(u - K[0, 2]) * z / f, (v - K[1, 2]) * z / f, z], 1).astype(np.float32)
    feat = np.zeros((n, 56), np.float32)
    feat[:, 0:4] = rng.normal(0, 1, (n, 4))
    feat[:, 4:7] = rng.normal(np.log(scale_factor * z)[:, None], 0.5, (n, 3))
    feat[:, 7] = rng.normal(0, 1.5, n)
    feat[:, 8:56] = rng.normal(0, 0.1, (n, 48))
    for c in (8, 24, 40):
        feat[:, c] = rng.normal(0, 1, n)
    invalid = (rng.uniform(0, 1, n) < invalid_fraction).astype(np.int8)
    obj = rng.integers(0, num_objects, n).astype(np.int32)
    q = np.zeros((num_objects, 4), np.float32)
    q[:, 3] = 1.0
    t = np.zeros((num_objects, 3), np.float32)
    if num_objects > 1:
        q[1:, :3] = rng.normal(0, 0.02, (num_objects - 1, 3))
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        t[1:] = rng.normal(0, 0.05, (num_objects - 1, 3))
    return SyntheticScene(xyz, feat, invalid, obj, K, q.astype(np.float32), t.astype(np.float32),
                          width, height)
