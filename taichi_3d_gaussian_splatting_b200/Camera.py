"""Camera records consumed by the rasteriser (same fields as the reference's
taichi_3d_gaussian_splatting/Camera.py:6-11; any object with these attributes works)."""
from dataclasses import dataclass

import torch


@dataclass
class CameraInfo:
    camera_intrinsics: torch.Tensor  # 3x3 matrix
    camera_height: int  # height of the image
    camera_width: int  # width of the image
    camera_id: int = 0  # camera id
