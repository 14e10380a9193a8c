"""B200-native (sm_100a) drop-in for the rasteriser of Wenri/taichi_3d_gaussian_splatting.

Like the reference package, the operator is imported from its module:
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import \
        GaussianPointCloudRasterisation
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
Importing that module loads libgs_b200.so (hand-written CUDA kernels behind the C ABI of
include/gs_b200.h); it raises if the library is missing and cannot be built.

The steps either side of the operator follow the same rule:
`LossFunction`, `optim.FusedAdam`, `GaussianPointAdaptiveController.AdaptiveControllerAccumulators`
(CUDA, through the same library) and `GaussianPointCloudScene` (host-side wire formats).
"""
__all__ = ["Camera", "GaussianPointCloudRasterisation", "GaussianPointCloudScene", "GaussianPointAdaptiveController",
           "LossFunction", "optim", "distributed", "synthetic"]
