"""Host side of the B200-native vision hot path (U-Net forward + curvature profile).

`rdp_b200._lib` binds the C ABI in `include/rdp_b200.h` with ctypes; `rdp_b200.unet` and `rdp_b200.geometry` are the
batched device-tensor entry points; the top-level `pkg` package next to this one mirrors the reference's module paths
(`pkg.segmentation_model.UNet`, `pkg.geometry_utils.compute_curvature_profile`) so its gRPC server runs unchanged.
PyTorch is used for device memory and streams only.  There is no CPU fallback: without the CUDA library or a GPU the
compute entry points raise.
"""
