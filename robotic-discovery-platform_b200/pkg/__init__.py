"""Drop-in mirror of the reference's `pkg` package for the vision hot path.

Put `robotic-discovery-platform_b200/` on `sys.path` ahead of the reference checkout and
`services/vision_analysis/server.py` imports these modules instead of its own (`server.py:47-48`).
Only the two hot-path modules are mirrored: `pkg.segmentation_model` and `pkg.geometry_utils`.
"""
