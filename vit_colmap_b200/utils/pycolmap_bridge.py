"""Glue to pycolmap for the parts of the pipeline that stay with COLMAP (geometric verification,
incremental mapping).  pycolmap is absent from the build image, so everything here degrades to a
clear ImportError; nothing on the GPU hot path imports this module's pycolmap branch."""
from __future__ import annotations

try:  # pragma: no cover - absent in the build image
    import pycolmap
except Exception:  # noqa: BLE001
    pycolmap = None


def require(what: str):
    if pycolmap is None:
        raise ImportError(f"{what} needs pycolmap, which is not installed")
    return pycolmap


def matching_options(use_gpu: bool, num_threads: int, max_ratio: float, max_distance: float, cross_check: bool):
    """pycolmap's matching options for either binding generation: 3.13 nests the SIFT thresholds
    under FeatureMatchingOptions.sift, 3.12 has a flat SiftMatchingOptions."""
    pc = require("to_pycolmap_options()")
    flat = not hasattr(pc, "FeatureMatchingOptions")
    outer = pc.SiftMatchingOptions() if flat else pc.FeatureMatchingOptions()
    for name, value in (("use_gpu", use_gpu), ("num_threads", num_threads)):
        if not flat and hasattr(outer, name):
            setattr(outer, name, value)
    sift = outer if flat else outer.sift
    for name, value in (("max_ratio", max_ratio), ("max_distance", max_distance), ("cross_check", cross_check)):
        setattr(sift, name, value)
    return outer


def mapper_options(**fields):
    pc = require("3D reconstruction (pycolmap.incremental_mapping)")
    out = pc.IncrementalPipelineOptions()
    for name, value in fields.items():
        setattr(out, name, value)
    return out
