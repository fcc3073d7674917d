"""Extractor plugins behind the reference's ``extractor_type`` switch
(vit_colmap/pipeline/run_pipeline.py:319-339)."""
from .base_extractor import BaseExtractor  # noqa: F401
from .colmap_sift_extractor import ColmapSiftExtractor  # noqa: F401
from .dummy_extractor import DummyExtractor  # noqa: F401


def __getattr__(name):  # the GPU extractors import torch lazily
    if name == "ViTExtractor":
        from .vit_extractor import ViTExtractor
        return ViTExtractor
    if name == "TrainableViTExtractor":
        from .trainable_vit_extractor import TrainableViTExtractor
        return TrainableViTExtractor
    if name == "HybridViTExtractor":
        from .hybrid_extractor import HybridViTExtractor
        return HybridViTExtractor
    raise AttributeError(name)
