"""Configuration dataclasses of the pipeline: same fields, defaults and methods as the reference's
``vit_colmap/utils/config.py`` (:9-198).  Where the reference hands pycolmap option objects out,
this mirror returns pycolmap's when pycolmap is importable and the equivalent plain dataclass of
``vit_colmap_b200.matching`` otherwise (pycolmap is not installed in the build image)."""
from __future__ import annotations

import logging
from dataclasses import dataclass, field
from typing import Optional

try:  # pragma: no cover - absent in the build image
    import pycolmap
except Exception:  # noqa: BLE001
    pycolmap = None


@dataclass
class LogConfig:
    level: int = logging.INFO
    format: str = "[%(asctime)s][%(filename)s:%(lineno)d][%(levelname)s] %(message)s"
    datefmt: str = "%H:%M:%S"

    def apply(self):
        logging.basicConfig(level=self.level, format=self.format, datefmt=self.datefmt, force=True)


@dataclass
class CameraConfig:
    model: str = "SIMPLE_PINHOLE"
    width: Optional[int] = None
    height: Optional[int] = None
    params: Optional[list[float]] = None

    def get_default_params(self, width: int, height: int) -> list[float]:
        if self.params is not None:
            return self.params
        f = max(width, height)
        if self.model == "SIMPLE_PINHOLE":
            return [f, width / 2.0, height / 2.0]
        if self.model == "PINHOLE":
            return [f, f, width / 2.0, height / 2.0]
        raise ValueError(f"Unsupported camera model: {self.model}")


@dataclass
class MatchingConfig:
    use_gpu: bool = True
    max_ratio: float = 0.8
    max_distance: float = 0.7
    cross_check: bool = True
    num_threads: int = -1

    def to_matching_options(self):
        """Options object for ``match_exhaustive`` (config.py:64-96 in the reference)."""
        from ..matching import FeatureMatchingOptions
        return FeatureMatchingOptions(use_gpu=self.use_gpu, num_threads=self.num_threads, max_ratio=self.max_ratio,
                                      max_distance=self.max_distance, cross_check=self.cross_check)

    def to_pycolmap_options(self):
        """The pycolmap object the reference builds; only available with pycolmap installed."""
        if pycolmap is None:
            raise ImportError("pycolmap is not installed")
        if hasattr(pycolmap, "FeatureMatchingOptions"):
            opts = pycolmap.FeatureMatchingOptions()
            if hasattr(opts, "use_gpu"):
                opts.use_gpu = self.use_gpu
            if hasattr(opts, "num_threads"):
                opts.num_threads = self.num_threads
            target = opts.sift
        else:
            opts = target = pycolmap.SiftMatchingOptions()
        target.max_ratio, target.max_distance, target.cross_check = self.max_ratio, self.max_distance, self.cross_check
        return opts


@dataclass
class ReconstructionConfig:
    min_num_matches: int = 15
    multiple_models: bool = True

    def to_mapper_options(self):
        if pycolmap is None:
            raise ImportError("3D reconstruction is delegated to pycolmap.incremental_mapping, which is not installed")
        opts = pycolmap.IncrementalPipelineOptions()
        opts.min_num_matches = self.min_num_matches
        opts.multiple_models = self.multiple_models
        return opts


@dataclass
class ExtractorConfig:
    extractor_type: str = "vit"  # "vit" | "trainable_vit" | "dummy" | "colmap_sift"
    vit_weights_path: Optional[str] = None
    # extensions (reference hard-codes these at the call site, run_pipeline.py:329-339)
    model_name: str = "dinov2_vitb14"
    num_keypoints: int = 2048
    descriptor_dim: int = 128
    batch_size: int = 32


@dataclass
class Config:
    log: LogConfig = field(default_factory=LogConfig)
    camera: CameraConfig = field(default_factory=CameraConfig)
    extractor: ExtractorConfig = field(default_factory=ExtractorConfig)
    matching: MatchingConfig = field(default_factory=MatchingConfig)
    reconstruction: ReconstructionConfig = field(default_factory=ReconstructionConfig)
    do_matching: bool = True
    do_reconstruction: bool = True

    def __post_init__(self):
        self.log.apply()

    @classmethod
    def from_args(cls, args):
        config = cls()
        if hasattr(args, "camera_model"):
            config.camera.model = args.camera_model
        if getattr(args, "extractor", None):
            config.extractor.extractor_type = args.extractor
        elif getattr(args, "use_colmap_sift", False):
            config.extractor.extractor_type = "colmap_sift"
        if getattr(args, "vit_weights", None):
            config.extractor.vit_weights_path = str(args.vit_weights)
        elif getattr(args, "model", None):
            config.extractor.vit_weights_path = str(args.model)
        if hasattr(args, "use_gpu"):
            config.matching.use_gpu = args.use_gpu
        if hasattr(args, "skip_matching"):
            config.do_matching = not args.skip_matching
        if hasattr(args, "skip_reconstruction"):
            config.do_reconstruction = not args.skip_reconstruction
        if getattr(args, "verbose", False):
            config.log.level = logging.DEBUG
            config.log.apply()
        return config

    def summary(self) -> str:
        onoff = lambda b: "enabled" if b else "disabled"  # noqa: E731
        return "\n".join([
            "Configuration:",
            f"  Extractor: {self.extractor.extractor_type}",
            f"  Camera model: {self.camera.model}",
            f"  Matching: {onoff(self.do_matching)}",
            f"  Reconstruction: {onoff(self.do_reconstruction)}",
            f"  GPU matching: {onoff(self.matching.use_gpu)}",
            f"  Min matches: {self.reconstruction.min_num_matches}",
        ])
