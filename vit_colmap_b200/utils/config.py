"""Pipeline configuration.  Field names, defaults and method names follow the reference's
``vit_colmap/utils/config.py`` (:9-198) because they ARE the interface (``Config.from_args``,
``config.matching.to_matching_options()``, ``config.extractor.extractor_type`` ...); the bodies are
this package's own: argparse mapping is table driven, camera defaults come from the extractor base
module, and pycolmap objects are only built in ``pycolmap_bridge`` (pycolmap is not in the image)."""
from __future__ import annotations

import logging
from dataclasses import dataclass, field
from typing import Optional

from . import pycolmap_bridge

pycolmap = pycolmap_bridge.pycolmap


@dataclass
class LogConfig:
    level: int = logging.INFO
    format: str = "[%(asctime)s][%(filename)s:%(lineno)d][%(levelname)s] %(message)s"
    datefmt: str = "%H:%M:%S"

    def apply(self):
        logging.basicConfig(force=True, **{k: getattr(self, k) for k in ("level", "format", "datefmt")})


@dataclass
class CameraConfig:
    model: str = "SIMPLE_PINHOLE"
    width: Optional[int] = None
    height: Optional[int] = None
    params: Optional[list[float]] = None

    def get_default_params(self, width: int, height: int) -> list[float]:
        """Explicit params win; otherwise f = max(w, h) and the principal point at the image centre
        (SIMPLE_PINHOLE / PINHOLE only -- anything else raises ValueError, as in the reference)."""
        if self.params is not None:
            return self.params
        from ..features.base_extractor import default_camera_params
        if self.model not in ("SIMPLE_PINHOLE", "PINHOLE"):
            raise ValueError(f"Unsupported camera model: {self.model}")
        return default_camera_params(self.model, width, height)


@dataclass
class MatchingConfig:
    use_gpu: bool = True
    max_ratio: float = 0.8
    max_distance: float = 0.7
    cross_check: bool = True
    num_threads: int = -1

    def _thresholds(self) -> dict:
        return dict(max_ratio=self.max_ratio, max_distance=self.max_distance, cross_check=self.cross_check)

    def to_matching_options(self):
        """What ``vit_colmap_b200.matching.match_exhaustive`` takes (reference: config.py:64-96 builds
        the pycolmap object; see to_pycolmap_options for that one)."""
        from ..matching import FeatureMatchingOptions
        return FeatureMatchingOptions(use_gpu=self.use_gpu, num_threads=self.num_threads, **self._thresholds())

    def to_pycolmap_options(self):
        return pycolmap_bridge.matching_options(self.use_gpu, self.num_threads, **self._thresholds())


@dataclass
class ReconstructionConfig:
    min_num_matches: int = 15
    multiple_models: bool = True

    def to_mapper_options(self):
        return pycolmap_bridge.mapper_options(min_num_matches=self.min_num_matches,
                                              multiple_models=self.multiple_models)


@dataclass
class ExtractorConfig:
    extractor_type: str = "vit"  # "vit" | "trainable_vit" | "dummy" | "colmap_sift"
    vit_weights_path: Optional[str] = None
    # not in the reference's dataclass: it hard-codes these where the extractor is built
    # (run_pipeline.py:329-339); same defaults
    model_name: str = "dinov2_vitb14"
    num_keypoints: int = 2048
    descriptor_dim: int = 128
    batch_size: int = 32


# argparse attribute -> (section, field, transform); first truthy source of a group wins
_ARG_TABLE = (
    (("camera_model",), "camera", "model", None, False),
    (("extractor",), "extractor", "extractor_type", None, True),
    (("vit_weights", "model"), "extractor", "vit_weights_path", str, True),
    (("use_gpu",), "matching", "use_gpu", None, False),
)


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
        """argparse namespace of ``run_pipeline.main`` -> Config (reference: config.py:152-181).
        ``--extractor`` beats the legacy ``--use-colmap-sift``; ``--vit-weights`` beats ``--model``."""
        cfg = cls()
        for sources, section, name, cast, need_truthy in _ARG_TABLE:
            for src in sources:
                if not hasattr(args, src):
                    continue
                value = getattr(args, src)
                if need_truthy and not value:
                    continue
                setattr(getattr(cfg, section), name, cast(value) if cast else value)
                break
        if not getattr(args, "extractor", None) and getattr(args, "use_colmap_sift", False):
            cfg.extractor.extractor_type = "colmap_sift"
        for flag, attr in (("skip_matching", "do_matching"), ("skip_reconstruction", "do_reconstruction")):
            if hasattr(args, flag):
                setattr(cfg, attr, not getattr(args, flag))
        if getattr(args, "verbose", False):
            cfg.log.level = logging.DEBUG
            cfg.log.apply()
        return cfg

    def summary(self) -> str:
        rows = (("Extractor", self.extractor.extractor_type), ("Camera model", self.camera.model),
                ("Matching", self.do_matching), ("Reconstruction", self.do_reconstruction),
                ("GPU matching", self.matching.use_gpu), ("Min matches", self.reconstruction.min_num_matches))
        text = ["Configuration:"]
        for label, value in rows:
            if isinstance(value, bool):
                value = "enabled" if value else "disabled"
            text.append(f"  {label}: {value}")
        return "\n".join(text)
