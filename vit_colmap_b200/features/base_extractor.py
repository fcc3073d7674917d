"""Extractor plugin interface (reference: vit_colmap/features/base_extractor.py:6-16)."""
from abc import ABC, abstractmethod
from pathlib import Path
from typing import Optional

IMAGE_EXTENSIONS = {".jpg", ".jpeg", ".png", ".bmp", ".tiff", ".tif"}


def list_images(image_dir: Path):
    """Sorted image files of a directory, the discovery rule every reference extractor uses
    (vit_extractor.py:684-687, dummy_extractor.py:39-42)."""
    image_dir = Path(image_dir)
    if not image_dir.exists():
        return []
    return sorted(f for f in image_dir.iterdir() if f.suffix.lower() in IMAGE_EXTENSIONS)


def default_camera_params(camera_model: str, width: int, height: int) -> list:
    """vit_extractor.py:703-716 / dummy_extractor.py:67-76."""
    f = max(width, height)
    if camera_model == "SIMPLE_PINHOLE":
        return [f, width / 2.0, height / 2.0]
    if camera_model == "PINHOLE":
        return [f, f, width / 2.0, height / 2.0]
    raise ValueError(f"Unsupported camera model: {camera_model}")


class BaseExtractor(ABC):
    @abstractmethod
    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:
        """Extract features for every image in ``image_dir`` and write them to the COLMAP database."""
        raise NotImplementedError
