"""``colmap_sift`` switch value (reference: vit_colmap/features/colmap_sift_extractor.py:9-43): a
pass-through to pycolmap's own SIFT extraction.  SIFT itself is outside the accelerated path."""
from pathlib import Path
from typing import Optional

from .base_extractor import BaseExtractor


class ColmapSiftExtractor(BaseExtractor):
    def __init__(self):
        pass

    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:
        try:
            import pycolmap
        except ImportError as e:
            raise ImportError("extractor_type='colmap_sift' delegates to pycolmap.extract_features; "
                              "pycolmap is not installed") from e
        reader = pycolmap.ImageReaderOptions()
        reader.camera_model = camera_model
        if camera_params is not None:
            reader.camera_params = ",".join(str(p) for p in camera_params)
        pycolmap.extract_features(database_path=str(db_path), image_path=str(image_dir), reader_options=reader)
