"""Fake backend of the smoke test (reference: vit_colmap/features/dummy_extractor.py:8-117).

Pure host code by contract: grid keypoints every ``step`` pixels and a 128-byte descriptor per grid
cell drawn from NumPy's legacy MT19937 stream seeded with ``seed + 1000*gx + gy``, so equal cells get
equal descriptors in every image.  Nothing here is a kernel: the RandomState byte stream IS the
interface the smoke test relies on."""
from pathlib import Path
from typing import Optional

import numpy as np

from .base_extractor import BaseExtractor, default_camera_params, list_images


def dummy_features(height: int, width: int, step: int = 32, seed: int = 42):
    """(keypoints f32 [N,2] (x,y), descriptors u8 [N,128]) of one image."""
    ys = np.arange(step // 2, height, step, dtype=np.float32)
    xs = np.arange(step // 2, width, step, dtype=np.float32)
    gy, gx = np.meshgrid(ys, xs, indexing="ij")
    kpts = np.stack([gx.ravel(), gy.ravel()], axis=1).astype(np.float32)
    cache = {}
    desc = np.empty((len(kpts), 128), np.uint8)
    for i, (x, y) in enumerate(kpts):
        key = (int(x / step), int(y / step))
        if key not in cache:
            cache[key] = np.random.RandomState(seed + key[0] * 1000 + key[1]).randint(0, 256, size=128, dtype=np.uint8)
        desc[i] = cache[key]
    return kpts, desc


class DummyExtractor(BaseExtractor):
    def __init__(self, step: int = 32, seed: int = 42):
        self.step = step
        self.seed = seed

    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:
        import cv2
        from ..database import ColmapDatabase, make_camera
        image_dir = Path(image_dir)
        files = list_images(image_dir)
        if not files:  # dummy_extractor.py:45-55
            print(f"No images found in {image_dir}, generating 10 dummy images...")
            image_dir.mkdir(parents=True, exist_ok=True)
            for i in range(10):
                path = image_dir / f"dummy_{i:03d}.png"
                cv2.imwrite(str(path), np.random.randint(0, 256, (480, 640, 3), dtype=np.uint8))
                files.append(path)
        db = ColmapDatabase(str(db_path))
        first = cv2.imread(str(files[0]))
        if first is None:
            return
        height, width = first.shape[:2]
        if camera_params is None:
            camera_params = default_camera_params(camera_model, width, height)
        camera_id = db.db.write_camera(make_camera(camera_model, width, height, camera_params))
        for f in files:
            img = cv2.imread(str(f))
            if img is None:
                continue
            image_id = db.add_image(f.name, camera_id=camera_id)
            kpts, desc = dummy_features(img.shape[0], img.shape[1], self.step, self.seed)
            db.add_keypoints(image_id, kpts)
            db.add_descriptors(image_id, desc)
        db.commit()
        db.close()
