"""Host-side mirror of the reference's plugin / Pipeline / Config interface (no GPU needed)."""
import logging
import sqlite3
from pathlib import Path

import numpy as np
import pytest


def _checkerboard(h=480, w=640, tile=40):
    """tests/test_smoke_e2e.py:10-17 pattern."""
    yy, xx = np.mgrid[0:h, 0:w]
    img = (((yy // tile) + (xx // tile)) % 2 * 255).astype(np.uint8)
    return np.stack([img] * 3, axis=-1)


def _write_shifted(tmp_path, n=3, h=480, w=640, tile=40):
    import cv2
    d = tmp_path / "images"
    d.mkdir()
    base = _checkerboard(h, w, tile)
    for i in range(n):
        m = np.float32([[1, 0, 50 * i], [0, 1, 30 * i]])
        cv2.imwrite(str(d / f"img_{i:03d}.png"), cv2.warpAffine(base, m, (w, h)))
    return d


def test_config_defaults_match_reference():
    from vit_colmap_b200.utils.config import Config
    c = Config()
    assert (c.camera.model, c.extractor.extractor_type, c.extractor.vit_weights_path) == ("SIMPLE_PINHOLE", "vit", None)
    assert (c.matching.use_gpu, c.matching.max_ratio, c.matching.max_distance, c.matching.cross_check,
            c.matching.num_threads) == (True, 0.8, 0.7, True, -1)
    assert (c.reconstruction.min_num_matches, c.reconstruction.multiple_models) == (15, True)
    assert c.do_matching and c.do_reconstruction and c.log.level == logging.INFO
    assert c.camera.get_default_params(640, 480) == [640, 320.0, 240.0]
    c.camera.model = "PINHOLE"
    assert c.camera.get_default_params(640, 480) == [640, 640, 320.0, 240.0]
    c.camera.model = "OPENCV"
    with pytest.raises(ValueError):
        c.camera.get_default_params(1, 1)
    o = c.matching.to_matching_options()
    assert (o.max_ratio, o.max_distance, o.cross_check, o.max_num_matches, o.guided_matching) == (0.8, 0.7, True, 32768, False)
    assert "Extractor: vit" in c.summary()


def test_config_from_args():
    from types import SimpleNamespace
    from vit_colmap_b200.utils.config import Config
    a = SimpleNamespace(camera_model="PINHOLE", extractor=None, use_colmap_sift=True, vit_weights=None, model=None,
                        use_gpu=False, skip_matching=True, skip_reconstruction=True, verbose=True)
    c = Config.from_args(a)
    assert c.extractor.extractor_type == "colmap_sift" and c.camera.model == "PINHOLE"
    assert not c.do_matching and not c.do_reconstruction and not c.matching.use_gpu and c.log.level == logging.DEBUG
    a.extractor, a.vit_weights = "dummy", Path("w.pth")
    c = Config.from_args(a)
    assert c.extractor.extractor_type == "dummy" and c.extractor.vit_weights_path == "w.pth"


def test_dummy_extractor_writes_reference_layout(tmp_path):
    """KA8: 640x480 / step 32 -> 20x15 = 300 keypoints per image, identical descriptors for identical
    grid cells in every image (dummy_extractor.py:94-109)."""
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.features.dummy_extractor import DummyExtractor, dummy_features
    d = _write_shifted(tmp_path)
    db_path = tmp_path / "db.db"
    DummyExtractor(step=32).extract(d, db_path, "PINHOLE", None)
    with ColmapDatabase.open_database(str(db_path)) as db:
        assert ColmapDatabase.get_db_count(db, "num_images") == 3 and ColmapDatabase.get_db_count(db, "num_cameras") == 1
        k1, k2 = db.read_keypoints(1), db.read_keypoints(2)
        d1, d2 = db.read_descriptors(1), db.read_descriptors(3)
        assert all(db.exists_keypoints(i) and db.exists_descriptors(i) for i in (1, 2, 3))
    assert k1.shape == (300, 2) and k1.dtype == np.float32 and np.array_equal(k1, k2)
    assert d1.shape == (300, 128) and d1.dtype == np.uint8 and np.array_equal(d1, d2)
    assert k1[0].tolist() == [16.0, 16.0] and k1[1].tolist() == [48.0, 16.0]
    # the descriptor of cell (gx, gy) is NumPy's legacy stream seeded with 42 + 1000*gx + gy
    assert np.array_equal(d1[21], np.random.RandomState(42 + 1 * 1000 + 1).randint(0, 256, size=128, dtype=np.uint8))
    kp, ds = dummy_features(480, 640)
    assert np.array_equal(kp, k1) and np.array_equal(ds, d1)
    con = sqlite3.connect(str(db_path))
    assert con.execute("SELECT rows, cols FROM keypoints WHERE image_id=1").fetchone() == (300, 2)
    assert con.execute("SELECT model, width, height FROM cameras").fetchone() == (1, 640, 480)


def test_dummy_extractor_generates_images_when_dir_is_empty(tmp_path):
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.features.dummy_extractor import DummyExtractor
    d = tmp_path / "empty"
    d.mkdir()
    DummyExtractor().extract(d, tmp_path / "db.db", "SIMPLE_PINHOLE")
    assert len(list(d.glob("dummy_*.png"))) == 10
    with ColmapDatabase.open_database(str(tmp_path / "db.db")) as db:
        assert db.num_images() == 10


def test_extractor_switch_and_error_contracts(tmp_path):
    import torch
    from vit_colmap_b200.features import BaseExtractor, ColmapSiftExtractor, DummyExtractor
    from vit_colmap_b200.pipeline.run_pipeline import Pipeline
    from vit_colmap_b200.utils.config import Config
    c = Config(do_reconstruction=False)
    c.extractor.extractor_type = "dummy"
    assert isinstance(Pipeline(c)._make_extractor(), DummyExtractor)
    c.extractor.extractor_type = "colmap_sift"
    ex = Pipeline(c)._make_extractor()
    assert isinstance(ex, ColmapSiftExtractor) and isinstance(ex, BaseExtractor)
    with pytest.raises(ImportError):
        ex.extract(tmp_path, tmp_path / "db.db", "PINHOLE")
    if not torch.cuda.is_available():  # the GPU extractors refuse to run without a device
        from vit_colmap_b200 import _lib
        for t in ("trainable_vit", "anything-else-falls-through-to-vit"):
            c.extractor.extractor_type = t
            with pytest.raises(_lib.VcError):
                Pipeline(c)._make_extractor()
    with pytest.raises(TypeError):
        BaseExtractor()
