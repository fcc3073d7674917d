"""BASELINE config C3 end to end through the plugin interface: 49 synthetic 1600 x 1200 images (DTU scan
shape: resized to 1596 x 1190 on the host, grid 85 x 114, 9691 tokens), ViT-B/14 (seeded random-init
weights: no checkpoint can be fetched here), num_keypoints 8192, `ViTExtractor.extract` into a COLMAP
database, then `match_exhaustive` over its 1176 pairs.  Wall-clock per stage, files on local disk.

    python tools/run_c3_pipeline.py [n_images=49] [model=dinov2_vitb14]"""
import json
import sys
import tempfile
import time
from pathlib import Path

import cv2
import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vit_colmap_b200.features.vit_extractor import ViTExtractor  # noqa: E402
from vit_colmap_b200.matching import FeatureMatchingOptions, match_exhaustive  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 49
    model = sys.argv[2] if len(sys.argv) > 2 else "dinov2_vitb14"
    rng = np.random.default_rng(49)
    with tempfile.TemporaryDirectory() as tmp:
        d = Path(tmp) / "scan"
        d.mkdir()
        big = cv2.GaussianBlur(rng.integers(0, 256, (1300 + 2 * n, 1800 + 4 * n, 3), dtype=np.uint8), (0, 0), 2.5)
        big = cv2.normalize(big, None, 0, 255, cv2.NORM_MINMAX)
        yy, xx = np.mgrid[0:big.shape[0], 0:big.shape[1]]
        big = np.where((((yy // 64) + (xx // 64)) % 2 == 0)[..., None], big, 255 - big).astype(np.uint8)
        for i in range(n):
            cv2.imwrite(str(d / f"rect_{i + 1:03d}_3_r5000.png"), big[2 * i:2 * i + 1200, 4 * i:4 * i + 1600])
        ex = ViTExtractor(model_name=model, num_keypoints=8192, descriptor_dim=128, synthetic_seed=0, batch_size=4)
        warm = Path(tmp) / "warm"
        warm.mkdir()
        for i in range(5):  # allocator pools, PCA fit, kernel opt-ins
            cv2.imwrite(str(warm / f"w{i}.png"), big[:1200, 8 * i:8 * i + 1600])
        ex.extract(warm, Path(tmp) / "warm.db", "SIMPLE_PINHOLE")
        torch.cuda.synchronize()
        db = Path(tmp) / "c3.db"
        t0 = time.perf_counter()
        ex.extract(d, db, "SIMPLE_PINHOLE")
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        stages = {}
        # mutual nearest neighbours: the clipped uint8 ViT rows accept nothing at COLMAP's default thresholds
        res = match_exhaustive(str(db), FeatureMatchingOptions(max_ratio=1.05, max_distance=3.2), stats=stages)
        t2 = time.perf_counter()
        import sqlite3
        con = sqlite3.connect(str(db))
        kp = [r for (r,) in con.execute("SELECT rows FROM keypoints")]
        print(json.dumps({"config": f"C3: {n} x 1600x1200 -> 1596x1190, {model}, num_keypoints 8192, exhaustive matching into the database",
                          "extract_seconds": t1 - t0, "images_per_s": n / (t1 - t0), "mean_keypoints": float(np.mean(kp)),
                          "match_seconds": t2 - t1, "pairs": len(res), "pairs_per_s": len(res) / (t2 - t1),
                          "match_stage_seconds": {k: round(v, 4) for k, v in stages.items()},
                          "match_rows": int(sum(len(m) for m in res.values())),
                          "includes": "png decode + cv2 resize on the host, H2D, forward, post-processing, D2H, sqlite writes"}))


if __name__ == "__main__":
    main()
