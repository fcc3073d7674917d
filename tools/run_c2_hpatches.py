"""BASELINE config C2 end to end through the plugin interface: HPatches-shaped data, 116 synthetic sequences of
6 images 800 x 600 (resized to 798 x 588 on the host, grid 42 x 57, 2395 tokens): a smoothed-noise base image
and five views of it through random homographies (corner jitter <= 15 %).  Per sequence, as the
reference's pipeline would run it: `ViTExtractor.extract(sequence_dir, db)` (ViT-B/14, seeded random-init
weights, num_keypoints 2048, 128-d PCA descriptors) then `match_exhaustive(db)` over its 15 pairs
(cross-checked nearest neighbours).  Wall-clock per stage, files on local disk.

    python tools/run_c2_hpatches.py [n_sequences=116] [model=dinov2_vitb14]"""
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


def make_sequence(rng, d: Path):
    base = cv2.GaussianBlur(rng.integers(0, 256, (600, 800, 3), dtype=np.uint8), (0, 0), 3.0)
    base = cv2.normalize(base, None, 0, 255, cv2.NORM_MINMAX)
    cv2.imwrite(str(d / "1.png"), base)
    src = np.float32([[0, 0], [800, 0], [800, 600], [0, 600]])
    for i in range(2, 7):
        dst = src + rng.uniform(-0.15, 0.15, (4, 2)).astype(np.float32) * np.float32([800, 600])
        cv2.imwrite(str(d / f"{i}.png"), cv2.warpPerspective(base, cv2.getPerspectiveTransform(src, dst), (800, 600)))


def main():
    n_seq = int(sys.argv[1]) if len(sys.argv) > 1 else 116
    model = sys.argv[2] if len(sys.argv) > 2 else "dinov2_vitb14"
    rng = np.random.default_rng(116)
    with tempfile.TemporaryDirectory() as tmp:
        root = Path(tmp)
        for s in range(n_seq + 1):
            (root / f"seq{s:03d}").mkdir()
            make_sequence(rng, root / f"seq{s:03d}")
        ex = ViTExtractor(model_name=model, num_keypoints=2048, descriptor_dim=128, synthetic_seed=0, batch_size=6)
        # mutual nearest neighbours (as tools/run_c3_pipeline.py): the clipped uint8 rows of a random-init backbone
        # pass COLMAP's default thresholds almost nowhere; the matcher's work per pair does not depend on them
        opts = FeatureMatchingOptions(max_ratio=1.05, max_distance=3.2)
        ex.extract(root / f"seq{n_seq:03d}", root / "warm.db", "SIMPLE_PINHOLE")   # allocator pools, PCA fit, opt-ins
        match_exhaustive(str(root / "warm.db"), opts)
        torch.cuda.synchronize()
        t_ex = t_m = 0.0
        pairs = rows = kps = 0
        stages = {}
        import sqlite3
        for s in range(n_seq):
            db = root / f"seq{s:03d}.db"
            t0 = time.perf_counter()
            ex.extract(root / f"seq{s:03d}", db, "SIMPLE_PINHOLE")
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            st = {}
            res = match_exhaustive(str(db), opts, stats=st)
            t2 = time.perf_counter()
            t_ex += t1 - t0
            t_m += t2 - t1
            pairs += len(res)
            rows += int(sum(len(m) for m in res.values()))
            for k, v in st.items():
                stages[k] = stages.get(k, 0.0) + v
            con = sqlite3.connect(str(db))
            kps += sum(r for (r,) in con.execute("SELECT rows FROM keypoints"))
            con.close()
        print(json.dumps({"config": f"C2: {n_seq} sequences x 6 images 800x600 -> 798x588, {model}, num_keypoints 2048, "
                                    "15 pairs per sequence, one extract() + match_exhaustive() call per sequence",
                          "images": 6 * n_seq, "extract_seconds": round(t_ex, 3), "images_per_s": round(6 * n_seq / t_ex, 1),
                          "mean_keypoints": round(kps / (6 * n_seq), 1), "pairs": pairs, "match_seconds": round(t_m, 3),
                          "pairs_per_s": round(pairs / t_m, 1), "match_rows": rows,
                          "match_stage_seconds": {k: round(v, 4) for k, v in stages.items()},
                          "includes": "png decode + cv2 resize on the host, H2D, forward, post-processing, D2H, sqlite "
                                      "writes / reads, one database per sequence"}))


if __name__ == "__main__":
    main()
