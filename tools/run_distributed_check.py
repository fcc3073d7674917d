"""torchrun entry: runs the whole hot path on WORLD_SIZE GPUs (vit_colmap_b200.distributed) and, on
rank 0, compares the database with the single-process ViTExtractor.extract + match_exhaustive run.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/run_distributed_check.py"""
import os
import sqlite3
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    import cv2
    from vit_colmap_b200 import distributed as vd
    from vit_colmap_b200.features.vit_extractor import ViTExtractor
    from vit_colmap_b200.matching import match_exhaustive
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tmp = Path(os.environ.get("VC_CHECK_DIR", tempfile.gettempdir())) / "vc_dist_check"
    img_dir = tmp / "images"
    n_images = 9
    if rank == 0:
        img_dir.mkdir(parents=True, exist_ok=True)
        rng = np.random.default_rng(0)
        base = cv2.GaussianBlur(rng.integers(0, 256, (700, 700, 3), dtype=np.uint8), (0, 0), 3)
        for i in range(n_images):
            cv2.imwrite(str(img_dir / f"im_{i:02d}.png"), base[14 * i:14 * i + 518, 14 * (i % 3):14 * (i % 3) + 518])  # patch-aligned shifts
        (img_dir / "im_04_broken.png").write_bytes(b"not an image")  # skipped by both runs (vit_extractor.py:733-737)
        for p in (tmp / "dist.db", tmp / "single.db"):
            if p.exists():
                p.unlink()
    if world > 1:
        dist.barrier()
    ex = ViTExtractor(model_name="dinov2_vits14", num_keypoints=512, descriptor_dim=128, device=f"cuda:{local}",
                      synthetic_seed=0, batch_size=4)
    from vit_colmap_b200.matching import FeatureMatchingOptions
    opts = FeatureMatchingOptions(max_ratio=1.05, max_distance=3.2, min_num_matches=1)  # mutual-NN only: clipped u8 ViT descriptors sit at acos(0.5) even from themselves
    res = vd.extract_and_match(img_dir, tmp / "dist.db", ex, block_size=4, matching_options=opts)
    if rank == 0:
        ex2 = ViTExtractor(model_name="dinov2_vits14", num_keypoints=512, descriptor_dim=128, device=f"cuda:{local}",
                           synthetic_seed=0, batch_size=4)
        # torch.svd signs are not reproducible run to run (in the reference too): share the fitted PCA
        ex2.descriptor_projection = ex.descriptor_projection
        ex2.extract(img_dir, tmp / "single.db", "SIMPLE_PINHOLE")
        from vit_colmap_b200.pairing import ExhaustivePairGenerator  # same block size as the sharded run
        class _P:  # noqa: E701
            block_size = 4
        ref = match_exhaustive(str(tmp / "single.db"), matching_options=opts, pairing_options=_P())
        a, b = sqlite3.connect(str(tmp / "dist.db")), sqlite3.connect(str(tmp / "single.db"))
        for table in ("images", "keypoints", "descriptors", "matches"):
            ra = a.execute(f"SELECT * FROM {table} ORDER BY 1").fetchall()
            rb = b.execute(f"SELECT * FROM {table} ORDER BY 1").fetchall()
            assert ra == rb, f"table {table} differs between the {world}-rank and the single-process run"
        n_match = sum(len(m) for m in res.values())
        assert len(res) == n_images * (n_images - 1) // 2 and n_match > 0, (len(res), n_match)
        print(f"distributed check ok: world={world}, {len(res)} pairs, {n_match} matches, databases identical")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
