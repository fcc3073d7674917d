"""Pipeline API (reference: vit_colmap/pipeline/run_pipeline.py Pipeline :20-418, main :420-510).

Same constructor, ``run`` signature, extractor switch and return contract.  Feature extraction and
exhaustive matching run on the sm_100a kernels; geometric verification and incremental mapping stay
with pycolmap (untouched by this project) and are invoked only when pycolmap is importable."""
from __future__ import annotations

import argparse
import logging
from pathlib import Path
from typing import Optional

from ..database import ColmapDatabase
from ..features.base_extractor import BaseExtractor
from ..utils.config import Config

try:  # pragma: no cover - absent in the build image
    import pycolmap
except Exception:  # noqa: BLE001
    pycolmap = None

logger = logging.getLogger(__name__)


class Pipeline:
    def __init__(self, config: Optional[Config] = None):
        self.config = config if config is not None else Config()

    # run_pipeline.py:319-339
    def _make_extractor(self) -> BaseExtractor:
        ex = self.config.extractor
        if ex.extractor_type == "dummy":
            logger.info("Using Dummy extractor")
            from ..features.dummy_extractor import DummyExtractor
            return DummyExtractor(step=32)
        if ex.extractor_type == "colmap_sift":
            logger.info("Using COLMAP SIFT extractor")
            from ..features.colmap_sift_extractor import ColmapSiftExtractor
            return ColmapSiftExtractor()
        if ex.extractor_type == "trainable_vit":
            logger.info("Using Trainable ViT extractor")
            from ..features.trainable_vit_extractor import TrainableViTExtractor
            return TrainableViTExtractor(weights_path=ex.vit_weights_path, num_keypoints=20480, nms_radius=1,
                                         score_threshold=0.4)
        logger.info("Using ViT extractor")
        from ..features.vit_extractor import ViTExtractor
        return ViTExtractor(weights_path=ex.vit_weights_path, model_name=ex.model_name,
                            num_keypoints=ex.num_keypoints, descriptor_dim=ex.descriptor_dim,
                            batch_size=ex.batch_size)

    def _print_summary(self, db_path, output_dir, do_matching, do_reconstruction, reconstructions) -> None:
        logger.info("")
        logger.info("=" * 60)
        logger.info("Pipeline Summary")
        logger.info("=" * 60)
        with ColmapDatabase.open_database(str(db_path)) as db:
            logger.info("Feature Extraction:")
            logger.info(f"  - Images processed: {ColmapDatabase.get_db_count(db, 'num_images')}")
            logger.info(f"  - Cameras: {ColmapDatabase.get_db_count(db, 'num_cameras')}")
            if do_matching:
                logger.info("Feature Matching:")
                logger.info(f"  - Matched image pairs: {ColmapDatabase.get_db_count(db, 'num_matched_image_pairs')}")
        if do_reconstruction and reconstructions:
            for idx, recon in reconstructions.items():
                logger.info(f"  Reconstruction {idx}: {len(recon.images)} images, {len(recon.points3D)} points")
        logger.info(f"Output directory: {output_dir}")

    def run(self, image_dir: Path, output_dir: Path, db_path: Path, dataset: Optional[str] = None,
            scene: Optional[str] = None, results_dir: Optional[Path] = None):
        """Feature extraction, exhaustive matching and (with pycolmap) reconstruction.  Returns the
        dictionary of reconstructions if do_reconstruction else None (run_pipeline.py:274-418)."""
        image_dir, output_dir, db_path = Path(image_dir), Path(output_dir), Path(db_path)
        camera_model = self.config.camera.model
        camera_params = self.config.camera.params
        output_dir.mkdir(parents=True, exist_ok=True)
        db_path.parent.mkdir(parents=True, exist_ok=True)

        extractor = self._make_extractor()
        logger.info("Extracting features...")
        extractor.extract(image_dir, db_path, camera_model, camera_params)
        with ColmapDatabase.open_database(str(db_path)) as db_check:
            logger.info(f"Extracted features for {ColmapDatabase.get_db_count(db_check, 'num_images')} images")

        if self.config.do_matching:
            logger.info("Running feature matching...")
            from ..matching import match_exhaustive
            opts = self.config.matching.to_matching_options()
            if self.config.extractor.extractor_type == "dummy":
                # uniform-random uint8 descriptors have norm >> 512, which saturates COLMAP's angular
                # rule (theta = 0 everywhere, nothing matches); its CPU index rule works in L2 space
                # and matches identical grid cells one-to-one
                opts.rule = "l2"
            match_exhaustive(database_path=str(db_path), matching_options=opts)
            if pycolmap is not None and hasattr(pycolmap, "verify_matches"):  # geometric verification: untouched
                self._verify_with_pycolmap(db_path, output_dir)
            with ColmapDatabase.open_database(str(db_path)) as db_check:
                logger.info(f"Matched {ColmapDatabase.get_db_count(db_check, 'num_matched_image_pairs')} image pairs")

        reconstructions = None
        if self.config.do_reconstruction:
            if pycolmap is None:
                raise ImportError("do_reconstruction=True delegates to pycolmap.incremental_mapping, which is not "
                                  "installed; run with do_reconstruction=False or install pycolmap")
            logger.info("Running 3D reconstruction...")
            sparse_dir = output_dir / "sparse"
            sparse_dir.mkdir(parents=True, exist_ok=True)
            reconstructions = pycolmap.incremental_mapping(
                database_path=str(db_path), image_path=str(image_dir), output_path=str(sparse_dir),
                options=self.config.reconstruction.to_mapper_options())
        self._print_summary(db_path, output_dir, self.config.do_matching, self.config.do_reconstruction,
                            reconstructions)
        return reconstructions

    def _verify_with_pycolmap(self, db_path: Path, output_dir: Path) -> None:  # pragma: no cover
        """pycolmap.verify_matches over the pairs we wrote (PYCOLMAP/pipeline/match_features.cc:53-71;
        pairs file format COLMAP/feature/pairing.cc:51-73)."""
        from ..database import Database
        pairs_path = output_dir / "matched_pairs.txt"
        with Database.open(str(db_path)) as db:
            names = {im.image_id: im.name for im in db.read_all_images()}
            with open(pairs_path, "w") as f:
                for (a, b), m in db.read_all_matches():
                    if len(m):
                        f.write(f"{names[a]} {names[b]}\n")
        pycolmap.verify_matches(str(db_path), str(pairs_path))


def main() -> None:
    ap = argparse.ArgumentParser(description="Run ViT-COLMAP Structure-from-Motion pipeline (B200 hot path)")
    ap.add_argument("--images", required=True, type=Path, help="Directory containing input images")
    ap.add_argument("--output", required=True, type=Path, help="Output directory for reconstruction")
    ap.add_argument("--db", default=Path("data/intermediate/database.db"), type=Path, help="COLMAP database file")
    ap.add_argument("--model", default=None, type=Path, help="Path to ViT model weights (optional)")
    ap.add_argument("--camera-model", default="SIMPLE_PINHOLE", type=str)
    ap.add_argument("--skip-matching", action="store_true")
    ap.add_argument("--skip-reconstruction", action="store_true")
    ap.add_argument("--verbose", "-v", action="store_true")
    ap.add_argument("--use-colmap-sift", action="store_true")
    ap.add_argument("--extractor", type=str, default=None, choices=["vit", "trainable_vit", "colmap_sift", "dummy"])
    ap.add_argument("--vit-weights", type=Path, default=None)
    ap.add_argument("--dataset", type=str, default=None)
    ap.add_argument("--scene", type=str, default=None)
    ap.add_argument("--export-metrics", type=Path, default=None)
    args = ap.parse_args()
    config = Config.from_args(args)
    logger.info("Configuration loaded")
    Pipeline(config=config).run(image_dir=args.images, output_dir=args.output, db_path=args.db,
                                dataset=args.dataset, scene=args.scene, results_dir=args.export_metrics)
    logger.info("Pipeline complete!")


if __name__ == "__main__":
    main()
