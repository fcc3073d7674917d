"""Hybrid extractor: OpenCV keypoint detector + ViT descriptors on the sm_100a kernels.

Same constructor, ``_run_inference`` and ``extract`` contracts as the reference class
(vit_colmap/features/hybrid_extractor.py:22-84, 326-420; the reference module also calls its class
``ViTExtractor`` - aliased at the bottom).  Detection stays on the host (cv2, hybrid_extractor.py:
131-180); the dense features come from vc_vit_forward and the descriptors from vc_sample_pack on its
general bilinear path (float feature-map coordinates, border clamp) with the RootSIFT normalisation
(L1 -> sqrt -> L2, :286-288) and the truncating uint8 conversion (:292) fused into the same pass.
"""
from __future__ import annotations

from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np

from ._pipelined import decode_ahead
from .base_extractor import default_camera_params, list_images
from .vit_extractor import PATCH, ViTExtractor as _GridViTExtractor

DETECTORS = ("sift", "fast", "gftt", "orb")


class HybridViTExtractor(_GridViTExtractor):
    def __init__(self, weights_path: Optional[str] = None, model_name: str = "dinov2_vitb14",
                 num_keypoints: int = 2048, descriptor_dim: int = 256, device: Optional[str] = None,
                 detector_type: str = "sift", *, state_dict=None, synthetic_seed: Optional[int] = None,
                 batch_size: int = 16):
        if detector_type not in DETECTORS:  # hybrid_extractor.py:127-128
            raise ValueError(f"Unknown detector type: {detector_type}")
        self.detector_type = detector_type
        print(f"Keypoint detector: {detector_type.upper()}")
        super().__init__(weights_path, model_name, num_keypoints, descriptor_dim, device, "harris",
                         state_dict=state_dict, synthetic_seed=synthetic_seed, batch_size=batch_size)
        self.detector = self._create_detector()
        print("✓ Hybrid model loaded successfully")

    # ------------------------------------------------------------------ host detector
    def _create_detector(self):
        """hybrid_extractor.py:107-128."""
        import cv2
        if self.detector_type == "sift":
            return cv2.SIFT_create(nfeatures=self.num_keypoints)
        if self.detector_type == "fast":
            return cv2.FastFeatureDetector_create(threshold=10, nonmaxSuppression=True)
        if self.detector_type == "gftt":
            return cv2.goodFeaturesToTrack
        return cv2.ORB_create(nfeatures=self.num_keypoints)

    def _thread_detector(self):
        """cv2 Feature2D objects are not documented as re-entrant; extract() detects on decode worker
        threads (decode_ahead), so every thread gets its own detector (same parameters, same results)."""
        import threading
        if getattr(self, "_tls", None) is None:
            self._tls = threading.local()
        if threading.current_thread() is threading.main_thread():
            return self.detector
        if getattr(self._tls, "detector", None) is None:
            self._tls.detector = self._create_detector()
        return self._tls.detector

    def _detect_keypoints(self, image_bgr: np.ndarray) -> np.ndarray:
        """hybrid_extractor.py:131-180: [n, 2] float32 (x, y) pixels of the ORIGINAL image."""
        import cv2
        gray = cv2.cvtColor(image_bgr, cv2.COLOR_BGR2GRAY)
        detector = self._thread_detector()
        if self.detector_type == "gftt":
            corners = detector(gray, maxCorners=self.num_keypoints, qualityLevel=0.01, minDistance=7,
                                    blockSize=7)
            if corners is None:
                return np.zeros((0, 2), np.float32)
            return corners.reshape(-1, 2).astype(np.float32)
        kpts = detector.detect(gray, None)
        keypoints = np.array([[kp.pt[0], kp.pt[1]] for kp in kpts], dtype=np.float32).reshape(-1, 2)
        if self.detector_type == "fast" and len(keypoints) > self.num_keypoints:
            responses = np.array([kp.response for kp in kpts])
            keypoints = keypoints[np.argsort(responses)[-self.num_keypoints:]]
        return keypoints

    # ------------------------------------------------------------------ device descriptors
    @staticmethod
    def grid_coords(keypoints_xy: np.ndarray, orig_wh, feat_wh, grid_hw) -> np.ndarray:
        """hybrid_extractor.py:247-255, same f32 array arithmetic; [n, 2] (y, x) feature-map units."""
        H, W = grid_hw
        kx = keypoints_xy[:, 0] * (feat_wh[0] / orig_wh[0]) * (W / feat_wh[0])
        ky = keypoints_xy[:, 1] * (feat_wh[1] / orig_wh[1]) * (H / feat_wh[1])
        return np.stack([ky, kx], axis=1).astype(np.float32)

    def _fit_projection_float(self, tokens, coords_yx, count: int, grid_hw):
        """First-call PCA / random projection of hybrid_extractor.py:296-324 on the raw bilinear
        rows of the first image (one-off; applied by vc_sample_pack afterwards)."""
        import torch
        import torch.nn.functional as F
        H, W = grid_hw
        C = tokens.shape[-1]
        fm = tokens[0].float().reshape(H, W, C).permute(2, 0, 1)[None]
        c = coords_yx[0, :count]
        gx = 2.0 * c[:, 1] / (W - 1) - 1.0
        gy = 2.0 * c[:, 0] / (H - 1) - 1.0
        grid = torch.stack([gx, gy], dim=1)[None, :, None]
        desc = F.grid_sample(fm, grid, mode="bilinear", padding_mode="border", align_corners=True)
        desc = desc[0, :, :, 0].t().cpu()
        if desc.shape[0] > self.descriptor_dim:
            centered = desc - desc.mean(dim=0, keepdim=True)
            _, S, V = torch.svd(centered)
            proj = V[:, : self.descriptor_dim].contiguous()
            print(f"Initialized PCA projection: {C} -> {self.descriptor_dim}")
            print(f"  Variance explained: {(S[:self.descriptor_dim].sum() / S.sum()).item():.2%}")
        else:
            proj = torch.randn(C, self.descriptor_dim, dtype=torch.float32) / np.sqrt(C)
            print(f"Initialized random projection: {C} -> {self.descriptor_dim}")
        self.descriptor_projection = proj.to(self.device)

    def _device_stage(self, batch, orig_wh, resized_wh, extra=None):
        """extra: per-image detector keypoints in ORIGINAL pixel coordinates."""
        import torch
        gh, gw = batch.shape[1] // PATCH, batch.shape[2] // PATCH
        tokens = self.model.forward_patch_tokens(batch, bgr=True)
        B, C = tokens.shape[0], tokens.shape[-1]
        n = np.array([len(k) for k in extra], np.int32)
        cap = max(int(n.max()), 1)
        host = np.zeros((B, cap, 2), np.float32)
        kp_host = np.zeros((B, cap, 2), np.float32)
        for i, k in enumerate(extra):
            if len(k):
                k = np.asarray(k, np.float32).reshape(-1, 2)
                kp_host[i, : len(k)] = k
                host[i, : len(k)] = self.grid_coords(k, orig_wh, resized_wh, (gh, gw))
        coords = torch.from_numpy(host).to(self.device)
        counts = torch.from_numpy(n).to(self.device)
        if C > self.descriptor_dim and self.descriptor_projection is None and n[0] > 0:
            self._fit_projection_float(tokens, coords, int(n[0]), (gh, gw))
        proj = self.descriptor_projection if C > self.descriptor_dim else None
        if C > self.descriptor_dim and proj is None:
            raise RuntimeError("no keypoints on the first image: the descriptor projection cannot be fitted")
        _, dbf, du8, _ = self.post.sample_pack(tokens, (gh, gw), coords, counts, orig_wh, resized_wh, projection=proj,
                                               rootsift=True)
        # keypoints stay the detector's own (hybrid_extractor.py:326-357)
        return torch.from_numpy(kp_host), du8, counts, (counts, dbf)

    def _extract_batch(self, images_bgr: List[np.ndarray], orig_wh, resized_wh, keypoints=None):
        return self._run_batch(images_bgr, orig_wh, resized_wh, keypoints)

    def _run_inference(self, image_bgr: np.ndarray):
        """hybrid_extractor.py:326-357."""
        keypoints = self._detect_keypoints(image_bgr)
        if len(keypoints) == 0:
            print("Warning: No keypoints detected")
            return keypoints, np.zeros((0, self.descriptor_dim), dtype=np.uint8)
        img, orig_wh, resized_wh = self._resize(image_bgr)
        out, _ = self._extract_batch([img], orig_wh, resized_wh, [keypoints])
        return out[0]

    # ------------------------------------------------------------------ plugin entry
    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:
        """hybrid_extractor.py:359-450: sorted files, one shared camera, per-image failures skipped."""
        import cv2
        from ..database import ColmapDatabase, make_camera
        image_dir = Path(image_dir)
        print(f"\n{'=' * 60}\nHybrid Feature Extraction\n{'=' * 60}")
        print(f"Keypoint detector: {self.detector_type.upper()}\nDescriptor: ViT ({self.model_name})\n{'=' * 60}\n")
        files = list_images(image_dir)
        if not files:
            raise ValueError(f"No images found in {image_dir}")
        print(f"Found {len(files)} images")
        db = ColmapDatabase(str(db_path))
        first = cv2.imread(str(files[0]))
        if first is None:
            raise ValueError(f"Failed to read first image: {files[0]}")
        height, width = first.shape[:2]
        if camera_params is None:
            camera_params = default_camera_params(camera_model, width, height)
        camera_id = db.db.write_camera(make_camera(camera_model, width, height, camera_params))
        print(f"Camera ID: {camera_id}\n")

        def prepare(path):  # worker threads: decode, host-side detector, resize (None: unreadable file)
            img = cv2.imread(str(path))
            if img is None:
                return None
            kps = self._detect_keypoints(img)
            return kps, self._resize(img)

        def items():
            for k, f, got in decode_ahead(files, prepare):
                if got is None:
                    print(f"[{k + 1}/{len(files)}] {f.name}: ⚠ Warning: Failed to read image, skipping")
                    continue
                image_id = db.add_image(f.name, camera_id=camera_id)
                if isinstance(got, Exception):
                    print(f"  ✗ Error: {got}")
                    continue
                kps, (img, orig_wh, resized_wh) = got
                yield image_id, img, orig_wh, resized_wh, kps

        def write(image_id, kps, desc):
            if len(kps) == 0:
                print("  ⚠ Warning: No keypoints extracted")
                return
            db.add_keypoints(image_id, kps)
            db.add_descriptors(image_id, desc)

        def on_error(k, e):
            print(f"  ✗ Error: {e}")

        self.extract_stream(items(), self.batch_size, write, on_error,
                            first_alone=lambda: self.descriptor_projection is None
                            and self.model.embed_dim > self.descriptor_dim)
        db.commit()
        db.close()
        print(f"\n{'=' * 60}\n✓ Hybrid extraction complete!\n{'=' * 60}\n")


ViTExtractor = HybridViTExtractor  # the reference module's class name
