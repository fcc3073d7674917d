"""``vit`` extractor on the sm_100a kernels.

Same constructor, ``extract`` and ``_run_inference`` contracts as the reference class
(vit_colmap/features/vit_extractor.py:17-116, 655-768); the per-image serial eager loop is replaced
by batches: decode on the host, one upload per batch, ViT forward (vc_vit_forward), score map,
selection, sampling / projection / packing (vc_score_map, vc_select_keypoints, vc_sample_pack), one
download per batch, rows written by a single database writer in sorted file order.
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np

from ._pipelined import BatchPipeline, decode_ahead
from .base_extractor import BaseExtractor, default_camera_params, list_images

PATCH = 14


class ViTExtractor(BatchPipeline, BaseExtractor):
    def __init__(self, weights_path: Optional[str] = None, model_name: str = "dinov2_vitb14",
                 num_keypoints: int = 2048, descriptor_dim: int = 128, device: Optional[str] = None,
                 detection_method: str = "harris", *, state_dict=None, synthetic_seed: Optional[int] = None,
                 batch_size: int = 32):
        import torch
        from .. import _lib
        from ..extraction import PostProcessor
        from ..vit import CONFIGS, VitBackbone
        self.weights_path = weights_path
        self.model_name = model_name
        self.num_keypoints = num_keypoints
        self.descriptor_dim = descriptor_dim
        self.detection_method = detection_method
        self.batch_size = int(batch_size)
        self.patch_size = PATCH
        if not torch.cuda.is_available():
            raise _lib.VcError("ViTExtractor needs an sm_100 GPU: vit_colmap_b200 has no CPU fallback")
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if dev.type != "cuda":
            raise _lib.VcError(f"device {dev} requested; vit_colmap_b200 runs on CUDA only")
        self.device = dev
        self.device_index = dev.index if dev.index is not None else torch.cuda.current_device()
        print(f"Initializing ViT extractor: {model_name} on {self.device}")
        if model_name not in CONFIGS:
            raise ValueError(f"Unsupported model: {model_name}. Currently only DINOv2 models are supported.")
        sd = self._load_state_dict(state_dict, synthetic_seed)
        self.model = VitBackbone(sd, model_name, self.device_index)
        self.post = PostProcessor(self.device_index, handle=self.model.handle)
        self.descriptor_projection = None  # [C, d] f32 on the device, fitted on the first image
        print("✓ ViT model loaded successfully")

    # ------------------------------------------------------------------ weights
    def _load_state_dict(self, state_dict, synthetic_seed):
        import torch
        if self.weights_path is not None:  # vit_extractor.py:88-92
            print(f"Loading custom weights from: {self.weights_path}")
            raise NotImplementedError("Custom weight loading not yet implemented")
        if state_dict is not None:
            return state_dict
        if synthetic_seed is not None:
            from ..vit import synthetic_state_dict
            return synthetic_state_dict(self.model_name, synthetic_seed)
        # the files torch.hub would have cached for facebookresearch/dinov2
        names = [f"{self.model_name}_pretrain.pth", f"{self.model_name}.pth"]
        dirs = [os.environ.get("VIT_COLMAP_B200_WEIGHTS_DIR"), os.path.join(torch.hub.get_dir(), "checkpoints")]
        for d in dirs:
            for n in names:
                if d and os.path.exists(os.path.join(d, n)):
                    return torch.load(os.path.join(d, n), map_location="cpu")
        raise FileNotFoundError(
            f"no DINOv2 checkpoint for {self.model_name}: put {names[0]} (the facebookresearch/dinov2 hub state "
            f"dict) under $VIT_COLMAP_B200_WEIGHTS_DIR or {dirs[1]}, or pass state_dict= / synthetic_seed=")

    # ------------------------------------------------------------------ host preprocessing
    def _resize(self, image_bgr: np.ndarray) -> Tuple[np.ndarray, Tuple[int, int], Tuple[int, int]]:
        """vit_extractor.py:118-129: only the resize stays on the host (cv2.INTER_LINEAR semantics);
        BGR->RGB, ToTensor and Normalize are fused into the patchify kernel."""
        import cv2
        h, w = image_bgr.shape[:2]
        h_new, w_new = (h // PATCH) * PATCH, (w // PATCH) * PATCH
        if h_new == 0 or w_new == 0:
            raise ValueError(f"image {w}x{h} is smaller than one 14x14 patch")
        if (h_new, w_new) != (h, w):
            image_bgr = cv2.resize(image_bgr, (w_new, h_new), interpolation=cv2.INTER_LINEAR)
        return np.ascontiguousarray(image_bgr), (w, h), (w_new, h_new)

    def _fit_projection(self, tokens, coords, count: int):
        """First-call PCA of vit_extractor.py:601-639 on the raw sampled rows of the first image
        (one-off host step; the projection itself is applied by vc_sample_pack)."""
        import torch
        C = tokens.shape[-1]
        W = self._grid[1]
        idx = (coords[0, :count, 0].long() * W + coords[0, :count, 1].long())
        desc = tokens[0].index_select(0, idx).float().cpu()
        if desc.shape[0] > self.descriptor_dim:
            centered = desc - desc.mean(dim=0, keepdim=True)
            _, S, V = torch.svd(centered)
            proj = V[:, : self.descriptor_dim].contiguous()
            print(f"Initialized PCA projection: {C} -> {self.descriptor_dim}")
            print(f"  Variance explained: {(S[:self.descriptor_dim].sum() / S.sum()).item():.2%}")
        else:
            proj = torch.randn(C, self.descriptor_dim, dtype=torch.float32) / np.sqrt(C)
            print(f"Initialized random projection (insufficient samples): {C} -> {self.descriptor_dim}")
        self.descriptor_projection = proj.to(self.device)

    # ------------------------------------------------------------------ batched core
    def _device_stage(self, batch, orig_wh, resized_wh, extra=None):
        """Device half of one batch: ViT forward, score map, selection, sampling / PCA / packing."""
        gh, gw = batch.shape[1] // PATCH, batch.shape[2] // PATCH
        self._grid = (gh, gw)
        tokens = self.model.forward_patch_tokens(batch, bgr=True)
        C = tokens.shape[-1]
        score = self.post.score_map(tokens, (gh, gw), self.detection_method)
        coords, scores, counts = self.post.select_keypoints(score, self.num_keypoints)
        if C > self.descriptor_dim and self.descriptor_projection is None:
            self._fit_projection(tokens, coords, int(counts[0]))
        proj = self.descriptor_projection if C > self.descriptor_dim else None
        kps, dbf, du8, _ = self.post.sample_pack(tokens, (gh, gw), coords, counts, orig_wh, resized_wh, projection=proj)
        return kps, du8, counts, (counts, dbf)

    def _extract_batch(self, images_bgr: List[np.ndarray], orig_wh, resized_wh):
        """images: equally sized BGR uint8 arrays (already multiples of 14).  Returns per-image
        (keypoints f32 [n,2], descriptors u8 [n,d]) plus the device-resident (counts, bf16 rows)."""
        return self._run_batch(images_bgr, orig_wh, resized_wh)

    def _run_inference(self, image_bgr: np.ndarray):
        """Single-image convenience over the batched path (vit_extractor.py:106-166)."""
        img, orig_wh, resized_wh = self._resize(image_bgr)
        out, _ = self._extract_batch([img], orig_wh, resized_wh)
        return out[0]

    # ------------------------------------------------------------------ plugin entry
    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:
        import cv2
        from ..database import ColmapDatabase, make_camera
        image_dir = Path(image_dir)
        print(f"\n{'=' * 60}\nViT Feature Extraction\n{'=' * 60}")
        print(f"Image directory: {image_dir}\nDatabase: {db_path}\nModel: {self.model_name}")
        print(f"Target keypoints per image: {self.num_keypoints}\n{'=' * 60}\n")
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
        print(f"Image dimensions: {width}x{height}\nCamera model: {camera_model}\nCamera ID: {camera_id}\n")

        # image ids follow the sorted file order (single writer); features are computed in batches
        # through BatchPipeline.extract_stream: decoding / staging of the next batch and the database
        # writes of the previous one run under the current batch's kernels
        def prepare(path):  # worker threads: decode + resize (None: unreadable file)
            img = cv2.imread(str(path))
            return None if img is None else (img, self._resize(img))

        def items():
            for k, f, got in decode_ahead(files, prepare):
                if got is None:
                    print(f"[{k + 1}/{len(files)}] {f.name}: ⚠ Warning: Failed to read image, skipping")
                    continue
                if isinstance(got, Exception) and not isinstance(got, ValueError):
                    raise got
                image_id = db.add_image(f.name, camera_id=camera_id)   # the image row exists even if extraction fails
                if isinstance(got, Exception):                          # _resize refused the image (vit_extractor.py:757-762)
                    print(f"  ✗ Error during feature extraction: {got}")
                    continue
                _, (img, orig_wh, resized_wh) = got
                yield image_id, img, orig_wh, resized_wh, None

        def write(image_id, kps, desc):
            if len(kps) == 0:
                print("  ⚠ Warning: No keypoints extracted")
                return
            db.add_keypoints(image_id, kps)
            db.add_descriptors(image_id, desc)

        def on_error(k, e):  # per-image skip semantics, vit_extractor.py:757-762
            print(f"  ✗ Error during feature extraction: {e}")
            import traceback
            traceback.print_exception(type(e), e, e.__traceback__)

        # the PCA projection is fitted on the first image alone, as in the reference
        self.extract_stream(items(), self.batch_size, write, on_error,
                            first_alone=lambda: self.descriptor_projection is None
                            and self.model.embed_dim > self.descriptor_dim)
        db.commit()
        db.close()
        print(f"\n{'=' * 60}\n✓ Feature extraction complete!\n{'=' * 60}\n")
