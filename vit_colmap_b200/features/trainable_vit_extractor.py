"""``trainable_vit`` extractor on the sm_100a kernels.

Same constructor, ``extract`` and ``_run_inference`` contracts as the reference class
(vit_colmap/features/trainable_vit_extractor.py:18-395).  ViTFeatureModel
(vit_colmap/model/vit_feature_model.py:32-232) runs as vc_vit_forward (DINOv2 backbone) +
vc_conv_head_forward (upsampler, trunk, heads); selection and packing (sigmoid, max-pool NMS,
threshold, top-K, sub-pixel offsets, (v+1)*127.5 quantisation; :173-269) run as vc_trainable_select
+ vc_sample_pack, which applies the descriptor head's final 1x1 convolution and the L2 normalisation
only at the selected keypoints.  Images are processed in batches; rows are written by a single
database writer in sorted file order.  No PyTorch fallback.
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import List, Optional, Tuple

import numpy as np

from ._pipelined import BatchPipeline, decode_ahead
from .base_extractor import BaseExtractor

PATCH = 14
IMAGE_EXTENSIONS = {".jpg", ".jpeg", ".png", ".bmp", ".tiff", ".tif", ".ppm"}  # trainable_vit_extractor.py:302


def split_checkpoint(checkpoint):
    """trainable_vit_extractor.py:100-110: accept {'model_state_dict': ...}, {'state_dict': ...} or a
    raw state dict; split ViTFeatureModel keys into (backbone state dict, head state dict)."""
    if isinstance(checkpoint, dict) and "model_state_dict" in checkpoint:
        sd = checkpoint["model_state_dict"]
    elif isinstance(checkpoint, dict) and "state_dict" in checkpoint:
        sd = checkpoint["state_dict"]
    else:
        sd = checkpoint
    backbone = {k[len("backbone."):]: v for k, v in sd.items() if k.startswith("backbone.")}
    head = {k: v for k, v in sd.items() if not k.startswith("backbone.")}
    return backbone, head


class TrainableViTExtractor(BatchPipeline, BaseExtractor):
    def __init__(self, weights_path: Optional[str] = None, model_name: str = "dinov2_vitb14",
                 num_keypoints: int = 2048, descriptor_dim: int = 128, device: Optional[str] = None,
                 score_threshold: float = 0.0, nms_radius: int = 4, *, state_dict=None, head_state_dict=None,
                 synthetic_seed: Optional[int] = None, batch_size: int = 16):
        import torch
        from .. import _lib
        from ..extraction import PostProcessor
        from ..head import ConvHead
        from ..vit import CONFIGS, VitBackbone
        self.weights_path = weights_path
        self.model_name = model_name
        self.num_keypoints = num_keypoints
        self.descriptor_dim = descriptor_dim
        self.score_threshold = score_threshold
        self.nms_radius = nms_radius
        self.batch_size = int(batch_size)
        self.patch_size = PATCH
        if not torch.cuda.is_available():
            raise _lib.VcError("TrainableViTExtractor needs an sm_100 GPU: vit_colmap_b200 has no CPU fallback")
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if dev.type != "cuda":
            raise _lib.VcError(f"device {dev} requested; vit_colmap_b200 runs on CUDA only")
        self.device = dev
        self.device_index = dev.index if dev.index is not None else torch.cuda.current_device()
        print(f"Initializing Trainable ViT extractor: {model_name} on {self.device}")
        if model_name not in CONFIGS:
            raise ValueError(f"Unsupported model: {model_name}. Currently only DINOv2 ViT-S/B/L/14 are supported.")
        backbone_sd, head_sd = self._load_weights(state_dict, head_state_dict, synthetic_seed)
        self.backbone = VitBackbone(backbone_sd, model_name, self.device_index)
        self.head = ConvHead(head_sd, self.backbone.embed_dim, descriptor_dim, self.device_index,
                             handle=self.backbone.handle)
        self.post = PostProcessor(self.device_index, handle=self.backbone.handle)
        print("✓ Model loaded successfully")

    # ------------------------------------------------------------------ weights
    def _load_weights(self, state_dict, head_state_dict, synthetic_seed):
        """Reference: ViTFeatureModel(...) = pretrained hub backbone + freshly initialised head, then
        load_state_dict(checkpoint, strict=False) (trainable_vit_extractor.py:89-112)."""
        import torch
        from ..head import synthetic_head_state_dict
        from ..vit import CONFIGS, synthetic_state_dict
        C = CONFIGS[self.model_name][0]
        backbone_sd, head_sd = state_dict, head_state_dict
        if self.weights_path is not None:
            print(f"Loading custom weights from: {self.weights_path}")
            ck_backbone, ck_head = split_checkpoint(torch.load(self.weights_path, map_location="cpu"))
            if ck_backbone and backbone_sd is None:
                backbone_sd = ck_backbone
            if ck_head:
                base = head_sd or synthetic_head_state_dict(C, self.descriptor_dim, synthetic_seed or 0)
                head_sd = {**base, **ck_head}  # strict=False: missing keys keep their initial values
            print("✓ Custom weights loaded")
        if backbone_sd is None:
            if synthetic_seed is not None:
                backbone_sd = synthetic_state_dict(self.model_name, synthetic_seed)
            else:
                names = [f"{self.model_name}_pretrain.pth", f"{self.model_name}.pth"]
                dirs = [os.environ.get("VIT_COLMAP_B200_WEIGHTS_DIR"), os.path.join(torch.hub.get_dir(), "checkpoints")]
                for d in dirs:
                    for n in names:
                        if backbone_sd is None and d and os.path.exists(os.path.join(d, n)):
                            backbone_sd = torch.load(os.path.join(d, n), map_location="cpu")
                if backbone_sd is None:
                    raise FileNotFoundError(
                        f"no DINOv2 checkpoint for {self.model_name}: put {names[0]} under "
                        f"$VIT_COLMAP_B200_WEIGHTS_DIR or {dirs[1]}, or pass state_dict= / synthetic_seed=")
        if head_sd is None:
            # the reference's head is randomly initialised when no checkpoint is given
            head_sd = synthetic_head_state_dict(C, self.descriptor_dim, synthetic_seed or 0)
        return backbone_sd, head_sd

    # ------------------------------------------------------------------ host preprocessing
    def _resize(self, image_bgr: np.ndarray) -> Tuple[np.ndarray, Tuple[int, int], Tuple[int, int]]:
        """trainable_vit_extractor.py:158-166 (cv2.INTER_LINEAR to a multiple of 14); BGR->RGB, ToTensor
        and Normalize are fused into the patchify kernel."""
        import cv2
        h, w = image_bgr.shape[:2]
        h_new, w_new = (h // PATCH) * PATCH, (w // PATCH) * PATCH
        if h_new == 0 or w_new == 0:
            raise ValueError(f"image {w}x{h} is smaller than one 14x14 patch")
        if (h_new, w_new) != (h, w):
            image_bgr = cv2.resize(image_bgr, (w_new, h_new), interpolation=cv2.INTER_LINEAR)
        return np.ascontiguousarray(image_bgr), (w, h), (w_new, h_new)

    # ------------------------------------------------------------------ batched core
    def _forward_device(self, batch_u8, orig_wh, resized_wh, want_f32: bool = False):
        """Device half: uint8 BGR batch [B, 14gh, 14gw, 3] -> (keypoints [B,cap,6], desc u8, desc bf16,
        desc f32|None, counts), all on the device."""
        from ..extraction import trainable_select_pack
        gh, gw = batch_u8.shape[1] // PATCH, batch_u8.shape[2] // PATCH
        tokens = self.backbone.forward_patch_tokens(batch_u8, bgr=True)
        maps = self.head.forward(tokens, (gh, gw), dense_descriptors=False)
        proj, bias = self.head.desc_projection
        return trainable_select_pack(self.post, maps["keypoints"], maps["desc_hidden"], orig_wh, resized_wh,
                                     self.num_keypoints, self.score_threshold, self.nms_radius, apply_tanh=False,
                                     want_f32=want_f32, projection=proj, projection_bias=bias)

    def _device_stage(self, batch, orig_wh, resized_wh, extra=None):
        kps, du8, dbf, _, counts = self._forward_device(batch, orig_wh, resized_wh)
        return kps, du8, counts, (counts, dbf)

    def _extract_batch(self, images_bgr: List[np.ndarray], orig_wh, resized_wh):
        return self._run_batch(images_bgr, orig_wh, resized_wh)

    def _run_inference(self, image_bgr: np.ndarray):
        """Single image -> (keypoints f32 [N,6] = x, y, 1, orientation, score, 0; descriptors u8 [N,d])
        (trainable_vit_extractor.py:140-269)."""
        img, orig_wh, resized_wh = self._resize(image_bgr)
        out, _ = self._extract_batch([img], orig_wh, resized_wh)
        return out[0]

    # ------------------------------------------------------------------ plugin entry
    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:
        import cv2
        from ..database import ColmapDatabase, make_camera
        image_dir = Path(image_dir)
        print(f"\n{'=' * 60}\nTrainable ViT Feature Extraction\n{'=' * 60}")
        print(f"Image directory: {image_dir}\nDatabase: {db_path}\nModel: {self.model_name}")
        print(f"Target keypoints per image: {self.num_keypoints}\nScore threshold: {self.score_threshold}")
        print(f"NMS radius: {self.nms_radius}\n{'=' * 60}\n")
        files = sorted(f for f in image_dir.iterdir() if f.suffix.lower() in IMAGE_EXTENSIONS) \
            if image_dir.exists() else []
        if not files:
            raise ValueError(f"No images found in {image_dir}")
        print(f"Found {len(files)} images")
        db = ColmapDatabase(str(db_path))
        first = cv2.imread(str(files[0]))
        if first is None:
            raise ValueError(f"Failed to read first image: {files[0]}")
        height, width = first.shape[:2]
        print(f"Image dimensions: {width}x{height}")
        if camera_params is None:  # trainable_vit_extractor.py:323-341
            f = max(width, height)
            defaults = {"SIMPLE_PINHOLE": [f, width / 2.0, height / 2.0],
                        "PINHOLE": [f, f, width / 2.0, height / 2.0],
                        "SIMPLE_RADIAL": [f, width / 2.0, height / 2.0, 0.0],
                        "RADIAL": [f, width / 2.0, height / 2.0, 0.0, 0.0]}
            if camera_model not in defaults:
                raise ValueError(f"Unsupported camera model: {camera_model}")
            camera_params = defaults[camera_model]
        print(f"Camera model: {camera_model}\nCamera params: {camera_params}")
        camera_id = db.db.write_camera(make_camera(camera_model, width, height, camera_params))
        print(f"Camera ID: {camera_id}\n")

        def prepare(path):  # worker threads: decode + resize (None: unreadable file)
            img = cv2.imread(str(path))
            return None if img is None else (img, self._resize(img))

        def items():
            for k, f, got in decode_ahead(files, prepare):
                if got is None:
                    print(f"[{k + 1}/{len(files)}] {f.name}: Warning: Failed to read image, skipping")
                    continue
                if isinstance(got, Exception) and not isinstance(got, ValueError):
                    raise got
                image_id = db.add_image(f.name, camera_id=camera_id)
                if isinstance(got, Exception):
                    print(f"  Error during feature extraction: {got}")
                    continue
                _, (img, orig_wh, resized_wh) = got
                yield (image_id, f.name), img, orig_wh, resized_wh, None

        def write(meta, kps, desc):
            image_id, name = meta
            print(f"{name}: extracted {len(kps)} keypoints, descriptor shape {desc.shape}")
            if len(kps) == 0:
                print("  Warning: No keypoints extracted")
                return
            print(f"  Score range: [{kps[:, 4].min():.3f}, {kps[:, 4].max():.3f}]")
            db.add_keypoints(image_id, kps)
            db.add_descriptors(image_id, desc)

        def on_error(k, e):  # per-image skip semantics, trainable_vit_extractor.py:381-386
            print(f"  Error during feature extraction: {e}")
            import traceback
            traceback.print_exception(type(e), e, e.__traceback__)

        self.extract_stream(items(), self.batch_size, write, on_error)
        db.commit()
        db.close()
        print(f"\n{'=' * 60}\nFeature extraction complete!\n{'=' * 60}\n")
