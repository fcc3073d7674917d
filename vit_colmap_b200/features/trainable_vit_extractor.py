"""``trainable_vit`` switch value (reference: vit_colmap/features/trainable_vit_extractor.py:18-395).

Status: the selection / packing stage of this extractor (sigmoid score, max-pool NMS, threshold,
top-K, sub-pixel offsets, direct descriptor gather, (v+1)*127.5 quantisation) and the convolutional
head of ViTFeatureModel (vit_colmap/model/vit_feature_model.py:173-232) are not built yet; the class
keeps the reference's constructor signature and fails loudly instead of falling back to PyTorch."""
from pathlib import Path
from typing import Optional

from .base_extractor import BaseExtractor


class TrainableViTExtractor(BaseExtractor):
    def __init__(self, weights_path: Optional[str] = None, model_name: str = "dinov2_vitb14",
                 num_keypoints: int = 2048, descriptor_dim: int = 128, device: Optional[str] = None,
                 score_threshold: float = 0.0, nms_radius: int = 4):
        self.weights_path = weights_path
        self.model_name = model_name
        self.num_keypoints = num_keypoints
        self.descriptor_dim = descriptor_dim
        self.score_threshold = score_threshold
        self.nms_radius = nms_radius
        raise NotImplementedError(
            "trainable_vit: the ViTFeatureModel conv head has no sm_100a implementation yet "
            "(DESIGN.md, 'not built'); there is no PyTorch fallback")

    def extract(self, image_dir: Path, db_path: Path, camera_model: str,
                camera_params: Optional[list[float]] = None) -> None:  # pragma: no cover
        raise NotImplementedError
