"""Device versions of the reference's training-time samplers (vit_colmap/dataloader): similarity
after homography warping, invariant-point selection, hard-negative mining.
The training loop itself (datasets, losses, optimiser) is out of scope."""
from .homography_utils import compute_feature_similarity  # noqa: F401
from .training_sampler import TrainingSampler  # noqa: F401
