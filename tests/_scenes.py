"""Synthetic scenes shared by the GPU test modules (not a test module itself)."""
import numpy as np


def unit_rows(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def two_view_scene(rng, n, kmax, d, kind):
    """Two images of n keypoints related by a homography (kind 'H') or a fundamental matrix (kind 'F'):
    image 2 holds a permuted subset of image 1's descriptors at the geometrically consistent positions,
    plus distractors: duplicated descriptors at WRONG positions (only the gate can tell them apart)."""
    from oracle import matching as om
    kp1 = rng.uniform(20, 620, (n, 2)).astype(np.float32)
    if kind == "H":
        G = np.array([[1.02, 0.03, 7.0], [-0.02, 0.98, -4.0], [2e-5, -1e-5, 1.0]])
        p = np.c_[kp1, np.ones(n)] @ G.T
        kp2 = (p[:, :2] / p[:, 2:]).astype(np.float32)
        config = 4
    else:
        # pure sideways translation: F = [t]_x with t = (1, 0, 0)  ->  corresponding points share y
        G = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
        kp2 = kp1.copy()
        kp2[:, 0] += rng.uniform(-60, 60, n).astype(np.float32)
        config = 3
    f1 = unit_rows(rng, n, d)
    perm = rng.permutation(n)
    f2 = f1[perm] + (0.25 / d ** 0.5) * rng.standard_normal((n, d)).astype(np.float32)
    kp2 = kp2[perm]
    # distractors: rows of image 2 re-using another row's descriptor (almost) at an inconsistent place
    for k in range(0, n, 7):
        src = (k + 3) % n
        f2[k] = f2[src] + (0.02 / d ** 0.5) * rng.standard_normal(d).astype(np.float32)
        kp2[k] = kp2[src] + np.float32([37.0, 41.0])
    kp2 += rng.normal(0, 0.8, kp2.shape).astype(np.float32)  # residuals on both sides of max_error = 2 px
    u1 = om.to_uint8_colmap(om.l2_normalize(np.abs(f1)))
    u2 = om.to_uint8_colmap(om.l2_normalize(np.abs(f2)))
    desc = np.zeros((2, kmax, d), np.uint8)
    desc[0, :n], desc[1, :n] = u1, u2
    kps = np.zeros((2, kmax, 2), np.float32)
    kps[0, :n], kps[1, :n] = kp1, kp2
    return desc, kps, config, G
