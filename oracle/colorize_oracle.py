"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's depth-image colouriser.

    utils.py:81-91   deep_to_img(deep_img[w,w,2]) -> uint8 [w,w,3]
    train_3d_match.py:212-213 / train_kitti.py:211-212 / visualize_depth_image.py:27-28
                     np.concatenate([deep_to_img(imgs[j]) for j in range(V)], axis=1) on imgs.permute(0,2,3,1)

numpy fp32, one rounding per reference op.  Pinned against the reference's own function: oracle/make_golden.py
executes the unmodified source of `deep_to_img` (extracted from /root/reference/utils.py, whose module-level
imports need open3d and cannot run here) and freezes tests/golden/colorize_strip.npz.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import numpy as np

COL_CH1 = np.array([1, 0.706, 0], dtype=np.float32)       # utils.py:87  (channel 1 won)
COL_CH0 = np.array([0, 0.651, 0.929], dtype=np.float32)   # utils.py:88


def deep_to_img(deep_img: np.ndarray) -> np.ndarray:
    """deep_img [w,w,2] fp32 -> uint8 [w,w,3] (utils.py:81-91)."""
    d = deep_img.astype(np.float32) * np.float32(0.5) + np.float32(0.5)                  # :83
    ind = (d[..., 1] > d[..., 0]).astype(np.float32)                                     # :84 first index on ties
    img = np.maximum(d[..., 0], d[..., 1])
    a = (img * ind)[..., None] * COL_CH1                                                 # :85,87
    b = (img * (np.float32(1) - ind))[..., None] * COL_CH0                               # :86,88
    return ((a + b) * np.float32(255)).astype(np.uint8)                                  # :89


def views_to_strip(imgs: np.ndarray) -> np.ndarray:
    """imgs [V,2,w,w] -> uint8 [w, V*w, 3]."""
    return np.concatenate([deep_to_img(np.transpose(imgs[j], (1, 2, 0))) for j in range(imgs.shape[0])], axis=1)
