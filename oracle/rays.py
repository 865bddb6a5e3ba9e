"""TEST INFRASTRUCTURE ONLY -- numpy restatement of `visu3d.Camera(spec, world_from_cam).rays()`
as the reference uses it (xunet.py:311-318).  Never imported by the product package.

PARITY UNPINNED: the arithmetic belongs to the un-vendored third-party module `visu3d`
(version not pinned by the reference); semantics restated from its published behaviour
(visu3d >= 1.3): OpenCV camera frame (x right, y down, z forward), pixel coords (u,v)=(col,row),
pixel centres at +0.5, float32 outputs, directions normalised.

    p      = [u+0.5, v+0.5, 1]
    d_cam  = inv(K) @ p
    d      = R @ d_cam ;  dir = d/||d|| ;  pos = t (broadcast over H,W)
"""
import numpy as np

# one-line switch recorded in DESIGN.md: 'uv' = (col,row) pixel convention (visu3d >= 1.3)
PIXEL_CONVENTION = "uv"


def camera_rays(K, R, t, H, W):
    """K,R: [...,3,3]; t: [...,3]  ->  pos, dir float32 [...,H,W,3]."""
    K = np.asarray(K, dtype=np.float64)
    R = np.asarray(R, dtype=np.float64)
    t = np.asarray(t, dtype=np.float64)
    v, u = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64), indexing="ij")
    if PIXEL_CONVENTION == "uv":
        p = np.stack([u + 0.5, v + 0.5, np.ones_like(u)], -1)       # [H,W,3]
    else:  # legacy (i,j)=(row,col)
        p = np.stack([v + 0.5, u + 0.5, np.ones_like(u)], -1)
    Kinv = np.linalg.inv(K)                                          # [...,3,3]
    M = R @ Kinv                                                     # world_from_px (direction part)
    d = np.einsum("...ij,hwj->...hwi", M, p)
    d = d / np.linalg.norm(d, axis=-1, keepdims=True)
    pos = np.broadcast_to(t[..., None, None, :], d.shape)
    return pos.astype(np.float32), d.astype(np.float32)
