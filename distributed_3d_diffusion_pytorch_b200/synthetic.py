"""Synthetic SRN-cars-shaped inputs (no dataset on the box).  Mirrors the *output contract* of the
reference loader (SRNdataset.py:90-95): imgs f32 [B,2,3,S,S] in [-1,1], R f64 [B,2,3,3],
T f64 [B,2,3], K f64 [B,3,3]; K is not rescaled with image size (SRNdataset.py:68-79)."""
import numpy as np
import torch

SRN_K = [[131.25, 0.0, 64.0], [0.0, 131.25, 64.0], [0.0, 0.0, 1.0]]
SRN_RADIUS = 1.3


def look_at_poses(n, rng, radius=SRN_RADIUS):
    """Camera centres uniform on the sphere, looking at the origin, world up = +z, OpenCV axes
    (x right, y down, z forward); R = [x|y|z] columns (world_from_cam), t = centre."""
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=-1, keepdims=True)
    c = v * radius
    z = -c / np.linalg.norm(c, axis=-1, keepdims=True)
    up = np.array([0.0, 0.0, 1.0])
    x = np.cross(z, up)
    bad = np.linalg.norm(x, axis=-1) < 1e-6
    x[bad] = np.array([1.0, 0.0, 0.0])
    x /= np.linalg.norm(x, axis=-1, keepdims=True)
    y = np.cross(z, x)
    R = np.stack([x, y, z], -1)
    return R, c


def make_views(B, size, seed=0, device="cpu"):
    """-> img [B,2,3,S,S] f32, R [B,2,3,3] f64, T [B,2,3] f64, K [B,3,3] f64."""
    rng = np.random.default_rng(seed)
    img = torch.from_numpy(rng.random((B, 2, 3, size, size), dtype=np.float32) * 2 - 1)
    R, c = look_at_poses(B * 2, rng)
    R = torch.from_numpy(R.reshape(B, 2, 3, 3))
    T = torch.from_numpy(c.reshape(B, 2, 3))
    K = torch.tensor(SRN_K, dtype=torch.float64)[None].repeat(B, 1, 1)
    return img.to(device), R.to(device), T.to(device), K.to(device)
