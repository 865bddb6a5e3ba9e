"""TEST INFRASTRUCTURE ONLY. Stand-in for the third-party module `visu3d` (PyPI; version NOT
pinned by the reference -- no requirements/lock file), restricted to the four call sites of
the reference: xunet.py:311 `v3d.Transform(R=, t=)`, :312 `v3d.PinholeCamera(resolution=, K=)`,
:313-314 `v3d.Camera(spec=, world_from_cam=).rays()`, consumed as `.pos`/`.dir` at :317-318.

PARITY UNPINNED at this boundary: visu3d is absent from /root/reference and from this image,
and the reference holds no test vector for it.  The arithmetic lives in oracle/rays.py
(restated from visu3d's published semantics)."""
import numpy as np

from oracle import rays as _rays


class Transform:
    def __init__(self, R, t):
        self.R = np.asarray(R)
        self.t = np.asarray(t)


class PinholeCamera:
    def __init__(self, resolution, K):
        self.resolution = tuple(resolution)
        self.K = np.asarray(K)


class _Rays:
    def __init__(self, pos, dir):
        self.pos = pos
        self.dir = dir


class Camera:
    def __init__(self, spec, world_from_cam):
        self.spec = spec
        self.world_from_cam = world_from_cam

    def rays(self):
        H, W = self.spec.resolution
        K = self.spec.K                      # [B,1,3,3]  (xunet.py:312 `K.unsqueeze(1)`)
        R = self.world_from_cam.R            # [B,2,3,3]
        t = self.world_from_cam.t            # [B,2,3]
        K = np.broadcast_to(K, R.shape)
        pos, dir = _rays.camera_rays(K, R, t, H, W)
        return _Rays(pos, dir)
