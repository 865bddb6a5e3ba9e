"""ctypes loader of the in-tree C-ABI library (libb3d.so).  There is no fallback: if the library is
missing or a call fails, the product path raises."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B3D_LIB") or os.path.join(_HERE, "libb3d.so")      # (B3D_LIB: A/B measurements against another build)
_lib = None


class B3DError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B3DError(f"{LIB_PATH} not built: run `python __graft_entry__.py` (there is no CPU/eager fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.b3d_last_error.restype = ctypes.c_char_p
        _lib.b3d_launch_count.restype = ctypes.c_int64
    return _lib


def check(rc, what):
    if rc != 0:
        raise B3DError(f"{what} failed ({rc}): {load().b3d_last_error().decode()}")


def launch_count():
    return int(load().b3d_launch_count())


def set_pdl(on):
    """programmatic dependent launch of the frequent kernels (include/b3d.h: b3d_set_pdl)"""
    load().b3d_set_pdl(ctypes.c_int(1 if on else 0))


def set_gn_chunk_mb(mb):
    """per-chunk budget of the image-chunked GroupNorm launches (include/b3d.h: b3d_set_gn_chunk_mb); 0 = off"""
    load().b3d_set_gn_chunk_mb(ctypes.c_int(int(mb)))


def set_gn_fuse_mb(mb):
    """largest input (MB) for which GroupNorm runs as one cluster-fused launch (include/b3d.h: b3d_set_gn_fuse_mb); 0 = never"""
    load().b3d_set_gn_fuse_mb(ctypes.c_int(int(mb)))


def set_gn_fuse_bwd_mb(mb):
    """the same threshold for the GroupNorm backward (0 by default: the fused backward measured slower)"""
    load().b3d_set_gn_fuse_bwd_mb(ctypes.c_int(int(mb)))


def set_conv_halo(on):
    """halo-patch form of the stride-1 3x3 convolutions (include/b3d.h: b3d_set_conv_halo)"""
    load().b3d_set_conv_halo(ctypes.c_int(int(on)))
