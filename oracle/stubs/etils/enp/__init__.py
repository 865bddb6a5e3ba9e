from . import numpy_utils, linalg  # noqa: F401
