import numpy as _np


class _Lazy:
    def get_xnp(self, x=None, **kw):
        return _np


lazy = _Lazy()
