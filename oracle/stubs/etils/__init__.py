"""TEST INFRASTRUCTURE ONLY. Stand-in for `etils` so `xunet.py:5-9` imports and its monkey patch
(`etils.enp.linalg._tf_or_xnp = ...`) succeeds."""
from . import enp  # noqa: F401
