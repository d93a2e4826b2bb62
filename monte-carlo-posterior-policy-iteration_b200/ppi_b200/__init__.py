"""ppi_b200 -- B200-native posterior policy iteration update path.

Same module / class names as the reference's `ppi` package (algorithms, policies, samplers) plus
`functions` (optimization/functions.py).  Everything computes through libppi_b200.so
(hand-written sm_100a CUDA behind the C ABI of include/ppi_b200.h); importing the package does
not need a GPU, using it does -- there is no CPU fallback.
"""
from . import _lib  # noqa: F401

__all__ = ["algorithms", "policies", "samplers", "functions", "ops"]


def __getattr__(name):
    if name in __all__:
        import importlib

        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
