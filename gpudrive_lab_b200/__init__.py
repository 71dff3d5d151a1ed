"""gpudrive_lab_b200 — B200-native step engine behind the `madrona_gpudrive.SimManager` surface.

Only what the hot path needs lives here: csrc/ (CUDA kernels + C ABI), the ctypes binding, the
`madrona_gpudrive` stand-in module, a host-only scenario loader check and the synthetic
Waymo-shaped scenario generator used by the tests and the benchmark.
"""
import sys as _sys

from . import madrona_gpudrive  # noqa: F401


def install():
    """Register the stand-in as `madrona_gpudrive` so `gpudrive/env` imports it unchanged."""
    _sys.modules["madrona_gpudrive"] = madrona_gpudrive
    _sys.modules["madrona_gpudrive.madrona"] = madrona_gpudrive.madrona
    return madrona_gpudrive
