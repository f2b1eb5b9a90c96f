"""delayed_speech_synthesis_b200 — the high-gamma-activity (HGA) front end of cronelab/delayed-speech-synthesis as
hand-written sm_100a CUDA kernels behind a C ABI (include/hga_b200.h), with the reference's Python call surface on top.

Importing the package does not need a GPU; constructing an extractor or calling a compute function does, and raises
when the CUDA library or device is missing (there is no CPU fallback).
"""
from . import filters  # noqa: F401
from ._lib import HgaError, lib_path, load  # noqa: F401

__all__ = ["HgaError", "lib_path", "load", "filters"]
__version__ = "0.1.0"
