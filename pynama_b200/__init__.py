"""pynama_b200 -- B200-native hot path of Pynama (KLE spectral-element assembly,
Krylov solve, vorticity-transport operators) behind the reference's Python API.

CUDA (sm_100a) lives in libpynama_b200.so, loaded through ctypes; see
include/pynama_b200.h, DESIGN.md and INTEGRATION.md.
"""
__version__ = "0.1.0"
