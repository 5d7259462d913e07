"""sph_b200 -- B200-native SPH step engine behind the Particle/Wall/SphUtils call surface of Dimitime/SPH.

The product is libsphb200.so (hand-written CUDA for sm_100a behind the C ABI of include/sphb200.h).
This package is the Python side of that boundary: ctypes bindings (capi), a mirror of the reference's
SphUtils interface (sphutils) and synthetic scenes.  Importing the package does not need a GPU; any
compute call does, and fails loudly without the library -- there is no CPU fallback.
"""
from .capi import (DOMAIN_BBOX, DOMAIN_FIXED, F_ACCEL, F_ALL, F_DENSITY, F_POS, F_VEL, PARTICLE, SPHERE, WALL, Engine,
                   Params, SphbError, default_params, load_library)
from .sphutils import SphUtils

__all__ = ["Engine", "Params", "SphUtils", "SphbError", "default_params", "load_library", "PARTICLE", "WALL", "SPHERE",
           "F_POS", "F_VEL", "F_ACCEL", "F_DENSITY", "F_ALL", "DOMAIN_BBOX", "DOMAIN_FIXED"]
