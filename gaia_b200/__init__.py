"""gaia_b200 — B200-native Vertex Block Descent step behind Gaia's VBDPhysics interface.

The product is the CUDA library gaia_b200/csrc/libgaia_b200.so (C ABI: include/gaia_b200.h) plus the C++ host
adapter gaia_b200/host/. This package is the Python-side binding and scene helpers used by tests and bench.
"""
from .capi import (Scene, GaiaError, PhysicsParams, CollisionParams, ObjectParams, default_physics_params,
                   default_collision_params, default_object_params, load_library)

__all__ = ["Scene", "GaiaError", "PhysicsParams", "CollisionParams", "ObjectParams", "default_physics_params",
           "default_collision_params", "default_object_params", "load_library"]
