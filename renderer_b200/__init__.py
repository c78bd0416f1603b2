"""renderer_b200 -- B200-native transform / cull / compact stage for
farnoy/renderer's per-frame hot path.

The compute lives in librcull.so (renderer_b200/csrc, C ABI in include/rcull.h);
this package is the thin host-side harness around it.  Importing the package
does not load the library; the first CullContext does, and fails loudly if the
shared object has not been built.
"""
from ._lib import CMD_DTYPE, MESH_DTYPE, PER_ENTITY, PER_MESH_INSTANCED, RcullError, load
from .context import CullContext, views_array

__all__ = [
    "CullContext", "RcullError", "views_array", "load",
    "CMD_DTYPE", "MESH_DTYPE", "PER_ENTITY", "PER_MESH_INSTANCED",
]
