"""radixmesh_b200 — B200-native radix-tree KV-cache index behind the reference's RadixCache / RadixMesh /
CacheAwareRouter surface.  The CUDA library (radixmesh_b200/_lib/libradixmesh_b200.so) is the product:
importing this package fails loudly if it has not been built; there is no CPU fallback."""
from . import capi

capi.lib()   # raises ImportError with build instructions if the shared library is missing

from .capi import DeviceIndex, RadixMeshError  # noqa: E402
from .host import CacheOplog, CacheOplogType, RadixMode, ServerArgs  # noqa: E402
from .radix_cache import MatchResult, RadixCache, TreeNode  # noqa: E402
from .radix_mesh import RadixMesh, RouterMatchResult  # noqa: E402
from .router import CacheAwareRouter, ConsistentHash, RouteResult  # noqa: E402

__all__ = ["DeviceIndex", "RadixMeshError", "RadixCache", "RadixMesh", "CacheAwareRouter", "ConsistentHash",
           "MatchResult", "RouterMatchResult", "RouteResult", "TreeNode", "ServerArgs", "CacheOplog",
           "CacheOplogType", "RadixMode"]
