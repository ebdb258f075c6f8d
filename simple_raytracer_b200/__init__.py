"""B200-native render core for simple-raytracer scenes.

The product is `lib/librt_b200.so` (hand-written CUDA for sm_100a behind the C ABI
of include/rt.h) plus the host front-end `lib/librt_host.so`; this package is the
thin Python host that mirrors the reference's seam:

    Scene.from_file(path) -> DeviceScene(scene).render() -> .download() -> ppm.write_ppm
"""
from . import abi  # noqa: F401
from .scene import Scene  # noqa: F401

__all__ = ["abi", "Scene"]
