"""b200rt — B200-native (sm_100a) backend for the per-pixel hot path of
JTtNinjaCode/CPU-Ray-Tracing-Implementation (camera::render's sample loop + ray_color)."""
from . import gltf  # noqa: F401
from . import _abi, camera, images, scene  # noqa: F401
from . import backend, distributed  # noqa: F401  (loading libb200rt.so itself is deferred to first use)
