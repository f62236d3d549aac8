"""henjou-renderer_b200 — B200-native (sm_100a) path tracer behind the C ABI of include/henjou_b200.h.

Python here is a thin driver for tests and bench: ctypes over libhenjou_b200.so (CUDA kernels + C ABI,
csrc/) and numpy containers for the reference's SceneData. There is no CPU fallback: creating a Renderer
without the built CUDA library or without a B200 raises."""
from . import abi                     # noqa: F401
from .scene import SceneData, NodeAnimation, make_material   # noqa: F401
from .camera import camera_basis, frame_params                # noqa: F401
from .sceneio import save_npz, load_npz                       # noqa: F401
from .renderer import Renderer, HenjouError, load_library, LIB_PATH   # noqa: F401
from .slices import sample_slice, weak_slice   # noqa: F401
from . import hdrio                  # noqa: F401
