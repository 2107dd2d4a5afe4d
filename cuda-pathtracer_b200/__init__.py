"""B200-native path-tracing core behind the reference's ``raytrace()`` boundary.

Python here is plumbing only (fixture reader, ctypes binding of the C ABI in
``include/ptcore.h``, torch device buffers / torch.distributed for multi-GPU);
the product is ``csrc/`` -> ``libptcore.so``.
"""
from . import ptscene  # noqa: F401
from . import synth  # noqa: F401
from . import tiles  # noqa: F401
from . import core  # noqa: F401,E402
from .core import Core, Group, HostImage, PtcError, SharedBuffer, host_bvh, load_library  # noqa: F401,E402
