"""llama.cpp.dsp_b200 -- B200-native (sm_100a) quantized MUL_MAT / MUL_MAT_ID hot path of the llama.cpp fork
DreamChaser-luzeyu/llama.cpp.dsp, behind a C ABI (include/b200_qmm.h).

This Python layer is plumbing only: it loads `lib/libb200qmm.so` with ctypes and passes raw device pointers of torch
tensors (device memory, streams) to the C entry points.  There is NO CPU fallback and no torch implementation of any
op: if the shared library is missing or no sm_100 device is present, calls raise.

The directory name contains dots, so import it through the repo-root shim:  `import b200pkg; qmm = b200pkg.load()`.
"""
from . import qmm  # noqa: F401
from .qmm import *  # noqa: F401,F403
