"""TEST INFRASTRUCTURE ONLY - runs the kernel SOURCE compiled for the host (tests/host_emul) behind the same
Go2Core interface, so that the CPU test-suite can exercise csrc/ and the host-side env logic without a GPU.
The package itself has no CPU path (go2_b200.core.Go2Core refuses non-CUDA devices)."""
import ctypes as C
import os

import numpy as np
import torch

from go2_b200.core import Go2Core

HOSTEMU_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul", "libgo2_hostemu.so")


class HostEmuCore(Go2Core):
    def __init__(self, cfg, device="cpu", model=None, lib_path=None):
        self._keep = []
        super().__init__(cfg, device="cpu", model=model, lib_path=lib_path or HOSTEMU_LIB)

    def _check_device(self):
        pass

    def _device_index(self):
        return 0

    def _view(self, ptr, shape, dtype):
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dtype))), shape=shape)
        self._keep.append(arr)
        return torch.from_numpy(arr)

    def _stream(self):
        return C.c_void_p(0)
