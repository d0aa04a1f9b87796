"""Handle + zero-copy tensor views over the native library's buffer table.

The library owns all memory (one device arena per handle); `Go2Core.buf[name]` are non-owning torch
views built through `__cuda_array_interface__` (lifetime = this object).  Every call is enqueued on
torch's current CUDA stream.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import abi
from .model import flatten_model, load_model


class _CudaBlob:
    """Minimal __cuda_array_interface__ carrier so torch.as_tensor() can alias library-owned memory."""

    def __init__(self, ptr, shape, dtype):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": np.dtype(dtype).str, "data": (int(ptr), False), "version": 2, "strides": None,
        }


class Go2Core:
    def __init__(self, cfg: abi.EnvConfig, device="cuda:0", model=None, lib_path=None):
        self.device = torch.device(device)
        self.on_gpu = self.device.type == "cuda"
        if lib_path is None:
            if not self.on_gpu:
                raise RuntimeError("go2_b200 runs on CUDA devices only (no CPU fallback); got device=%r" % (device,))
            lib_path = abi.CUDA_LIB_PATH
        self.L = abi.Go2Lib(lib_path)
        self.lib = self.L.lib
        self.model = model or load_model()
        iv, fv = flatten_model(self.model)
        self.cfg = cfg
        dev_index = (self.device.index or 0) if self.on_gpu else 0
        if self.on_gpu:
            torch.cuda.init()
            torch.zeros(1, device=self.device)  # make sure the primary context exists before the library uses it
        self.h = self.L.create(iv, fv, cfg, dev_index)
        self.B = cfg.num_envs
        self.buf = {}
        self._keep = []
        for name, bid in abi.BUF.items():
            ptr, dt, r, c = self.L.buffer_info(self.h, bid)
            if self.on_gpu:
                t = torch.as_tensor(_CudaBlob(ptr, (r, c), dt), device=self.device)
            else:
                arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(r, c))
                self._keep.append(arr)
                t = torch.from_numpy(arr)
            self.buf[name] = t
        self._stats = np.zeros(self.L.stats_len, dtype=np.float32)

    def close(self):
        if getattr(self, "h", None):
            self.buf.clear()
            self.lib.go2_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream) if self.on_gpu else C.c_void_p(0)

    def step(self, actions: torch.Tensor, sp: abi.StepParams):
        a = actions
        if a.dtype != torch.float32 or not a.is_contiguous() or a.device != self.device:
            a = a.to(device=self.device, dtype=torch.float32).contiguous()
        self._last_actions_ref = a  # keep alive until the next call
        self.L.check(self.h, self.lib.go2_step(self.h, C.c_void_p(a.data_ptr()), C.byref(sp), C.sizeof(sp), self._stream()), "go2_step")

    def step_host(self, h_actions, sp, h_obs, h_rew, h_done, h_priv=None, h_timeouts=None):
        def p(x):
            return C.c_void_p(x.ctypes.data) if isinstance(x, np.ndarray) else (C.c_void_p(x.data_ptr()) if x is not None else C.c_void_p(0))

        self.L.check(self.h, self.lib.go2_step_host(self.h, p(h_actions), C.byref(sp), C.sizeof(sp), p(h_obs), p(h_rew), p(h_done),
                                                    p(h_priv), p(h_timeouts), self._stream()), "go2_step_host")

    def reset_all(self, sp: abi.StepParams):
        self.L.check(self.h, self.lib.go2_reset_all(self.h, C.byref(sp), C.sizeof(sp), self._stream()), "go2_reset_all")

    def read_stats(self):
        self.L.check(self.h, self.lib.go2_read_stats(self.h, self._stats.ctypes.data_as(C.c_void_p), len(self._stats), self._stream()),
                     "go2_read_stats")
        return self._stats

    def export_links(self, sp: abi.StepParams):
        self.L.check(self.h, self.lib.go2_export_links(self.h, C.byref(sp), C.sizeof(sp), self._stream()), "go2_export_links")

    def set_terrain(self, hf_m, h_scale, origin):
        """Heightfield terrain (metres, [rows][cols]) replacing the ground box - go2_set_terrain."""
        hf = np.ascontiguousarray(hf_m, dtype=np.float32)
        org = np.asarray(origin, dtype=np.float32)
        self._terrain_keep = (hf, org)
        self.L.check(self.h, self.lib.go2_set_terrain(self.h, hf.ctypes.data_as(C.c_void_p), int(hf.shape[0]), int(hf.shape[1]),
                                                      C.c_float(float(h_scale)), org.ctypes.data_as(C.c_void_p)), "go2_set_terrain")

    def launch_count(self):
        return int(self.lib.go2_launch_count(self.h))
