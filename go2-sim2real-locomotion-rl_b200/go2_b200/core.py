"""Handle + zero-copy tensor views over the native library's buffer table.

The library owns all memory (one device arena per handle); `Go2Core.buf[name]` are non-owning torch
views built through `__cuda_array_interface__` (lifetime = this object).  Every call is enqueued on
torch's current CUDA stream of the handle's device.  CUDA only: there is no CPU execution path in the
package (the test-suite's host emulation of the kernel source lives under tests/).
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
        self._check_device()
        # lib_path: another build of the SAME CUDA library (kernel-variant A/B runs); default = the in-tree one
        self.L = abi.Go2Lib(lib_path or abi.CUDA_LIB_PATH)
        self.lib = self.L.lib
        self.model = model or load_model()
        iv, fv = flatten_model(self.model)
        self.cfg = cfg
        self.h = self.L.create(iv, fv, cfg, self._device_index())
        self.B = cfg.num_envs
        self.buf = {}
        for name, bid in abi.BUF.items():
            ptr, dt, r, c = self.L.buffer_info(self.h, bid)
            self.buf[name] = self._view(ptr, (r, c), dt)
        self._stats = np.zeros(self.L.stats_len, dtype=np.float32)

    # ---- device plumbing (overridden only by the test-suite's host emulation)
    def _check_device(self):
        if self.device.type != "cuda":
            raise RuntimeError("go2_b200 runs on CUDA devices only (no CPU fallback); got device=%r" % (self.device,))
        torch.cuda.init()
        torch.zeros(1, device=self.device)  # make sure the primary context exists before the library uses it

    def _device_index(self):
        return self.device.index if self.device.index is not None else torch.cuda.current_device()

    def _view(self, ptr, shape, dtype):
        return torch.as_tensor(_CudaBlob(ptr, shape, dtype), device=self.device)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

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

    # ---- the tick
    def step(self, actions: torch.Tensor, sp: abi.StepParams):
        a = actions
        if a.numel() != self.B * self.cfg.num_actions:  # the kernel reads num_actions floats per env: a short tensor would be read past its end
            raise ValueError(f"actions must be [{self.B}, {self.cfg.num_actions}], got {tuple(a.shape)}")
        if a.dtype != torch.float32 or not a.is_contiguous() or a.device != self.device:
            a = a.to(device=self.device, dtype=torch.float32).contiguous()
        self._last_actions_ref = a  # keep alive until the next call
        self.L.check(self.h, self.lib.go2_step(self.h, C.c_void_p(a.data_ptr()), C.byref(sp), C.sizeof(sp), self._stream()), "go2_step")

    def post(self, sp: abi.StepParams):
        """Deferred post-pass of the tick launched with sp.defer_post = 1 (multi-rank statistics reduction in between)."""
        self.L.check(self.h, self.lib.go2_post(self.h, C.byref(sp), C.sizeof(sp), self._stream()), "go2_post")

    def tick_parity(self):
        return int(self.lib.go2_tick_parity(self.h))

    def step_host(self, h_actions, sp, h_obs, h_rew, h_done, h_priv=None, h_timeouts=None):
        def p(x):
            return C.c_void_p(x.ctypes.data) if isinstance(x, np.ndarray) else (C.c_void_p(x.data_ptr()) if x is not None else C.c_void_p(0))

        self.L.check(self.h, self.lib.go2_step_host(self.h, p(h_actions), C.byref(sp), C.sizeof(sp), p(h_obs), p(h_rew), p(h_done),
                                                    p(h_priv), p(h_timeouts), self._stream()), "go2_step_host")

    def step_host_begin(self, h_actions, sp, h_obs, h_rew, h_done):
        """H2D actions + step kernel (sp.defer_post decides about the post-pass); returns True if obs / rew / done still have to
        be copied from the device buffers (pageable host buffers), False if the kernel writes them itself (pinned buffers)."""
        def p(x):
            return C.c_void_p(x.ctypes.data) if isinstance(x, np.ndarray) else C.c_void_p(x.data_ptr())

        staged = C.c_int(1)
        self.L.check(self.h, self.lib.go2_step_host_begin(self.h, p(h_actions), C.byref(sp), C.sizeof(sp), p(h_obs), p(h_rew), p(h_done),
                                                          self._stream(), C.byref(staged)), "go2_step_host_begin")
        return bool(staged.value)

    def reset_all(self, sp: abi.StepParams):
        self.L.check(self.h, self.lib.go2_reset_all(self.h, C.byref(sp), C.sizeof(sp), self._stream()), "go2_reset_all")

    # ---- statistics accumulated on the device since the previous read
    def read_stats(self):
        self.L.check(self.h, self.lib.go2_read_stats(self.h, self._stats.ctypes.data_as(C.c_void_p), len(self._stats), self._stream()),
                     "go2_read_stats")
        return self._stats

    def read_stats_async(self, slot):
        self.L.check(self.h, self.lib.go2_read_stats_async(self.h, int(slot), self._stream()), "go2_read_stats_async")

    def stats_fetch(self, slot, block=True):
        """The numbers of an earlier read_stats_async(slot); None if block=False and the copy has not completed."""
        rc = self.lib.go2_stats_fetch(self.h, int(slot), self._stats.ctypes.data_as(C.c_void_p), len(self._stats), 1 if block else 0)
        if rc == 1:
            return None
        self.L.check(self.h, rc, "go2_stats_fetch")
        return self._stats

    def export_links(self):
        self.L.check(self.h, self.lib.go2_export_links(self.h, self._stream()), "go2_export_links")

    def set_terrain(self, hf_m, h_scale, origin):
        """Heightfield terrain (metres, [rows][cols]) replacing the ground box - go2_set_terrain."""
        hf = np.ascontiguousarray(hf_m, dtype=np.float32)
        org = np.asarray(origin, dtype=np.float32)
        self._terrain_keep = (hf, org)
        self.L.check(self.h, self.lib.go2_set_terrain(self.h, hf.ctypes.data_as(C.c_void_p), int(hf.shape[0]), int(hf.shape[1]),
                                                      C.c_float(float(h_scale)), org.ctypes.data_as(C.c_void_p)), "go2_set_terrain")

    def launch_count(self):
        return int(self.lib.go2_launch_count(self.h))

    def launch_shape(self):
        out = (C.c_int * 3)()
        self.lib.go2_launch_shape(self.h, out)
        return {"blocks": out[0], "threads": out[1], "small_kernel": bool(out[2])}
