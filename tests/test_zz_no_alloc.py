"""go2_mlp_forward_noalloc (include/go2_policy.h): the MLP forward for callers that cannot allocate - packed weights live in a
file-scope `static __device__` array of 1 MiB instead of cudaMalloc'ed buffers.  No counterpart in the reference.
Checked through the C ABI against the existing entry points go2_mlp_create + go2_mlp_forward on
the same parameters and inputs, on calls in a row that share the array (wide net, then a narrow one whose zero padding overlaps the
first one's weights, then the wide one again).  Tolerance: 2e-3 absolute, the same-arithmetic bound of tests/test_policy_gpu.py
(both paths run the same bf16 x bf16 -> fp32 kernel; in practice the outputs are bit-identical).

Collected last on purpose: a failure here must not hide another test.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 2e-3


def _lib():
    from go2_b200 import abi

    lib = abi.Go2Lib(abi.CUDA_LIB_PATH).lib  # raises if the CUDA library is missing: no fallback
    lib.go2_mlp_create.restype = C.c_int
    lib.go2_mlp_forward.restype = C.c_int
    lib.go2_mlp_forward_noalloc.restype = C.c_int
    lib.go2_mlp_forward_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
    return lib


def _params(dims, seed):
    rng = np.random.default_rng(seed)
    parts = []
    for i in range(len(dims) - 1):
        bound = 1.0 / np.sqrt(dims[i])
        parts.append(rng.uniform(-bound, bound, dims[i + 1] * dims[i]).astype(np.float32))  # Linear.weight [out][in]
        parts.append(rng.uniform(-bound, bound, dims[i + 1]).astype(np.float32))  # Linear.bias
    return np.concatenate(parts)


def _existing(lib, dims, flat, x):
    """The existing entry points: host parameters packed by go2_mlp_create, go2_mlp_forward."""
    h = C.c_void_p()
    cd = (C.c_int32 * len(dims))(*dims)
    assert lib.go2_mlp_create(cd, len(dims) - 1, flat.ctypes.data_as(C.c_void_p), 0, C.byref(h)) == 0
    out = torch.full((x.shape[0], dims[-1]), float("nan"), device="cuda:0")
    st = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    assert lib.go2_mlp_forward(h, C.c_void_p(x.data_ptr()), x.shape[0], C.c_void_p(out.data_ptr()), st) == 0
    torch.cuda.synchronize()
    lib.go2_mlp_destroy(h)
    return out


def _noalloc(lib, dims, d_flat, x, out):
    err = C.create_string_buffer(256)
    cd = (C.c_int32 * len(dims))(*dims)
    st = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    rc = lib.go2_mlp_forward_noalloc(C.cast(cd, C.c_void_p), len(dims) - 1, C.c_void_p(d_flat.data_ptr()), 0, C.c_void_p(x.data_ptr()), x.shape[0],
                                     C.c_void_p(out.data_ptr()), st, err, 256)
    return rc, err.value.decode()


def test_noalloc_forward_matches_existing_entry_point_on_calls_in_a_row():
    lib = _lib()
    torch.cuda.set_device(0)
    cases = [([104, 512, 256, 128, 1], 300, 1), ([49, 128, 16], 77, 2), ([104, 512, 256, 128, 1], 300, 1), ([49, 512, 256, 128, 12], 1000, 3)]
    # every buffer the calls touch exists before the first call: the entry point itself allocates nothing
    work = []
    for dims, B, seed in cases:
        flat = _params(dims, seed)
        x = torch.randn(B, dims[0], device="cuda:0", generator=torch.Generator(device="cuda:0").manual_seed(seed))
        work.append((dims, flat, torch.as_tensor(flat, device="cuda:0"), x, torch.full((B, dims[-1]), float("nan"), device="cuda:0")))
    refs = [_existing(lib, dims, flat, x) for dims, flat, _, x, _ in work]
    # set-up call (first use on the device: symbol address + kernel attribute), then the calls in a row
    rc, msg = _noalloc(lib, *[work[0][i] for i in (0, 2, 3, 4)])
    assert rc == 0, msg
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for dims, flat, d_flat, x, out in work:
        out.fill_(float("nan"))
        rc, msg = _noalloc(lib, dims, d_flat, x, out)
        assert rc == 0, msg
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0, "device memory in use changed across no-alloc calls"
    for (dims, _, _, _, out), ref in zip(work, refs):
        assert torch.isfinite(out).all() and torch.isfinite(ref).all()
        err = float((out - ref).abs().max())
        assert err <= TOL, (dims, err)
    # and against plain fp32 torch at the kernel's stated bf16 tolerance (tests/test_policy_gpu.py)
    dims, flat, _, x, out = work[3]
    h, p = x, 0
    for i in range(len(dims) - 1):
        W = torch.as_tensor(flat[p:p + dims[i + 1] * dims[i]].reshape(dims[i + 1], dims[i]), device="cuda:0")
        p += dims[i + 1] * dims[i]
        b = torch.as_tensor(flat[p:p + dims[i + 1]], device="cuda:0")
        p += dims[i + 1]
        h = h @ W.T + b
        if i + 2 < len(dims):
            h = torch.nn.functional.elu(h)
    assert ((out - h).abs() <= 3e-2 + 3e-2 * h.abs()).all()


def test_noalloc_rejects_networks_beyond_the_static_array():
    lib = _lib()
    torch.cuda.set_device(0)
    dims = [512, 512, 512, 512, 16]  # 1.6 MB of packed weights > GO2_MLP_NOALLOC_SCRATCH_BYTES = 1 MiB
    d_flat = torch.zeros(sum(dims[i] * dims[i + 1] + dims[i + 1] for i in range(4)), device="cuda:0")
    x = torch.zeros(4, 512, device="cuda:0")
    out = torch.full((4, 16), 7.0, device="cuda:0")
    rc, msg = _noalloc(lib, dims, d_flat, x, out)
    torch.cuda.synchronize()
    assert rc == 8 and "scratch" in msg and str(1 << 20) in msg, (rc, msg)
    assert bool((out == 7.0).all())  # nothing was launched
    rc, msg = _noalloc(lib, [49, 600, 16], d_flat, x, out)
    assert rc == 2 and "600" in msg, (rc, msg)
