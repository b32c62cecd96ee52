import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu() -> bool:
    try:
        import ctypes

        cuda = ctypes.CDLL("libcuda.so.1")
        n = ctypes.c_int(0)
        return cuda.cuInit(0) == 0 and cuda.cuDeviceGetCount(ctypes.byref(n)) == 0 and n.value > 0
    except OSError:
        return False


HAS_GPU = _has_gpu()


def pytest_collection_modifyitems(config, items):
    if HAS_GPU:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def vertices_equal(a: np.ndarray, b: np.ndarray, tangent_ulp: int = 0):
    """Bit-exact comparison of two 64-byte vertex arrays; NaN matches NaN regardless of sign/payload
    (x86 produces 0xFFC00000, the GPU 0x7FFFFFFF).  `tangent_ulp` > 0 relaxes only tangent.xyz."""
    ua = np.ascontiguousarray(a).view(np.uint32).reshape(-1, 16)
    ub = np.ascontiguousarray(b).view(np.uint32).reshape(-1, 16)
    if ua.shape != ub.shape:
        return False, f"vertex count {ua.shape[0]} != {ub.shape[0]}"
    fa, fb = ua.view(np.float32), ub.view(np.float32)
    ok = (ua == ub) | (np.isnan(fa) & np.isnan(fb))
    if tangent_ulp:
        ia, ib = ua.astype(np.int64), ub.astype(np.int64)
        same_sign = (ia >> 31) == (ib >> 31)
        close = same_sign & (np.abs(ia - ib) <= tangent_ulp)
        ok[:, 12:15] |= close[:, 12:15]
    if ok.all():
        return True, ""
    r, c = np.argwhere(~ok)[0]
    return False, f"{(~ok).sum()} float mismatches; first at vertex {r} float {c}: {fa[r, c]!r} vs {fb[r, c]!r}"


def assert_shapes_equal(got, want, tangent_ulp: int = 0):
    assert len(got) == len(want), f"shape count {len(got)} != {len(want)}"
    for i, (g, w) in enumerate(zip(got, want)):
        assert len(g.indices) == len(w.indices), f"shape {i}: index count"
        assert np.array_equal(g.indices, w.indices), f"shape {i}: index buffer differs at {np.argwhere(g.indices != w.indices)[:3].ravel()}"
        ok, msg = vertices_equal(g.vertices, w.vertices, tangent_ulp)
        assert ok, f"shape {i}: {msg}"
        assert np.array_equal(np.asarray(g.extent, np.float32), np.asarray(w.extent, np.float32), equal_nan=True), \
            f"shape {i}: extent {g.extent} vs {w.extent}"
