"""CPU tests of the drop-in boundary: the C-ABI library builds, loads, exports every symbol that
include/vkmesh_b200.h declares, and refuses to compute without a CUDA device (there is no CPU path)."""
import ctypes
import os
import re

import numpy as np
import pytest
from conftest import HAS_GPU, ROOT

from vulkanengine_b200 import capi


def header_symbols():
    text = open(os.path.join(ROOT, "include", "vkmesh_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vkm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = capi.lib()
    declared = header_symbols()
    assert declared, "no declarations parsed"
    assert sorted(capi.SYMBOLS) == declared
    raw = ctypes.CDLL(capi.lib_path())
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in include/vkmesh_b200.h but not exported"
    assert b"sm_100a" in L.vkm_version()


def test_struct_layouts_match_reference_types():
    assert capi.VERTEX_DTYPE.itemsize == 64  # sizeof(Vertex3D), renderer_types.hpp:13-19
    assert ctypes.sizeof(capi.Extent) == 36  # GeometryExtent, mesh_resource.hpp:8-12
    off = {n: capi.VERTEX_DTYPE.fields[n][1] for n in capi.VERTEX_DTYPE.names}
    assert off == {"position": 0, "normal": 12, "texcoord": 24, "color": 32, "tangent": 48}


def test_cuda_sass_is_sm100a():
    import subprocess

    out = subprocess.run(["cuobjdump", "-lelf", capi.lib_path()], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


@pytest.mark.skipif(HAS_GPU, reason="checks the no-device behaviour")
def test_no_cpu_fallback_without_device():
    L = capi.lib()
    assert L.vkm_device_count() == 0
    v = np.zeros(3, capi.VERTEX_DTYPE)
    idx = np.array([0, 1, 2], np.uint32)
    with pytest.raises(capi.VkmError) as e:
        capi.generate_tangents(v, idx)
    assert e.value.code == 5
    with pytest.raises(capi.VkmError) as e:
        capi.import_shapes(np.zeros((3, 3), np.float32), np.zeros((1, 3), np.float32), np.zeros((1, 2), np.float32),
                           np.zeros((3, 3), np.int32), np.array([0, 3], np.uint64))
    assert e.value.code == 5
    with pytest.raises(capi.VkmError) as e:
        capi.Plan(0, 3)
    assert e.value.code == 5


def test_obj_number_parser_has_no_fused_multiply_add():
    """The decimal routine must round every product and sum separately, as the CPU parser does (built without
    contraction): the SASS of the record-parsing kernel may not contain DFMA."""
    import shutil
    import subprocess

    from vulkanengine_b200 import build

    if not shutil.which("cuobjdump"):
        import pytest

        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", build.LIB_PATH], capture_output=True, text=True).stdout
    fn, hits, seen = None, 0, False
    for line in sass.splitlines():
        if "Function :" in line:
            fn = line.split("Function :")[1].strip()
            seen |= "k_parse_records" in fn
        elif fn and "k_parse_records" in fn and "DFMA" in line:
            hits += 1
    assert seen, "k_parse_records not found in the library's SASS"
    assert hits == 0, f"{hits} DFMA instructions in k_parse_records"
