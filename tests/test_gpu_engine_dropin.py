"""GPU acceptance test of the drop-in boundary: the ENGINE'S OWN MeshLoader::Load (unmodified reference object
code in oracle/_ref/libvkref.so) runs with integration/engine_shims.cpp linked in front of it, so its LoadOBJ is the
CUDA-backed shim; the engine's own WriteToE3DM then writes the cache file.  That file must be byte-identical to
the one the pure reference writes for the same OBJ text."""
import os
import subprocess

import numpy as np
import pytest
from conftest import ROOT

from oracle import refcpu
from vulkanengine_b200 import synth

pytestmark = pytest.mark.gpu

DEMO = os.path.join(ROOT, "oracle", "_ref", "dropin_demo")


@pytest.mark.skipif(not (os.path.exists(DEMO) and refcpu.available()), reason="oracle/_ref drop-in demo not built")
@pytest.mark.parametrize("mesh", [lambda: synth.uv_sphere(24, 40), lambda: synth.concat([synth.grid(30), synth.grid(7)])],
                         ids=["sphere", "two_shapes"])
def test_engine_load_through_cuda_shim_writes_the_reference_e3dm(tmp_path, mesh):
    m = mesh()
    d_ours, d_ref = tmp_path / "ours", tmp_path / "ref"
    d_ours.mkdir()
    d_ref.mkdir()
    text = synth.to_obj_text(m)
    (d_ours / "mesh.obj").write_text(text)
    (d_ref / "mesh.obj").write_text(text)
    r = subprocess.run([DEMO, str(d_ours), "mesh"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "vertices" in r.stdout
    refcpu.obj_to_e3dm(str(d_ref / "mesh.obj"), str(d_ref / "mesh.e3dm"), "mesh")
    ours = (d_ours / "mesh.e3dm").read_bytes()
    ref = (d_ref / "mesh.e3dm").read_bytes()
    assert len(ours) == len(ref)
    # Byte-identical except for NaN payloads (degenerate pole triangles: x86 writes 0xFFC00000, the GPU 0x7FFFFFFF):
    # compare the parsed files field by field with NaN == NaN.
    from conftest import vertices_equal

    from vulkanengine_b200 import capi

    name_a, shapes_a, names_a, mats_a = capi.e3dm_read(str(d_ours / "mesh.e3dm"))
    name_b, shapes_b, names_b, mats_b = capi.e3dm_read(str(d_ref / "mesh.e3dm"))
    assert (name_a, names_a, mats_a) == (name_b, names_b, mats_b) and len(shapes_a) == len(shapes_b) > 0
    for x, y in zip(shapes_a, shapes_b):
        assert np.array_equal(x.indices, y.indices)
        ok, msg = vertices_equal(x.vertices, y.vertices)
        assert ok, msg
    if not np.isnan(np.concatenate([x.vertices["tangent"].ravel() for x in shapes_b])).any():
        assert ours == ref
