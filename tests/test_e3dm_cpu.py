"""CPU tests of the E3DM cache code: the writer reproduces the reference-SHIPPED file
assets/models/falcon.e3dm byte for byte from the golden buffers, and reader and writer are inverses."""
import hashlib

import numpy as np
from conftest import load_golden

from vulkanengine_b200 import capi


def _golden_shapes(g):
    return [capi.ShapeResult(np.ascontiguousarray(g[f"vertices{i}"]).view(capi.VERTEX_DTYPE).reshape(-1),
                             g[f"indices{i}"], g[f"extent{i}"]) for i in range(int(g["nshapes"]))]


def test_writer_reproduces_reference_shipped_falcon_e3dm(tmp_path):
    g = load_golden("falcon")
    p = str(tmp_path / "falcon.e3dm")
    capi.e3dm_write(p, str(g["e3dm_name"]), _golden_shapes(g), [str(s) for s in g["shape_names"]],
                    [str(s) for s in g["material_names"]])
    raw = open(p, "rb").read()
    assert len(raw) == int(g["e3dm_size"]) == 446725
    assert hashlib.sha256(raw).hexdigest() == str(g["e3dm_sha256"])


def test_reader_inverts_writer(tmp_path):
    g = load_golden("viking_room")
    shapes = _golden_shapes(g) * 2
    p = str(tmp_path / "two.e3dm")
    capi.e3dm_write(p, "viking", shapes, ["a", ""], ["m0", ""])
    name, got, names, mats = capi.e3dm_read(p)
    assert name == "viking" and names == ["a", ""] and mats == ["m0", ""]
    for a, b in zip(got, shapes):
        assert a.vertices.tobytes() == b.vertices.tobytes() and np.array_equal(a.indices, b.indices)


def test_reader_rejects_truncated_file(tmp_path):
    import pytest

    g = load_golden("viking_room")
    p = str(tmp_path / "t.e3dm")
    capi.e3dm_write(p, "v", _golden_shapes(g), ["s"], [""])
    raw = open(p, "rb").read()
    open(p, "wb").write(raw[: len(raw) // 2])
    with pytest.raises(capi.VkmError):
        capi.e3dm_read(p)
