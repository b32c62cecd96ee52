"""GPU tests of the OBJ text parser: vkm_parse_obj against the reference's own tinyobj
(oracle/_ref/libvkref.so, ref_parse_obj) on the same text -- floats bit for bit, index streams, shape split,
names.  The expected values come from running the reference parser on the GPU box itself."""
import numpy as np
import pytest

from oracle import refcpu
from vulkanengine_b200 import capi, synth

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refcpu.available(), reason="oracle/_ref/libvkref.so not built")]


def compare(text: str, tmp_path, name="m", mtls=None):
    p = tmp_path / (name + ".obj")
    p.write_bytes(text.encode())
    for fn, body in (mtls or {}).items():
        (tmp_path / fn).write_bytes(body)
    want = refcpu.parse_obj(str(p))
    got = capi.parse_obj(text.encode(), mtls=mtls or {})
    for a, b, what in [(got.pos, want.pos, "v"), (got.nrm, want.nrm, "vn"), (got.uv, want.uv, "vt")]:
        assert a.shape == b.shape, what
        ua, ub = a.view(np.uint32), b.view(np.uint32)
        same = (ua == ub) | (np.isnan(a) & np.isnan(b))
        assert same.all(), f"{what}: first difference at {np.argwhere(~same)[0]}: {a[~same][0]!r} vs {b[~same][0]!r}"
    assert np.array_equal(got.idx, want.idx)
    assert np.array_equal(got.shape_offsets, want.shape_offsets)
    assert got.shape_names == want.shape_names
    assert got.material_names == want.material_names  # materials[material_ids[0]].name, "" for id -1
    return got, want


def test_synthetic_meshes_round_trip(tmp_path):
    for i, m in enumerate([synth.uv_sphere(12, 20), synth.concat([synth.grid(9), synth.grid(4)]), synth.adversarial(12)]):
        got, want = compare(synth.to_obj_text(m), tmp_path, f"s{i}")
        assert np.array_equal(got.idx, m.idx)


NUMBERS = ["1", "-1.", ".5", "+.25e1", "1e-3", "1E+2", "123456789.123456789", "0.1234567890123456789", "1e400", "1e-400",
           "-0", "-0.0", "abc", "1.5x", "3.14159265358979", "2.718281828e0", "-7.25E-2", "1e", "1e+", "000.125",
           "4.9406564584124654e-324", "1.7976931348623157e308", "0.30000000000000004", "16777217", "1e22", "1e23",
           "9007199254740993", "0.000001", "0.0000001", "0.00000001", "123.456e-7", "+5", "-.75", "5.", "1.0000001"]


def test_number_formats_match_tinyobj_bit_for_bit(tmp_path):
    rng = np.random.default_rng(3)
    lines = []
    for i in range(0, len(NUMBERS) - 2, 1):
        lines.append(f"v {NUMBERS[i]} {NUMBERS[i + 1]} {NUMBERS[i + 2]}")
    for _ in range(400):  # random decimal strings with 1..25 significant digits and odd exponents
        parts = []
        for _ in range(3):
            nd = int(rng.integers(1, 26))
            digs = "".join(str(int(d)) for d in rng.integers(0, 10, nd))
            dot = int(rng.integers(0, nd + 1))
            s = digs[:dot] + "." + digs[dot:] if rng.random() < 0.8 else digs
            if rng.random() < 0.3:
                s += ("e" if rng.random() < 0.5 else "E") + str(int(rng.integers(-45, 45)))
            parts.append(("-" if rng.random() < 0.3 else "") + s)
        lines.append("vn " + " ".join(parts))
        lines.append("vt " + " ".join(parts[:2]))
    lines.append("v 1 2")            # missing z -> 0
    lines.append("v 1 2 3 0.5 0.25 1")  # vertex colours are ignored here
    lines.append("vt 0.5 0.5 0.0")     # w ignored
    lines += ["f 1 2 3"]
    compare("\n".join(lines) + "\n", tmp_path)


TRICKY = """# comment line
mtllib tricky.mtl
\tv 0 0 0
v 1 0 0\r
v 1 1 0
v 0 1 0
v 0.5 0.5 1
v 2 0 0
v 2 3 0
vt 0 0
vt 1 0
vt 1 1
vt 0 1
vn 0 0 1
vn 0 1 0

f 1 2 3
f 1/1 2/2 3/3
f 1//1 2//1 3//2
f 1/1/1 2/2/1 3/3/2
o first object  with spaces
usemtl matA
f -5 -4 -3
f -5/-4/-2 -4/-3/-2 -3/-2/-1
f 1 2 3 4
f 2 6 7 3
f 1/1/1 2/2/1 3/3/1 4/4/1
g grp1   grp2\tgrp3
f 1 2 5
usemtl matB
f 2 3 5
g
g empty_then_named
o
usemtl matA
o second
f 3 4 5
s off
f 4 1 5
"""


def test_tricky_text_matches_tinyobj(tmp_path):
    mtl = {"tricky.mtl": b"newmtl matA\nKd 1 0 0\nnewmtl matB\nKd 0 1 0\n"}
    got, want = compare(TRICKY, tmp_path, "tricky", mtl)
    assert len(got.shape_names) >= 4
    assert got.usemtl_names == want.material_names  # here every usemtl names a material the .mtl defines


def test_material_resolution_follows_tinyobj(tmp_path):
    """usemtl is looked up when it is READ (tinyobj.hpp:2812-2826): undefined names, a usemtl ahead of its mtllib, a
    .mtl that cannot be read, several file names on one mtllib line, CRLF and blank-led .mtl lines."""
    tri = "v 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\n"
    cases = {
        "undefined": ("mtllib a.mtl\n" + tri + "o s1\nusemtl nope\nf 1 2 3\no s2\nusemtl red\nf 2 3 4\n", {"a.mtl": b"newmtl red\n"}),
        "early_usemtl": (tri + "usemtl red\no s1\nf 1 2 3\nmtllib a.mtl\no s2\nf 2 3 4\no s3\nusemtl red\nf 1 3 4\n", {"a.mtl": b"newmtl red\n"}),
        "missing_file": ("mtllib gone.mtl\n" + tri + "usemtl red\nf 1 2 3\n", {}),
        "two_files": ("mtllib gone.mtl b.mtl c.mtl\n" + tri + "o s1\nusemtl blue\nf 1 2 3\no s2\nusemtl green\nf 2 3 4\n",
                      {"b.mtl": b"  newmtl blue\r\nKd 0 0 1\r\n", "c.mtl": b"newmtl green\n"}),
        "no_usemtl": ("mtllib a.mtl\n" + tri + "f 1 2 3\n", {"a.mtl": b"newmtl red\n"}),
        "mid_shape_change": ("mtllib a.mtl\n" + tri + "o s1\nf 1 2 3\nusemtl red\nf 2 3 4\n", {"a.mtl": b"newmtl red\nnewmtl other\n"}),
    }
    for name, (text, mtl) in cases.items():
        compare(text, tmp_path, name, mtl)


def _polygon_file(rings, rng, planar=True):
    """OBJ text with one face per ring (lists of 2-D points lifted into a tilted plane), plus vt / vn so that all three
    index columns are exercised."""
    lines, faces, base = [], [], 0
    for ring in rings:
        for (x, y) in ring:
            z = 0.3 * x - 0.2 * y + (0.0 if planar else 0.05 * float(rng.standard_normal()))
            lines.append("v %.9g %.9g %.9g" % (x, y, z))
            lines.append("vt %.9g %.9g" % (x * 0.1, y * 0.1))
        lines.append("vn 0 0 1")
        faces.append("f " + " ".join("%d/%d/%d" % (base + k + 1, base + k + 1, len(faces) + 1) for k in range(len(ring))))
        base += len(ring)
    return "\n".join(lines + faces) + "\n"


def test_polygons_with_more_than_four_corners_follow_tinyobj_ear_clipping(tmp_path):
    rng = np.random.default_rng(11)
    rings = []
    for n in (5, 6, 7, 8, 12, 31):  # convex: regular n-gons, both windings
        ang = np.linspace(0, 2 * np.pi, n, endpoint=False)
        ring = [(float(3 * np.cos(a)), float(3 * np.sin(a))) for a in ang]
        rings += [ring, ring[::-1]]
    rings.append([(0, 0), (4, 0), (4, 4), (2, 1), (0, 4)])                      # concave: notch
    rings.append([(0, 0), (6, 0), (6, 2), (2, 2), (2, 4), (6, 4), (6, 6), (0, 6)])  # concave: C shape
    rings.append([(0, 0), (2, 0), (4, 0), (4, 3), (2, 3), (0, 3)])              # collinear runs
    rings.append([(0, 0), (1, 0), (1, 0), (1, 1), (0, 1)])                      # repeated point
    rings.append([(0, 0), (1, 1), (2, 2), (3, 3), (4, 4)])                      # all collinear (zero area)
    for _ in range(40):  # star-shaped random polygons (simple, often concave)
        n = int(rng.integers(5, 20))
        ang = np.sort(rng.uniform(0, 2 * np.pi, n))
        rad = rng.uniform(0.5, 3.0, n)
        ring = [(float(r * np.cos(a)), float(r * np.sin(a))) for a, r in zip(ang, rad)]
        rings.append(ring if rng.random() < 0.5 else ring[::-1])
    for _ in range(20):  # arbitrary point orders: self-intersecting, the bail-out counters decide
        n = int(rng.integers(5, 12))
        rings.append([(float(x), float(y)) for x, y in rng.uniform(-2, 2, (n, 2))])
    got, want = compare(_polygon_file(rings, rng), tmp_path, "poly")
    assert len(want.idx) > 3 * len(rings)
    compare(_polygon_file(rings, rng, planar=False), tmp_path, "poly_bent")  # non-planar: other projection axes


def test_mixed_triangles_quads_polygons_and_shapes(tmp_path):
    text = ("v 0 0 0\nv 1 0 0\nv 2 0.5 0\nv 2 2 0\nv 1 3 0\nv 0 2 0\nv -1 1 0\nv 0.5 1 1\n"
            "vt 0 0\nvt 1 1\nvn 0 0 1\n"
            "o first\nf 1 2 3\nf 1 2 3 4\nf 1 2 3 4 5\nf 1/1/1 2/2/1 3/1/1 4/2/1 5/1/1 6/2/1 7/1/1\n"
            "g second\nf -8 -7 -6 -5 -4 -3\nf 2 3 8\nusemtl m\nf 7//1 6//1 5//1 4//1 3//1 2//1 1//1\n")
    compare(text, tmp_path, "mixed")


def test_asset_text_to_e3dm_is_the_reference_shipped_file(tmp_path):
    """falcon.obj TEXT (+ falcon.mtl) -> GPU parse -> GPU geometry stage -> E3DM writer == the byte-identical file the
    reference ships as assets/models/falcon.e3dm, material name `falc_wreck` included, with no engine-side lookup.
    viking_room.obj names a .mtl that is not shipped: its material name must come out empty."""
    import hashlib

    from conftest import load_golden

    txt = load_golden("asset_text")
    g = load_golden("falcon")
    shapes, names, mats = capi.import_obj_text(txt["falcon_obj"].tobytes(), mtls={"falcon.mtl": txt["falcon_mtl"].tobytes()})
    assert names == [str(s) for s in g["shape_names"]] and mats == ["falc_wreck"]
    p = str(tmp_path / "falcon.e3dm")
    capi.e3dm_write(p, "falcon", shapes, names, mats)
    assert hashlib.sha256(open(p, "rb").read()).hexdigest() == str(g["e3dm_sha256"])
    v = load_golden("viking_room")
    shapes, names, mats = capi.import_obj_text(txt["viking_room_obj"].tobytes(), mtls={})
    assert names == [str(s) for s in v["shape_names"]] and mats == [""] == [str(s) for s in v["material_names"]]
    assert len(shapes[0].vertices) == len(v["vertices0"]) and np.array_equal(shapes[0].indices, v["indices0"])


def test_crlf_and_lone_cr_line_endings(tmp_path):
    body = "v 0 0 0\nv 1 0 0\nv 0 1 0\nvt 0 0\nvn 0 0 1\nf 1/1/1 2/1/1 3/1/1\n"
    for name, text in [("crlf", body.replace("\n", "\r\n")), ("cr", body.replace("\n", "\r")), ("nofinal", body.rstrip("\n"))]:
        compare(text, tmp_path, name)


def test_unsupported_and_bad_input():
    with pytest.raises(capi.VkmError) as e:  # a polygon corner beyond the file's records
        capi.parse_obj(b"v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv 0.5 2 0\nf 1 2 3 4 9\n")
    assert e.value.code == 7
    big = "".join("v %d %d 0\n" % (k, k * k) for k in range(300)) + "f " + " ".join(str(k + 1) for k in range(300)) + "\n"
    with pytest.raises(capi.VkmError) as e:  # more than 255 corners
        capi.parse_obj(big.encode())
    assert e.value.code == 7
    with pytest.raises(capi.VkmError) as e:  # polyline record
        capi.parse_obj(b"v 0 0 0\nv 1 0 0\nl 1 2\n")
    assert e.value.code == 7
    with pytest.raises(capi.VkmError) as e:  # zero vertex index: tinyobj fails the load
        capi.parse_obj(b"v 0 0 0\nv 1 0 0\nv 1 1 0\nf 0 1 2\n")
    assert e.value.code == 2
    empty = capi.parse_obj(b"")
    assert len(empty.pos) == 0 and len(empty.idx) == 0 and empty.shape_names == []


def test_parse_then_import_equals_reference_loadobj(tmp_path):
    """OBJ text -> GPU parse -> GPU import == the reference's LoadOBJ on the same file."""
    from conftest import assert_shapes_equal

    m = synth.concat([synth.uv_sphere(10, 16), synth.grid(12)])
    p = tmp_path / "mesh.obj"
    p.write_text(synth.to_obj_text(m))
    want = refcpu.load_obj(str(p), "mesh")
    s = capi.parse_obj(p.read_bytes())
    got = capi.import_shapes(s.pos, s.nrm, s.uv, s.idx, s.shape_offsets)
    assert_shapes_equal(got, want)
    assert s.shape_names == [w.name for w in want]


def test_import_obj_text_equals_reference_loadobj(tmp_path):
    """The whole of LoadOBJ from text on the GPU (streams never visit the host) == the reference on the same file."""
    from conftest import assert_shapes_equal

    (tmp_path / "tricky.mtl").write_text("newmtl matA\nKd 1 0 0\nnewmtl matB\nKd 0 1 0\n")
    m = synth.concat([synth.uv_sphere(9, 14), synth.grid(11), synth.adversarial(10)])
    for name, text in [("mesh", synth.to_obj_text(m))]:
        p = tmp_path / (name + ".obj")
        p.write_text(text)
        want = refcpu.load_obj(str(p), name)
        got, names, mats = capi.import_obj_text(text.encode())
        assert_shapes_equal(got, want)
        assert names == [w.name for w in want]


def test_cached_parser_state_across_sizes_and_shutdown(tmp_path):
    """The parser keeps its device buffers, stream and staging between calls: a large file (above the staged-upload
    threshold), then small ones, then the large one again, with a vkm_shutdown in between, all match tinyobj."""
    big = synth.to_obj_text(synth.grid(260))          # about 15 MB of text: uploaded through the page-locked staging
    small = synth.to_obj_text(synth.uv_sphere(6, 9))
    tiny = "v 0 0 0\nv 1 0 0\nv 0 1 0\nvt 0 0\nvn 0 0 1\nf 1/1/1 2/1/1 3/1/1\n"
    for i, text in enumerate([big, small, tiny, big]):
        compare(text, tmp_path, f"c{i}")
        if i == 1:
            capi.shutdown()
    got, _, _ = capi.import_obj_text(big.encode())
    m = synth.grid(260)
    from oracle import port
    from conftest import assert_shapes_equal
    assert_shapes_equal(got, port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets))


def test_decimal_fuzz_matches_tinyobj_bit_for_bit(tmp_path):
    """60 000 random decimal literals (many digits, long fractions, exponents): every float equal to tinyobj's.  A fused
    multiply-add anywhere in the accumulation shows up here as a float that is 1 ulp off."""
    rng = np.random.default_rng(2024)
    lines = []
    for _ in range(20000):
        parts = []
        for _ in range(3):
            nd = int(rng.integers(1, 30))
            digs = "".join(str(int(d)) for d in rng.integers(0, 10, nd))
            dot = int(rng.integers(0, nd + 1))
            s = (digs[:dot] + "." + digs[dot:]) if rng.random() < 0.85 else digs
            if rng.random() < 0.35:
                s += ("e" if rng.random() < 0.5 else "E") + ("+" if rng.random() < 0.2 else "") + str(int(rng.integers(-60, 50)))
            parts.append(("-" if rng.random() < 0.3 else "") + s)
        lines.append("v " + " ".join(parts))
    lines.append("f 1 2 3")
    compare("\n".join(lines) + "\n", tmp_path, "fuzz")
