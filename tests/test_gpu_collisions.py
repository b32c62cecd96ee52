"""Keys engineered to collide in the build's OWN table hashes (generator:
vulkanengine_b200/hash_collide.py).  Collisions may cost re-runs -- the plan statistics must show that the overflow /
re-seed / false-alarm machinery really ran -- but never change a result: every case is compared with the reference.
"""
import os

import numpy as np
import pytest
from conftest import assert_shapes_equal

from oracle import port, refcpu
from vulkanengine_b200 import capi, hash_collide, synth

pytestmark = pytest.mark.gpu


def _ref(m):
    fn = refcpu.import_arrays if refcpu.available() else port.import_arrays
    return fn(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)


def _fresh_import(m, records: bool):
    """Import through a NEW host plan (the plan reads VKM_RECORDS when it is created)."""
    capi.shutdown()
    old = os.environ.get("VKM_RECORDS")
    os.environ["VKM_RECORDS"] = "1" if records else "0"
    try:
        got = capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
        return got, capi.last_import_stats()
    finally:
        if old is None:
            os.environ.pop("VKM_RECORDS", None)
        else:
            os.environ["VKM_RECORDS"] = old
        capi.shutdown()


def test_restated_hashes_are_the_device_hashes():
    """The generator's numpy hashes must be the kernels': a value cluster built with them has to overflow the
    fingerprint table of the record schedule (statistics say so); if the restatement drifted, nothing would."""
    m = hash_collide.inject(synth.grid(60), value_keys=hash_collide.K_MAX_PROBE + 200)
    got, st = _fresh_import(m, records=True)
    assert st["table_overflows"] >= 1 and st["reseeds"] >= 1, st
    assert_shapes_equal(got, _ref(m))


@pytest.mark.parametrize("records", [True, False], ids=["record_schedule", "general_schedule"])
def test_value_cluster_beyond_the_probe_limit_reseeds_and_matches_reference(records):
    """> kMaxProbe distinct value keys in a handful of neighbouring slots at ANY table size: growing the table cannot
    help, only the next hash generation does."""
    m = hash_collide.inject(synth.grid(708), value_keys=hash_collide.K_MAX_PROBE + 100)  # 1.0 M triangles
    got, st = _fresh_import(m, records)
    assert st["table_overflows"] >= 1 and st["reseeds"] >= 1 and st["hash_seed"] >= 1, st
    assert_shapes_equal(got, _ref(m))


@pytest.mark.parametrize("records", [True, False], ids=["record_schedule", "general_schedule"])
def test_triplet_cluster_overflows_the_side_table_and_matches_reference(records):
    """> kMaxProbe distinct triplets that share one position record and cluster in hash_triplet: side table of the
    record schedule, scrambled mode of the general one."""
    m = hash_collide.inject(synth.grid(300), triplet_keys=hash_collide.K_MAX_PROBE + 100)
    got, st = _fresh_import(m, records)
    assert st["table_overflows"] >= 1, st
    assert_shapes_equal(got, _ref(m))


def test_fingerprint_twins_cost_a_rerun_not_a_merge():
    """Two different value keys with one fingerprint in one bucket: the record schedule must call the import
    'maybe not clean' and let the general schedule decide -- which finds no duplicate."""
    base = synth.grid(40)
    m = hash_collide.inject(base, twin_pairs=3)
    got, st = _fresh_import(m, records=True)
    assert st["mis_speculated"] == 1 and st["schedule"] == 2, st
    want = _ref(m)
    assert len(want[0].vertices) == len(base.pos) + 6  # six new vertices: nothing merged
    assert_shapes_equal(got, want)
    # the false alarm does not stick: the next clean mesh speculates again and stands
    capi.shutdown()
    g2 = capi.import_shapes(base.pos, base.nrm, base.uv, base.idx, base.shape_offsets)
    st2 = capi.last_import_stats()
    assert st2["schedule"] == 0 and st2["runs"] == 1, st2
    assert_shapes_equal(g2, _ref(base))


def test_adversarial_with_all_collision_kinds_matches_reference():
    m = synth.adversarial(220, frac=2e-3, nan_frac=1e-3)
    m = hash_collide.inject(m, value_keys=600, triplet_keys=600, twin_pairs=2)
    got, st = _fresh_import(m, records=True)
    assert_shapes_equal(got, _ref(m))


def test_plan_statistics_and_pointer_alignment_checks():
    """vkm_plan_get_stats tells which schedule stood; vkm_plan_import refuses device pointers its vector loads cannot
    take (a sub-range of a packed arena) instead of faulting."""
    m = synth.grid(64)
    d = [capi.DeviceArray(0, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
    na = max(len(m.pos), len(m.nrm), len(m.uv))
    plan = capi.Plan(0, m.corners, max_attrib=na + 8, max_vertices=na + 1024)
    plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
    plan.finish()
    st = plan.stats()
    assert st == {"runs": 1, "schedule": 0, "mis_speculated": 0, "table_overflows": 0, "reseeds": 0, "hash_seed": 0}, st
    dirty = synth.with_duplicate_record(m)
    dp, di = capi.DeviceArray(0, dirty.pos), capi.DeviceArray(0, dirty.idx)
    plan.import_async(dp.ptr, len(dirty.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), di.ptr, dirty.shape_offsets)
    r = plan.finish()
    st = plan.stats()
    assert st["runs"] == 2 and st["schedule"] == 2 and st["mis_speculated"] == 1, st
    assert int(r.total_vertices) == len(m.pos)  # the duplicated record merged
    # misaligned index stream (4 bytes into the buffer) and texcoords (4 bytes in): rejected, nothing launched
    with pytest.raises(capi.VkmError) as e:
        plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr + 4, m.shape_offsets)
    assert e.value.code == 1 and "aligned" in str(e.value)
    with pytest.raises(capi.VkmError) as e:
        plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr + 4, len(m.uv) - 1, d[3].ptr, m.shape_offsets)
    assert e.value.code == 1
    plan.destroy()
    for x in d + [dp, di]:
        x.free()
