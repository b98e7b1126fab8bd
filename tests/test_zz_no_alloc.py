"""GPU: rt_forest_trees_with_static -- the entry point that allocates nothing (static device
scratch, rust_tracer_b200/csrc/no_alloc.cu) -- against rt_forest_trees_with through the C ABI.
Counts are integers: they must be equal.

Collected last on purpose; passes alone in a fresh process and
after the other tests.
"""
import ctypes as C

import pytest

import rust_tracer_b200 as rt
from rust_tracer_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_TREES_MAX = 4 * 1024 * 1024  # include/rt_b200.h: RT_STATIC_TREES_MAX


def _both(forest, shape_id):
    a, b = C.c_uint64(0xDEAD), C.c_uint64(0xBEEF)
    assert _lib.lib().rt_forest_trees_with(forest._h, shape_id, C.byref(a)) == 0
    assert _lib.lib().rt_forest_trees_with_static(forest._h, shape_id, C.byref(b)) == 0
    return a.value, b.value


def test_static_scratch_entry_point_equals_the_allocating_one_on_calls_in_a_row():
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    forest = sc.build_forest(320, 200, 6)
    seen = []
    # ids 0..11 alias cube triangles with top-level shapes; 12 and -1 are in no tree.  Each id twice in a
    # row and a many-trees id right before an empty one: what a call leaves in the scratch must not reach the next.
    for shape_id in [0, 0, 1, 12, 1, 1, 2, -1, 3, 4, 5, 11, 0]:
        want, got = _both(forest, shape_id)
        assert got == want, shape_id
        seen.append(got)
    assert max(seen) > 0 and min(seen) == 0
    # a second, smaller forest on the same device shares the array
    small = sc.build_forest(64, 48, 3)
    for shape_id in [0, 1, 0]:
        want, got = _both(small, shape_id)
        assert got == want and got <= 64 * 48
    want, got = _both(forest, 0)
    assert got == want == seen[0]
    small.close()
    forest.close()
    sc.close()


def test_static_scratch_entry_point_rejects_a_forest_beyond_its_maximum():
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    w, h = 2048, 2049  # one row of trees more than RT_STATIC_TREES_MAX
    assert w * h > STATIC_TREES_MAX
    forest = sc.build_forest(w, h, 1)
    n = C.c_uint64(7)
    rc = _lib.lib().rt_forest_trees_with_static(forest._h, 0, C.byref(n))
    assert rc == -1 and n.value == 7  # RT_ERR_INVALID, output untouched
    assert b"RT_STATIC_TREES_MAX" in _lib.lib().rt_last_error()
    assert forest.trees_with(0) > 0  # the allocating entry point still serves it
    forest.close()
    # exactly the maximum is accepted
    forest = sc.build_forest(2048, 2048, 1)
    want, got = _both(forest, 0)
    assert got == want > 0
    forest.close()
    sc.close()
