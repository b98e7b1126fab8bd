"""GPU tests of the sweep kernel's own machinery (rust_tracer_b200/csrc/sweep.cu): the
conservative bound filter must never change the result of Scene::intersect's linear scan
(reference src/scene/mod.rs:98-116), whatever path a (ray, shape) pair takes through it --
several shared-memory chunks of the bound table, a full candidate queue, rays or bounds the
filter cannot bound, anisotropic-transform segments.  Every case is checked against the
all-pairs audit (every pair through the exact test) and, where the oracle finishes in
seconds, against the oracle.
"""
import ctypes as C

import numpy as np
import pytest

import orc
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi

pytestmark = pytest.mark.gpu


def _render_pair(desc, w, h, depth, cam=None):
    sc = rt.Scene(desc)
    try:
        a = sc.render(w, h, depth, cam=cam)
        b = sc.render(w, h, depth, cam=cam, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    finally:
        sc.close()
    return a, b


def _assert_same(a, b):
    (f1, u1, s1), (f2, u2, s2) = a, b
    assert np.array_equal(f1, f2, equal_nan=True)
    assert (u1 == u2).all()
    assert s1["rays"] == s2["rays"] and s1["nodes"] == s2["nodes"]


def _sphere(L, tx, ty, tz, sx, sy, sz, material, rz=0.0):
    s = abi.rt_shape()
    s.kind, s.material = abi.RT_SHAPE_SPHERE, material
    m = (C.c_float * 16)()
    L.rt_matrix_compose(*[C.c_float(v) for v in (tx, ty, tz, 0.0, 0.0, rz, sx, sy, sz)], m)
    inv = (C.c_float * 16)()
    assert L.rt_matrix_inverse(m, inv) == 0
    s.transform[:] = list(m)
    s.inv_transform[:] = list(inv)
    return s


def _scene_with_spheres(spheres):
    base = rt.SceneDesc.builtin("default")
    n = len(spheres)
    shapes = (abi.rt_shape * (n + 1))(*spheres, base.desc.shapes[4])  # + the checkerboard floor
    shapes[n].material = 3
    mats = (abi.rt_material * 4)(base.desc.materials[0], base.desc.materials[1], base.desc.materials[2],
                                 base.desc.materials[4])
    lights = (abi.rt_light * 3)(*[base.desc.lights[i] for i in range(3)])
    return rt.SceneDesc.from_arrays(shapes, mats, lights, (0.1, 0.1, 0.1)), base


def test_chunked_bound_table_equals_resident_table():
    """A bound table larger than shared memory is swept in chunks, one launch each, the keys
    accumulating in global memory; forced here with a tiny chunk size."""
    desc = rt.SceneDesc.synthetic(400, 400, 4242, 0.03, 0.1)
    w, h = 192, 108
    sc = rt.Scene(desc)
    f0, u0, s0 = sc.render(w, h, 5)
    with rt.tuned(table_cap=512):
        f1, u1, s1 = sc.render(w, h, 5)
    # every sweep finds the nearest hit: the exact tests are then nearly the same set of pairs (not exactly: what the distance
    # clip culls depends on which hits were found first, and chunks are separate launches that start without a clip)
    with rt.tuned(shadow_anyhit=0):
        f2, u2, s2 = sc.render(w, h, 5)
        with rt.tuned(table_cap=512):
            f3, u3, s3 = sc.render(w, h, 5)
    sc.close()
    assert np.array_equal(f0, f1, equal_nan=True) and (u0 == u1).all()
    assert np.array_equal(f0, f2, equal_nan=True) and np.array_equal(f0, f3, equal_nan=True)
    assert s0["rays"] == s1["rays"] and s2["rays"] == s3["rays"] and s2["nodes"] == s3["nodes"]
    assert abs(s2["exact_tests"] - s3["exact_tests"]) <= 0.25 * s2["exact_tests"]
    assert s0["exact_tests"] <= s2["exact_tests"]  # shadow rays that stop at the first occluder test fewer pairs
    assert s1["kernel_launches"] > s0["kernel_launches"]  # the chunks really were separate launches
    of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), 5)
    assert (u1 == ou).all() and s1["rays"] == oc.rays


def test_full_candidate_queue_is_flushed_not_dropped():
    """700 nested spheres: every ray through the middle passes every bound, far more than a
    warp's queue holds between two scheduled flushes."""
    L = rt.lib()
    spheres = [_sphere(L, 0.0, 0.0, 1.0, 0.3 + 0.002 * i, 0.3 + 0.002 * i, 0.3 + 0.002 * i, i % 3) for i in range(700)]
    desc, _keep = _scene_with_spheres(spheres)
    w, h = 64, 36
    a, b = _render_pair(desc, w, h, 3)
    _assert_same(a, b)
    assert a[2]["exact_tests"] > 100 * a[2]["rays"] // 4  # the filter let the nest through
    of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), 3)
    assert (a[1] == ou).all() and a[2]["rays"] == oc.rays


def test_rays_the_filter_cannot_bound_pass_every_shape():
    """A camera 4e15 away: |O| + cmax is beyond the filter's f32 range, so its rays must be
    handed to the exact test for every shape instead of being filtered."""
    desc = rt.SceneDesc.synthetic(60, 40, 77, 0.05, 0.2)
    w, h = 48, 27
    cam = abi.default_camera(w, h)
    cam.origin[:] = [0.0, 0.0, -4.0e15]
    a, b = _render_pair(desc, w, h, 3, cam=cam)
    _assert_same(a, b)
    of, ou, oc = orc.render(desc.desc, cam, 3)
    assert (a[1] == ou).all() and a[2]["rays"] == oc.rays


def test_bound_beyond_the_filter_range_is_always_a_candidate():
    """One sphere of radius 3e15 around everything: its slot carries W = -FLT_MAX."""
    L = rt.lib()
    spheres = [_sphere(L, 0.0, 0.0, 0.0, 3.0e15, 3.0e15, 3.0e15, 2),
               _sphere(L, -1.0, 0.0, 0.0, 1.0, 0.25, 1.0, 0, rz=75.0), _sphere(L, 1.0, -1.0, 0.0, 1.0, 1.0, 1.0, 1)]
    desc, _keep = _scene_with_spheres(spheres)
    w, h = 64, 36
    a, b = _render_pair(desc, w, h, 4)
    _assert_same(a, b)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), 4)
    assert (a[1] == ou).all() and a[2]["rays"] == oc.rays


def test_anisotropic_and_similarity_spheres_use_separate_segments():
    """Spheres squashed up to 16:1 go to the segment with the larger margin; images still equal
    the audit and the oracle."""
    L = rt.lib()
    rng = np.random.default_rng(5)
    spheres = []
    for i in range(150):
        c = rng.uniform([-3, -1.5, -1], [3, 2.5, 5])
        r = float(rng.uniform(0.05, 0.3))
        if i % 2:
            spheres.append(_sphere(L, *c, r, r / float(rng.uniform(2, 16)), r, i % 3, rz=float(rng.uniform(0, 180))))
        else:
            spheres.append(_sphere(L, *c, r, r, r, i % 3))
    desc, _keep = _scene_with_spheres(spheres)
    w, h = 160, 90
    a, b = _render_pair(desc, w, h, 4)
    _assert_same(a, b)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), 4)
    assert (a[1] == ou).all() and a[2]["rays"] == oc.rays


def test_ray_pools_are_reused_across_scenes_and_trim_frees_them():
    desc = rt.SceneDesc.builtin("default")
    _, u0, _ = rt.Scene(desc).render(96, 54, 5, want_f32=False)
    for _ in range(3):  # create -> render -> destroy, pools handed from scene to scene
        sc = rt.Scene(desc)
        _, u, _ = sc.render(96, 54, 5, want_f32=False)
        sc.close()
        assert (u == u0).all()
    rt.trim()
    _, u, _ = rt.Scene(desc).render(96, 54, 5, want_f32=False)
    assert (u == u0).all()


def _moved_camera(w, h, origin):
    cam = abi.default_camera(w, h)
    cam.origin[:] = list(origin)
    return cam


@pytest.mark.parametrize("origin", [(0.0, 0.0, -8.0), (0.5, 0.7, 1.5), (300.0, -120.0, -4000.0)])
def test_origin_tables_equal_general_filter_and_audit(origin):
    """Primary rays are swept against the bound table re-centred on the camera, the shadow rays
    of each light against the table re-centred on that light (4 FMA per pair instead of 7);
    the nearest hits must be those of the general filter and of the all-pairs audit -- also
    for a camera inside the cloud of shapes and one far outside it."""
    desc = rt.SceneDesc.synthetic(500, 500, 777, 0.03, 0.1)
    w, h, depth = 224, 126, 4
    cam = _moved_camera(w, h, origin)
    sc = rt.Scene(desc)
    try:
        a = sc.render(w, h, depth, cam=cam)
        b = sc.render(w, h, depth, cam=cam, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
        a4 = sc.render(w, h, depth, cam=cam, spp=4, seed=11)
        b4 = sc.render(w, h, depth, cam=cam, spp=4, seed=11, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    finally:
        sc.close()
    with rt.tuned(origin_tables=0):
        sc = rt.Scene(desc)
        try:
            g = sc.render(w, h, depth, cam=cam)
        finally:
            sc.close()
    _assert_same(a, b)
    _assert_same(a4, b4)
    _assert_same(a, g)
    # the origin tables really were used: fewer filter flops for the same pairs
    assert a[2]["filter_flops"] < 0.8 * g[2]["filter_flops"]
    assert a[2]["kernel_launches"] > g[2]["kernel_launches"]


def test_origin_tables_follow_the_camera():
    """The camera table is rebuilt when the origin changes between calls on one scene."""
    desc = rt.SceneDesc.synthetic(300, 300, 99, 0.04, 0.12)
    w, h = 160, 90
    sc = rt.Scene(desc)
    try:
        for origin in ((0.0, 0.0, -8.0), (2.0, 1.0, -3.0), (0.0, 0.0, -8.0)):
            cam = _moved_camera(w, h, origin)
            _assert_same(sc.render(w, h, 3, cam=cam), sc.render(w, h, 3, cam=cam, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS))
    finally:
        sc.close()


def test_origin_tables_with_a_light_inside_a_shape_and_far_lights():
    """A light at the centre of a sphere (W' < 0: every shadow ray of that light is a candidate
    of that sphere), one on a surface, one 1e6 away, one beyond the filter's range."""
    base = rt.SceneDesc.synthetic(300, 300, 5, 0.04, 0.12)
    n = base.desc.n_shapes
    shapes = (abi.rt_shape * n)(*[base.desc.shapes[i] for i in range(n)])
    mats = (abi.rt_material * base.desc.n_materials)(*[base.desc.materials[i] for i in range(base.desc.n_materials)])
    s1 = shapes[1]  # a sphere: centre = translation column of its transform
    centre = (s1.transform[3], s1.transform[7], s1.transform[11])
    lights = (abi.rt_light * 4)(*[base.desc.lights[i] for i in range(4)])
    lights[0].pos[:] = list(centre)
    lights[1].pos[:] = [centre[0] + abs(s1.transform[0]), centre[1], centre[2]]
    lights[2].pos[:] = [1.0e6, 2.0e6, -1.5e6]
    lights[3].pos[:] = [0.0, 4.0e15, 0.0]
    desc = rt.SceneDesc.from_arrays(shapes, mats, lights, (0.1, 0.1, 0.1))
    a, b = _render_pair(desc, 160, 90, 3)
    _assert_same(a, b)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(160, 90), 3)
    assert (a[1] == ou).all() and a[2]["rays"] == oc.rays


def test_work_item_shapes_do_not_change_the_image():
    """The sweep's scheduling knobs (rt_debug_set) -- table slices per work item (want_light, slices),
    the threshold between sweeping a group ray-by-ray or pair-by-pair (item_max) -- and
    the node order must not change a single key: same image, same ray and node counts, and
    the same as the all-pairs audit (which uses none of them)."""
    desc = rt.SceneDesc.synthetic(600, 600, 31, 0.03, 0.1)  # floor plane: hit points far outside the shapes' box
    w, h, depth = 256, 144, 4
    ref = None
    for knobs in ({}, {"want_light": 1}, {"want_light": 64}, {"slices": 1}, {"slices": 5}, {"item_max": 0},
                  {"item_max": 32}, {"rays_per_thread": 1}, {"rays_per_thread": 2}, {"shadow_anyhit": 0},
                  {"shadow_anyhit": 0, "slices": 3}, {"node_bits": 3}, {"node_bits": 5}, {"node_bits": 6}, {"node_bits": 7}):
        with rt.tuned(**knobs):
            sc = rt.Scene(desc)
            try:
                out = sc.render(w, h, depth)
                if ref is None:
                    ref = out
                    audit = sc.render(w, h, depth, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
                    _assert_same(ref, audit)
            finally:
                sc.close()
        _assert_same(ref, out)


def test_far_hit_points_take_the_heavy_path():
    """A camera low over the floor: most primary rays hit the plane far outside the shapes' box,
    their shadow rays carry wide filter margins (many candidates) and are scheduled first in
    finer work items.  Result = audit = oracle."""
    desc = rt.SceneDesc.synthetic(300, 300, 8, 0.04, 0.12)
    w, h = 192, 108
    cam = abi.default_camera(w, h)
    cam.origin[:] = [0.0, -1.9, -8.0]
    cam.y_min, cam.y_max = -2.2, -1.2
    a, b = _render_pair(desc, w, h, 3, cam=cam)
    _assert_same(a, b)
    of, ou, oc = orc.render(desc.desc, cam, 3)
    assert (a[1] == ou).all() and a[2]["rays"] == oc.rays and a[2]["nodes"] == oc.nodes


def test_bands_do_not_change_the_image():
    """A frame larger than one band of samples is cast band by band (each with its own ray pools,
    node order and counters); cutting a small frame into many bands must give the same image,
    ray and node counts as one band, for 1 and 4 samples per pixel."""
    desc = rt.SceneDesc.synthetic(300, 300, 77, 0.04, 0.12)
    w, h, depth = 200, 120, 4
    sc = rt.Scene(desc)
    try:
        one = sc.render(w, h, depth)
        one4 = sc.render(w, h, depth, spp=4, seed=5)
        with rt.tuned(band_samples=3000):
            many = sc.render(w, h, depth)
            many4 = sc.render(w, h, depth, spp=4, seed=5)
    finally:
        sc.close()
    _assert_same(one, many)
    _assert_same(one4, many4)
    assert many[2]["kernel_launches"] > 4 * one[2]["kernel_launches"]


def test_tested_pairs_counter():
    """rt_stats.tested_pairs (credit only the tests actually resolved): every ray x shape
    pair under the all-pairs audit, a small share of them under the group / quad culling."""
    desc = rt.SceneDesc.synthetic(600, 600, 3, 0.03, 0.1)
    a, b = _render_pair(desc, 160, 90, 3)
    n_shapes = desc.desc.n_shapes
    assert b[2]["tested_pairs"] == b[2]["rays"] * n_shapes
    assert 0 < a[2]["tested_pairs"] < 0.5 * a[2]["rays"] * n_shapes
    assert a[2]["tested_pairs"] >= a[2]["plane_tests"]


def test_merged_light_sweeps_and_tail_items_do_not_change_the_image():
    """One launch for every light's shadow rays (launch_sweep_jobs: CTA b serves light b % L) against
    one launch per light (merge_lights = 0), and the finer work items at the end of a launch
    (tail_pct, tail_mult): scheduling only -- the same image and counts,
    and the same as the all-pairs audit.  Also at fewer rays per thread, where the jobs kernel is
    another instantiation; with and without CTAs moving on to another light's items when theirs run out
    (job_rotate); levels pipelined or joined one by one (level_pipeline); primary rays in tiles or row by row
    (primary_tiles); reflected / refracted rays at 1, 2 or 4 per thread."""
    desc = rt.SceneDesc.synthetic(500, 500, 77, 0.03, 0.1)
    w, h, depth = 224, 126, 4
    ref = None
    for knobs in ({}, {"merge_lights": 0}, {"tail_pct": 0}, {"tail_pct": 100, "tail_mult": 8},
                  {"tail_pct": 50, "tail_mult": 64}, {"rays_per_thread": 1},
                  {"rays_per_thread": 2, "want_light": 16}, {"shadow_anyhit": 0}, {"shadow_anyhit": 0, "merge_lights": 0},
                  {"merge_lights": 0, "rays_per_thread": 2}, {"merge_lights": 0, "table_cap": 512},
                  {"job_rotate": 0}, {"job_rotate": 0, "rays_per_thread": 1}, {"job_rotate": 1, "want_heavy_jobs": 64},
                  {"level_pipeline": 0}, {"level_pipeline": 0, "merge_lights": 0}, {"level_pipeline": 1, "merge_lights": 0},
                  {"primary_tiles": 0}, {"primary_tiles": 1}, {"tree_rays_per_thread": 4}, {"tree_rays_per_thread": 1, "want_light": 16},
                  {"small_shadow_rays": 1 << 30}, {"small_tree_rays": 0}):
        with rt.tuned(**knobs):
            sc = rt.Scene(desc)
            try:
                out = sc.render(w, h, depth)
                if ref is None:
                    ref = out
                    _assert_same(ref, sc.render(w, h, depth, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS))
            finally:
                sc.close()
        _assert_same(ref, out)
    assert ref[2]["kernel_launches"] > 0


@pytest.mark.parametrize("n_lights", [1, 2, 5, 8, 9])
def test_any_number_of_point_lights(n_lights):
    """1 light: nothing to merge; 2..8: one jobs launch per level (8 = MAX_SWEEP_JOBS); 9: more lights
    than a jobs launch takes, back to one launch per light.  Against the oracle and the audit."""
    base = rt.SceneDesc.synthetic(150, 150, 5, 0.05, 0.14)
    d = base.desc
    lights = (abi.rt_light * n_lights)()
    rng = np.random.default_rng(n_lights)
    for i in range(n_lights):
        lights[i].kind = abi.RT_LIGHT_POINT
        lights[i].pos[:] = [float(rng.uniform(-5, 5)), float(rng.uniform(2, 8)), float(rng.uniform(-7, 2))]
        lights[i].color[:] = [float(x) for x in rng.uniform(0.1, 0.5, 3)]
    shapes = (abi.rt_shape * d.n_shapes)(*[d.shapes[i] for i in range(d.n_shapes)])
    mats = (abi.rt_material * d.n_materials)(*[d.materials[i] for i in range(d.n_materials)])
    desc = rt.SceneDesc.from_arrays(shapes, mats, lights, (0.1, 0.1, 0.1))
    w, h, depth = 128, 72, 4
    a, b = _render_pair(desc, w, h, depth)
    _assert_same(a, b)
    _, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), depth)
    diff = np.abs(a[1].astype(np.int32) - ou.astype(np.int32)).max(axis=-1)
    assert float((diff <= 1).mean()) >= 0.999, int(diff.max())
    assert a[2]["rays"] == oc.rays and a[2]["shadow_rays"] == oc.shadow_rays


def test_scene_tables_kept_for_an_identical_description():
    """rt_scene_create keeps the host-side tables of the last description it built and hands them to the
    next scene if its description is the same byte for byte (csrc/api_scene.cu: host_scene_for); a
    description that differs in one float gets its own build.  Same images either way."""
    desc = rt.SceneDesc.synthetic(200, 200, 17, 0.04, 0.12)
    w, h, depth = 160, 90, 4
    with rt.tuned(scene_cache=0):
        ref = _render_once(desc, w, h, depth)
    a = _render_once(desc, w, h, depth)  # builds and keeps
    b = _render_once(desc, w, h, depth)  # reuses
    _assert_same(ref, a)
    _assert_same(ref, b)
    d = desc.desc
    shapes = (abi.rt_shape * d.n_shapes)(*[d.shapes[i] for i in range(d.n_shapes)])
    mats = (abi.rt_material * d.n_materials)(*[d.materials[i] for i in range(d.n_materials)])
    lights = (abi.rt_light * d.n_lights)(*[d.lights[i] for i in range(d.n_lights)])
    lights[0].color[0] += 0.25
    other = rt.SceneDesc.from_arrays(shapes, mats, lights, list(d.ambient))
    c = _render_once(other, w, h, depth)
    assert not np.array_equal(c[0], a[0], equal_nan=True)  # not served from the kept tables
    with rt.tuned(scene_cache=0):
        _assert_same(c, _render_once(other, w, h, depth))
    _assert_same(a, _render_once(desc, w, h, depth))  # and back
    rt.trim(-1)
    _assert_same(a, _render_once(desc, w, h, depth))


def _render_once(desc, w, h, depth):
    sc = rt.Scene(desc)
    try:
        return sc.render(w, h, depth)
    finally:
        sc.close()
