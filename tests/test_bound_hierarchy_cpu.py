"""CPU: the bound hierarchy the library builds from a scene description (api_scene.cu: build_host_scene), read back
through the test hooks of include/rt_b200_debug.h.  No GPU.

What the sweep's filter and its distance clip rest on ("filter margins"): every bound slot lies inside its
group sphere, every group sphere inside its quad sphere, every quad sphere inside its super sphere; a level's largest
radius is what DevScene::level_rmax says; cmax reaches beyond every sphere of every level."""
import ctypes as C

import numpy as np
import pytest

import rust_tracer_b200 as rt

GROUP, QUAD, SUPER, SEG_PAD = 16, 64, 256, 512


def _level(desc, level):
    L = rt.lib()
    n = L.rt_debug_scene_spheres(C.byref(desc.desc), level, None, 0)
    assert n > 0
    buf = np.zeros((n, 5), np.float64)
    assert L.rt_debug_scene_spheres(C.byref(desc.desc), level, buf.ctypes.data_as(C.POINTER(C.c_double)), n) == n
    return buf


def _scalars(desc):
    out = (C.c_double * 6)()
    assert rt.lib().rt_debug_scene_scalars(C.byref(desc.desc), out) == 0
    return dict(zip(["n_slots", "cmax", "slot_rmax", "group_rmax", "quad_rmax", "super_rmax"], out))


def _contained(children, parents, fan):
    """Every child sphere (kind 0) inside the parent sphere of its block of `fan`; returns the largest overshoot / parent radius."""
    worst = 0.0
    for j, par in enumerate(parents):
        kids = children[j * fan:(j + 1) * fan]
        real = kids[kids[:, 4] == 0]
        if par[4] == 1:  # a parent of pads only
            assert len(real) == 0 and not (kids[:, 4] == 2).any()
            continue
        if (kids[:, 4] == 2).any():
            assert par[4] == 2  # a member every ray must pass makes its parent one
            continue
        assert par[4] == 0 and len(real) > 0
        reach = np.linalg.norm(real[:, :3] - par[:3], axis=1) + real[:, 3]
        worst = max(worst, float((reach - par[3]).max() / par[3]))
    return worst


@pytest.mark.parametrize("name", ["default", "s1k", "m10k", "synthetic"])
def test_every_level_contains_the_level_below(name):
    desc = rt.SceneDesc.synthetic(700, 900, 31, 0.02, 0.3) if name == "synthetic" else rt.SceneDesc.builtin(name)
    slots, groups, quads, supers = (_level(desc, k) for k in range(4))
    sc = _scalars(desc)
    n = int(sc["n_slots"])
    assert len(slots) == n and n % SEG_PAD == 0
    assert len(groups) == n // GROUP and len(quads) == n // QUAD and len(supers) == n // SUPER
    n_bounded = sum(1 for i in range(desc.desc.n_shapes) if desc.desc.shapes[i].kind != 1)  # planes have no bound
    assert int((slots[:, 4] != 1).sum()) == n_bounded
    # float32 centres: containment holds up to their rounding, far below the 1e-6 relative slack asked here
    assert _contained(slots, groups, GROUP) <= 1e-6
    assert _contained(groups, quads, QUAD // GROUP) <= 1e-6
    assert _contained(quads, supers, SUPER // QUAD) <= 1e-6
    for level, key in ((slots, "slot_rmax"), (groups, "group_rmax"), (quads, "quad_rmax"), (supers, "super_rmax")):
        real = level[level[:, 4] == 0]
        if (level[:, 4] == 2).any():
            assert sc[key] >= 1e38  # a sphere every ray passes: no clip at this level
        else:
            assert real[:, 3].max() <= sc[key] <= real[:, 3].max() * 1.001 + 1e-4 * sc["cmax"]
        reach = np.linalg.norm(real[:, :3], axis=1) + real[:, 3]
        assert reach.max() <= sc["cmax"] * (1 + 1e-6)


def test_group_spheres_are_tight_enough_to_cull():
    """Not a correctness property but the point of the ordering: the 16 bounds of a group are neighbours in space, so a
    group sphere of m10k is far smaller than the scene."""
    desc = rt.SceneDesc.builtin("m10k")
    groups, supers = _level(desc, 1), _level(desc, 3)
    sc = _scalars(desc)
    real = groups[groups[:, 4] == 0]
    assert np.median(real[:, 3]) < 0.1 * sc["cmax"]
    assert np.median(supers[supers[:, 4] == 0][:, 3]) < 0.35 * sc["cmax"]


def test_hooks_refuse_bad_arguments():
    desc = rt.SceneDesc.builtin("default")
    L = rt.lib()
    assert L.rt_debug_scene_spheres(C.byref(desc.desc), 4, None, 0) < 0
    assert L.rt_debug_scene_spheres(None, 0, None, 0) < 0
    assert L.rt_debug_scene_scalars(C.byref(desc.desc), None) < 0


def _light_table(desc, light):
    L = rt.lib()
    origin = (C.c_double * 5)()
    n = L.rt_debug_scene_light_table(C.byref(desc.desc), light, None, 0, origin)
    assert n > 0
    buf = np.zeros((n, 4), np.float64)
    assert L.rt_debug_scene_light_table(C.byref(desc.desc), light, buf.ctypes.data_as(C.POINTER(C.c_double)), n, origin) == n
    return buf, np.array(origin[:3]), origin[3], origin[4]


@pytest.mark.parametrize("name", ["default", "m10k"])
def test_origin_tables_hold_the_same_spheres_seen_from_the_light(name):
    """A light's table is the hierarchy re-centred on the light: C' = C - L rounded to f32, W' = |C'|^2 - R'^2 with R' no
    smaller than the radius plus the centre's rounding displacement (so the re-centred sphere contains the original), pads
    stay pads, and rmax / reach bound the slots' radii / every sphere's reach."""
    desc = rt.SceneDesc.builtin(name)
    levels = [_level(desc, k) for k in range(4)]
    allspheres = np.concatenate(levels)
    n_slots = len(levels[0])
    n_point = sum(1 for i in range(desc.desc.n_lights) if desc.desc.lights[i].kind == 0)
    assert n_point >= 1
    for light in range(n_point):
        tab, pos, reach, rmax = _light_table(desc, light)
        assert len(tab) == len(allspheres)
        real = allspheres[:, 4] == 0
        assert (tab[allspheres[:, 4] == 1, 3] >= 3e38).all()  # pads pass no ray
        want = allspheres[real, :3] - pos
        got = tab[real, :3]
        disp = np.linalg.norm(got - want, axis=1)
        assert (disp <= 1e-6 * (np.linalg.norm(want, axis=1) + 1.0)).all()  # one f32 rounding of the re-centred centre
        r2 = (got ** 2).sum(1) - tab[real, 3]  # R'^2 as the filter sees it
        assert (r2 > 0).all()
        rprime = np.sqrt(r2)
        need = allspheres[real, 3] + disp
        # W' is rounded down in f32 (R' only grows); cancellation in |C'|^2 - W' costs a few ulp of |C'|^2 here
        slack = 4e-7 * (got ** 2).sum(1) / np.maximum(rprime, 1e-30)
        assert (rprime + slack >= need).all()
        assert (rprime <= need * 1.001 + slack + 1e-6).all()  # ... and not much larger: the table still culls
        slots_real = allspheres[:n_slots, 4] == 0
        assert rmax >= allspheres[:n_slots][slots_real, 3].max()
        # reach = |L| + cmax: no sphere of the table reaches further from the light (|C - L| + R <= |L| + |C| + R)
        assert (np.linalg.norm(got, axis=1) + rprime).max() <= reach * (1 + 1e-5)
