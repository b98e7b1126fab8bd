"""GPU parity tests proper: the CUDA path (through the C ABI of librt_b200.so) against the
oracle on the same inputs, against the committed golden fixtures, and -- at BASELINE.json's
full sizes -- through size-independent properties.

Bar (BASELINE.json north_star): per-channel |gpu - reference| <= 1/255 on >= 99.9 % of
pixels, max abs error stated.  In practice the path is bit-exact on the u8 image because
every decision uses the reference's operand order; the tolerance is still what is asserted.
"""
import os

import numpy as np
import pytest

import orc
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi
from golden.make_golden import CASES

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL_FRACTION = 0.999  # >= 99.9 % of pixels within 1/255


def assert_parity(u8, ref8, what):
    d = np.abs(u8.astype(np.int32) - ref8.astype(np.int32)).max(axis=-1)
    frac = float((d <= 1).mean())
    print(f"{what}: {frac:.6f} of pixels within 1/255, max abs error {int(d.max())}/255, "
          f"bit-exact pixels {float((d == 0).mean()):.6f}")
    assert frac >= TOL_FRACTION, f"{what}: only {frac:.5f} of pixels within 1/255 (max {d.max()})"


def gpu_render(desc, w, h, depth, **kw):
    sc = rt.Scene(desc)
    try:
        return sc.render(w, h, depth, **kw)
    finally:
        sc.close()


@pytest.mark.parametrize("name", sorted(CASES))
def test_matches_golden_fixture(name):
    mk, w, h, depth, mode, spp, seed = CASES[name]
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    flags = abi.RT_FLAG_FOREST_SHADING if mode else 0
    f, u, st = gpu_render(mk(), w, h, depth, spp=spp, seed=seed, flags=flags)
    assert_parity(u, g["rgb_u8"], name)
    assert [st["rays"], st["shadow_rays"], st["nodes"]] == list(g["counters"])
    fin = np.isfinite(g["rgb_f32"]) & np.isfinite(f)
    assert np.abs(f[fin] - g["rgb_f32"][fin]).max() < 1e-4


def test_config1_default_512_depth8_full_frame():
    """BASELINE config 1: the reference CLI's defaults (src/cli.rs:43,51,59)."""
    desc = rt.SceneDesc.builtin("default")
    f, u, st = gpu_render(desc, 512, 512, 8)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(512, 512), 8)
    assert_parity(u, ou, "config 1")
    assert st["rays"] == oc.rays and st["nodes"] == oc.nodes and st["shadow_rays"] == oc.shadow_rays
    assert st["sphere_tests"] == oc.sphere_tests and st["cube_tests"] == oc.cube_tests


def test_config2_default_1080p_depth3_full_frame():
    """BASELINE config 2, including the pixel rows/columns that sit exactly on checkerboard
    edges (hard part 1: rows 108, 252, 396, 684 and columns 192, 448, ...)."""
    desc = rt.SceneDesc.builtin("default")
    f, u, st = gpu_render(desc, 1920, 1080, 3)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(1920, 1080), 3)
    assert_parity(u, ou, "config 2")
    for v in (108, 252, 396, 684, 828, 1020, 900, 860):
        assert_parity(u[v:v + 1], ou[v:v + 1], f"config 2 aligned row {v}")
    for c in (192, 448, 704, 1216, 1472, 1728):
        assert_parity(u[:, c:c + 1], ou[:, c:c + 1], f"config 2 aligned column {c}")
    assert st["rays"] == oc.rays


def _row_subset(h):
    rows = sorted(set(range(7, h, 16)) | {h // 2, h - 1})
    return np.array(rows, dtype=np.int32)


def aligned_rows(h):
    """Pixel rows whose primary rays meet the floor of the synthetic scenes (y = -2, camera at z = -8)
    at an integer z: the ray through row v has y = -(v - h/2) * 6/h at z = 0, so it reaches y = -2 at
    z = 8 h / (3 (v - h/2)) - 8; for v - h/2 = (8h/3) / m that is z = m - 8.  There
    the checkerboard (src/my_scene.rs:26-43) is decided by the last bit of p.z (hard part 1)."""
    return sorted({h // 2 + (8 * h // 3) // m for m in (10, 9, 8, 6) if (8 * h // 3) % m == 0})


@pytest.mark.parametrize("name,depth", [("s1k", 5), ("m10k", 5)])
def test_config3_4_true_resolution_row_subset(name, depth):
    """BASELINE configs 3 / 4 at 3840x2160: the oracle renders rows v = 7 mod 16, the
    checkerboard-aligned floor rows 1656 / 1720 / 1800 / 2040, the middle row and the last row at the
    true resolution; every aligned row is asserted on its own, bit for bit."""
    w, h = 3840, 2160
    desc = rt.SceneDesc.builtin(name)
    f, u, st = gpu_render(desc, w, h, depth, want_f32=False)
    special = aligned_rows(h)
    assert special == [1656, 1720, 1800, 2040]
    rows = np.array(sorted(set(range(7, h, 16)) | set(special) | {h // 2, h - 1}), dtype=np.int32)
    _, ou, _ = orc.render(desc.desc, abi.default_camera(w, h), depth, rows=rows, want_f32=False, threads=os.cpu_count())
    assert_parity(u[rows], ou, f"{name} 4K rows 7 mod 16 + aligned rows")
    for v in special + [h // 2, h - 1]:
        i = int(np.searchsorted(rows, v))
        assert (u[v] == ou[i]).all(), f"{name} 4K row {v}: {int((u[v] != ou[i]).any(axis=-1).sum())} pixels differ"
    assert (u[rows] == ou).all(), f"{name} 4K: {int((u[rows] != ou).any(axis=-1).sum())} pixels differ from the oracle"


def test_config5_true_size_8k_spp4_depth8():
    """BASELINE config 5 at its real size: 7680x4320, 4 spp jittered, depth 8, m10k -- a frame of 133 M
    samples, cast in 16 bands.  The oracle renders 10 rows of it, among them the rows where the floor
    is reached at an integer z (3312, 3440, 3600, 4080); bit-exact on the RGB8 rows."""
    w, h, depth, spp, seed = 7680, 4320, 8, 4, 2024
    desc = rt.SceneDesc.builtin("m10k")
    f, u, st = gpu_render(desc, w, h, depth, spp=spp, seed=seed, want_f32=False)
    special = aligned_rows(h)
    assert special == [3312, 3440, 3600, 4080]
    rows = np.array(sorted(set(special) | {7, 1000, 2160, 2700, 3000, h - 1}), dtype=np.int32)
    _, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), depth, rows=rows, spp=spp, seed=seed, want_f32=False,
                           threads=os.cpu_count())
    assert_parity(u[rows], ou, "config 5 at 7680x4320")
    assert (u[rows] == ou).all(), f"config 5: {int((u[rows] != ou).any(axis=-1).sum())} pixels differ from the oracle"
    assert st["pixels"] == w * h and st["rays"] > 10 * w * h * spp


def test_filter_equals_all_pairs_audit_on_a_4k_share():
    """One rank's share of the 4K m10k frame as 8 GPUs split it (row tiles of 8, rank 3): the image and
    the counters with the bound filter, the group / quad culling, the node order and shadow rays that
    stop at their first occluder must equal those with every (ray, shape) pair tested exactly in shape
    order -- the reference's literal loop (src/scene/mod.rs:98-116)."""
    desc = rt.SceneDesc.builtin("m10k")
    w, h, depth = 3840, 2160, 5
    sc = rt.Scene(desc)
    try:
        f1, u1, s1 = sc.render(w, h, depth, rank=3, world=8, rows_per_tile=8)
        f2, u2, s2 = sc.render(w, h, depth, rank=3, world=8, rows_per_tile=8, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    finally:
        sc.close()
    assert (u1 == u2).all(), f"{int((u1 != u2).any(axis=-1).sum())} pixels differ"
    assert np.array_equal(f1, f2, equal_nan=True)
    for k in ("pixels", "rays", "shadow_rays", "nodes"):
        assert s1[k] == s2[k], k
    assert s2["exact_tests"] == s2["rays"] * desc.desc.n_shapes
    assert s1["exact_tests"] < s2["exact_tests"] / 100


@pytest.mark.parametrize("name,depth", [("s1k", 5), ("m10k", 5)])
def test_config3_4_full_frame_at_eighth_resolution(name, depth):
    desc = rt.SceneDesc.builtin(name)
    w, h = 480, 270
    f, u, st = gpu_render(desc, w, h, depth)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), depth)
    assert_parity(u, ou, f"{name} {w}x{h}")
    assert st["rays"] == oc.rays and st["nodes"] == oc.nodes


def test_config5_spp4_depth8_row_subset():
    """BASELINE config 5 (4 spp jitter, depth 8) on m10k, reduced to 960x540 (the real size: test_config5_true_size_8k_spp4_depth8)."""
    desc = rt.SceneDesc.builtin("m10k")
    w, h = 960, 540
    f, u, st = gpu_render(desc, w, h, 8, spp=4, seed=2024, want_f32=False)
    rows = _row_subset(h)[::2]
    _, ou, _ = orc.render(desc.desc, abi.default_camera(w, h), 8, rows=rows, spp=4, seed=2024, want_f32=False)
    assert_parity(u[rows], ou, "config 5 (reduced)")


def test_filter_equals_all_pairs_audit():
    """The bound filter may only ever reject pairs the exact test rejects: the image with the
    filter must equal the image with every (ray, shape) pair tested exactly."""
    desc = rt.SceneDesc.builtin("m10k")
    w, h = 320, 180
    f1, u1, s1 = gpu_render(desc, w, h, 5)
    f2, u2, s2 = gpu_render(desc, w, h, 5, flags=abi.RT_FLAG_AUDIT_ALL_PAIRS)
    assert np.array_equal(f1, f2, equal_nan=True) and (u1 == u2).all()
    assert s1["rays"] == s2["rays"] and s1["nodes"] == s2["nodes"]
    assert s1["exact_tests"] < s2["exact_tests"] / 20  # the filter does remove the bulk of the pairs


def test_row_tiles_reassemble_to_the_full_frame():
    """Interleaved row tiles of 2 / 4 / 8 ranks cover the frame once and give the same pixels."""
    desc = rt.SceneDesc.builtin("s1k")
    w, h = 640, 360
    sc = rt.Scene(desc)
    _, full, _ = sc.render(w, h, 5, want_f32=False)
    for world in (2, 8):
        out = np.zeros_like(full)
        rays = 0
        for rank in range(world):
            _, tile, st = sc.render(w, h, 5, rank=rank, world=world, rows_per_tile=8, want_f32=False)
            out[rt.tile_rows(h, rank, world, 8)] = tile
            rays += st["rays"]
        assert (out == full).all()
    sc.close()


def test_deterministic_and_depth_semantics():
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    f1, u1, s1 = sc.render(200, 150, 6)
    f2, u2, s2 = sc.render(200, 150, 6)
    assert np.array_equal(f1, f2, equal_nan=True) and s1["rays"] == s2["rays"]
    f0, u0, s0 = sc.render(64, 48, 0)  # depth 0 -> BLACK, no rays (src/render.rs:43-45)
    assert (u0 == 0).all() and s0["rays"] == 0
    fd1, _, sd1 = sc.render(64, 48, 1)  # depth 1 -> primary + shadow rays only
    assert sd1["rays"] == 64 * 48 + sd1["shadow_rays"] and sd1["nodes"] <= 64 * 48
    sc.close()


def test_unclamped_f32_and_saturating_u8():
    """Colours above 1 and below 0 survive in f32 and only saturate at as_u8 (color.rs:43-46)."""
    desc = rt.SceneDesc.builtin("default")
    f, u, _ = gpu_render(desc, 256, 256, 8)
    assert (f > 1.0).any() and (f < 0.0).any()
    assert (u[f > 1.0] == 255).all() and (u[f < 0.0] == 0).all()


def test_material_update_changes_pixels_without_new_geometry():
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    _, before, _ = sc.render(160, 120, 4, want_f32=False)
    mats = desc.materials_copy()
    mats[1].diffuse.rgb[:] = [0.0, 1.0, 0.0]  # the "blue" sphere turns green (gui.rs:221-235)
    sc.update_materials(mats)
    _, after, _ = sc.render(160, 120, 4, want_f32=False)
    ref_desc = rt.SceneDesc.from_arrays(
        (abi.rt_shape * 6)(*[desc.desc.shapes[i] for i in range(6)]), mats,
        (abi.rt_light * 3)(*[desc.desc.lights[i] for i in range(3)]), list(desc.desc.ambient))
    _, ref, _ = orc.render(ref_desc.desc, abi.default_camera(160, 120), 4, want_f32=False)
    assert (before != after).any()
    assert_parity(after, ref, "after set_diffuse")
    sc.close()


def test_ragged_and_tiny_frames():
    desc = rt.SceneDesc.builtin("default")
    for w, h in ((1, 1), (3, 2), (33, 17), (257, 5)):
        f, u, st = gpu_render(desc, w, h, 4)
        of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), 4)
        assert (u == ou).all() and st["rays"] == oc.rays


def test_scene_without_shapes_or_lights():
    empty = rt.SceneDesc.from_arrays((abi.rt_shape * 0)(), (abi.rt_material * 0)(), (abi.rt_light * 0)(), (0.1, 0.1, 0.1))
    f, u, st = gpu_render(empty, 32, 16, 5)
    assert (u == 0).all() and st["rays"] == 32 * 16 and st["nodes"] == 0


def test_error_codes():
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    with pytest.raises(rt.RtError) as e:
        sc.render(0, 10, 3)
    assert e.value.code == abi.RT_ERR_INVALID
    with pytest.raises(rt.RtError) as e:
        sc.render(8, 8, 3, spp=2)
    assert e.value.code == abi.RT_ERR_UNSUPPORTED
    with pytest.raises(rt.RtError):
        sc.update_materials((abi.rt_material * 2)())
    with pytest.raises(rt.RtError):
        rt.Scene(desc, device=99)
    sc.close()


def _nan_scene():
    """A light placed exactly on a shadow-ray origin: norm(0) = NaN direction (vector3.rs:91-94).
    The reference then 'hits' the first sphere with t = NaN and the pixel becomes NaN -> 0."""
    base = rt.SceneDesc.builtin("default")
    shapes = (abi.rt_shape * 2)(base.desc.shapes[3], base.desc.shapes[1])  # back wall plane, blue sphere
    shapes[0].material, shapes[1].material = 0, 1
    mats = (abi.rt_material * 2)(base.desc.materials[3], base.desc.materials[1])
    lights = (abi.rt_light * 2)(base.desc.lights[0], base.desc.lights[1])
    zs = np.float32(2.0) + np.float32(0.0002) * np.float32(-1.0)  # point + 0.0002 * normal for the centre pixel
    lights[1].pos[:] = [0.0, 0.0, float(zs)]
    return rt.SceneDesc.from_arrays(shapes, mats, lights, (0.1, 0.1, 0.1)), base


def test_nan_rays_follow_the_reference():
    desc, _keep = _nan_scene()
    w, h = 64, 64  # pixel (32, 32) looks straight down +z
    f, u, st = gpu_render(desc, w, h, 4)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(w, h), 4)
    assert np.isnan(of[32, 32]).all(), "the oracle scene no longer produces the NaN pixel"
    assert (np.isnan(f) == np.isnan(of)).all()
    assert (u == ou).all() and (u[32, 32] == 0).all()
    assert st["rays"] == oc.rays


def test_ambient_light_source_is_nan_like_the_reference():
    """AmbientLight as a LightSource has direction (0,0,0) (scene/mod.rs:236-238); the Blinn
    half vector normalises it -> NaN specular in the reference.  Same here."""
    base = rt.SceneDesc.builtin("default")
    lights = (abi.rt_light * 1)()
    lights[0].kind = abi.RT_LIGHT_AMBIENT
    lights[0].color[:] = [0.3, 0.3, 0.3]
    desc = rt.SceneDesc.from_arrays((abi.rt_shape * 6)(*[base.desc.shapes[i] for i in range(6)]), base.materials_copy(),
                                    lights, (0.1, 0.1, 0.1))
    f, u, st = gpu_render(desc, 48, 48, 3)
    of, ou, oc = orc.render(desc.desc, abi.default_camera(48, 48), 3)
    assert (np.isnan(f) == np.isnan(of)).all() and np.isnan(of).any()
    assert (u == ou).all() and st["rays"] == oc.rays and st["shadow_rays"] == 0


def test_frame_rows_flag_writes_rows_in_place():
    """RT_FLAG_FRAME_ROWS (rt_render_device): each rank's rows land at their own place of a
    whole-frame buffer, so two ranks writing into one buffer assemble the frame with no gather
    (the multi-GPU path maps that buffer across processes: distributed.PeerFrame)."""
    import torch
    desc = rt.SceneDesc.builtin("default")
    w, h, depth = 200, 131, 4  # 131 rows: a ragged last tile
    sc = rt.Scene(desc)
    try:
        ref = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda")
        sc.render_device(w, h, depth, d_rgb_u8=ref.data_ptr())
        full = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda")
        ref_f = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
        sc.render_device(w, h, depth, d_rgb_f32=ref_f.data_ptr())
        full_f = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
        for rank in range(3):
            sc.render_device(w, h, depth, d_rgb_u8=full.data_ptr(), d_rgb_f32=full_f.data_ptr(), rank=rank, world=3,
                             rows_per_tile=8, flags=abi.RT_FLAG_FRAME_ROWS)
        torch.cuda.synchronize()
        assert bool((full == ref).all())
        assert torch.equal(full_f, ref_f)
        with pytest.raises(Exception):
            sc.render(w, h, depth, flags=abi.RT_FLAG_FRAME_ROWS)  # host-buffer call: compacted rows only
    finally:
        sc.close()


def test_frame_signal_words():
    """rt_frame_signal / rt_frame_wait: stream-ordered words behind the pixels of a frame; a wait for a
    value nobody signals ends with RT_ERR_TIMEOUT instead of hanging the GPU."""
    import ctypes as C
    import time
    import torch
    from rust_tracer_b200._lib import RtError, check, lib
    L = lib()
    nbytes = 100 * 30 * 3  # not a multiple of 256: the words start at the next 256-byte boundary
    ptr = C.c_void_p()
    check(L.rt_frame_alloc(0, nbytes, C.byref(ptr)))
    try:
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(L.rt_frame_wait(0, ptr, nbytes, 0, abi.RT_FRAME_SIGNAL_WORDS, 0, 100, st))  # zeroed at allocation
        for word in range(40):  # more words than one warp has lanes
            check(L.rt_frame_signal(0, ptr, nbytes, word, 7, st))
        check(L.rt_frame_signal(0, ptr, nbytes, abi.RT_FRAME_WORD_RELEASED, 6, st))
        check(L.rt_frame_wait(0, ptr, nbytes, 0, 40, 7, 1000, st))
        check(L.rt_frame_wait(0, ptr, nbytes, abi.RT_FRAME_WORD_RELEASED, 1, 6, 1000, st))
        t0 = time.perf_counter()
        with pytest.raises(RtError) as e:
            check(L.rt_frame_wait(0, ptr, nbytes, 0, 41, 7, 50, st))  # word 40 was never signalled
        assert e.value.code == abi.RT_ERR_TIMEOUT and time.perf_counter() - t0 < 5.0
        with pytest.raises(RtError):
            check(L.rt_frame_wait(0, ptr, nbytes, 250, 10, 1, 50, st))  # range past the last word
        # the pixels are untouched by the signalling
        from rust_tracer_b200.distributed import _RawCudaBuffer
        px = torch.as_tensor(_RawCudaBuffer(ptr.value, (nbytes,)), device="cuda")
        px.fill_(9)
        check(L.rt_frame_signal(0, ptr, nbytes, 0, 8, st))
        check(L.rt_frame_wait(0, ptr, nbytes, 0, 1, 8, 1000, st))
        assert bool((px == 9).all())
    finally:
        L.rt_frame_release(0, ptr, 0)


def test_scene_loaded_from_a_file_renders_the_same_frame(tmp_path):
    """Scene description file (rt_scene_desc_save / _load): a scene that went
    through the file is the same scene to the device, byte for byte, and matches the oracle
    rendering the loaded description."""
    desc = rt.SceneDesc.synthetic(120, 120, 11, 0.04, 0.12)
    path = tmp_path / "s.rtscene"
    desc.save(path)
    back = rt.SceneDesc.load(path)
    f0, u0, st0 = gpu_render(desc, 160, 90, 4)
    f1, u1, st1 = gpu_render(back, 160, 90, 4)
    assert (u0 == u1).all() and np.array_equal(f0, f1, equal_nan=True) and st0["rays"] == st1["rays"]
    _, ou, oc = orc.render(back.desc, abi.default_camera(160, 90), 4)
    assert_parity(u1, ou, "scene file")
    assert st1["rays"] == oc.rays


def test_multi_gpu_frame_equals_single_gpu_frame():
    """rt_multi (one host call drives n devices, include/rt_b200.h) against the single-device frame:
    host buffers (every device copies its own rows home) and the peer-mapped device frame (every
    device's resolve kernel stores into device 0's memory).  Needs two devices; with one, the same
    calls run on a one-device rt_multi."""
    import torch
    n_dev = rt.device_count()
    devices = list(range(min(n_dev, 8)))
    desc = rt.SceneDesc.builtin("s1k")
    w, h, depth = 1280, 715, 5  # 715 rows: a ragged last tile
    sc = rt.Scene(desc)
    f0, u0, s0 = sc.render(w, h, depth)
    sc.close()
    for devs in ([0], devices[:2], devices):
        m = rt.MultiScene(desc, devs)
        try:
            f1, u1, s1 = m.render(w, h, depth)
            assert (u1 == u0).all() and np.array_equal(f1, f0, equal_nan=True), devs
            for k in ("pixels", "rays", "shadow_rays", "nodes"):
                assert s1[k] == s0[k], (k, devs)
            frame = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda:%d" % devs[0])
            m.render_device(w, h, depth, d_rgb_u8=frame.data_ptr())
            torch.cuda.synchronize(devs[0])
            assert (frame.cpu().numpy() == u0).all(), devs
            # 4 spp, other tile height
            _, u4, _ = m.render(w, h, depth, spp=4, seed=9, rows_per_tile=5, want_f32=False)
        finally:
            m.close()
        sc = rt.Scene(desc)
        _, r4, _ = sc.render(w, h, depth, spp=4, seed=9, want_f32=False)
        sc.close()
        assert (u4 == r4).all(), devs
    u5, st5 = rt.render_multi(desc, devices, 320, 180, 3)
    sc = rt.Scene(desc)
    _, r5, _ = sc.render(320, 180, 3, want_f32=False)
    sc.close()
    assert (u5 == r5).all()
    with pytest.raises(rt.RtError):
        rt.MultiScene(desc, [0, 0])
    with pytest.raises(rt.RtError):
        rt.MultiScene(desc, [99])


def test_scene_may_be_destroyed_before_its_forest():
    """rt_scene_destroy with a live forest defers the free to the last rt_forest_destroy (either order
    of the two calls is fine); reflectivity / refraction_index are frozen while a forest is alive."""
    desc = rt.SceneDesc.builtin("default")
    sc = rt.Scene(desc)
    forest = sc.build_forest(96, 64, 4)
    f0, u0, _ = forest.eval()
    mats = desc.materials_copy()
    mats[1].refraction_index = 1.7
    with pytest.raises(rt.RtError) as e:
        sc.update_materials(mats)
    assert e.value.code == abi.RT_ERR_UNSUPPORTED
    mats = desc.materials_copy()
    mats[1].diffuse.rgb[:] = [0.0, 1.0, 0.0]
    sc.update_materials(mats)  # colours may change
    sc.close()                 # scene first ...
    f1, u1, _ = forest.eval()  # ... the forest still works
    assert (u1 != u0).any()
    forest.close()
