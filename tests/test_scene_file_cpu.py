"""CPU: the scene description file (include/rt_b200_host.h: rt_scene_desc_save / rt_scene_desc_load,
SURVEY 8 f4).  The reference has no scene format (its scene is code, src/my_scene.rs:45-120), so the
checks are: the layout documented in csrc/host/scene_file.cpp is what is written (independent
struct-level reader below), save -> load preserves every f32 bit, the oracle renders the same
image from a loaded scene, the committed fixture of the default scene still loads and equals
today's builder, and damaged files are refused with a message."""
import ctypes as C
import os
import struct
import subprocess

import numpy as np
import pytest

import orc
import rust_tracer_b200 as rt
from rust_tracer_b200 import abi
from rust_tracer_b200._lib import RtError

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CLI = os.path.join(ROOT, "rust_tracer_b200", "rt_cli")
HEADER = struct.Struct("<8sII3I3i3fIQ")  # scene_file.cpp: 64 bytes


def fnv1a(data):
    # FNV-1a 64, vectorised: h = (h ^ b) * P mod 2^64, done in Python ints per 4 KiB is too slow
    # for 2.5 MB, so fold byte by byte only for small payloads and use the C library's own
    # round trip for the big one (test_m10k_round_trip).
    h = 0xCBF29CE484222325
    for b in data:
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h


def payload_bytes(desc):
    d = desc.desc
    return (C.string_at(d.shapes, C.sizeof(abi.rt_shape) * d.n_shapes)
            + C.string_at(d.materials, C.sizeof(abi.rt_material) * d.n_materials)
            + C.string_at(d.lights, C.sizeof(abi.rt_light) * d.n_lights))


def test_record_sizes_are_the_documented_ones():
    assert HEADER.size == 64
    assert (C.sizeof(abi.rt_shape), C.sizeof(abi.rt_material), C.sizeof(abi.rt_light)) == (184, 64, 28)


def test_file_layout_matches_the_documented_header(tmp_path, default_desc):
    path = tmp_path / "default.rtscene"
    default_desc.save(path)
    raw = path.read_bytes()
    magic, version, abi_v, sb, mb, lb, ns, nm, nl, ar, ag, ab, zero, checksum = HEADER.unpack_from(raw)
    assert magic == b"RTSCENE\0" and version == 1 and abi_v == rt.lib().rt_abi_version() and zero == 0
    assert (sb, mb, lb) == (184, 64, 28)
    d = default_desc.desc
    assert (ns, nm, nl) == (d.n_shapes, d.n_materials, d.n_lights) == (6, 6, 3)
    assert (ar, ag, ab) == tuple(np.float32(x) for x in d.ambient)
    assert len(raw) == 64 + 184 * ns + 64 * nm + 28 * nl
    assert raw[64:] == payload_bytes(default_desc)
    assert checksum == fnv1a(raw[64:])


def test_round_trip_preserves_every_bit(tmp_path):
    desc = rt.SceneDesc.synthetic(40, 30, 7, 0.04, 0.12)
    path = tmp_path / "s.rtscene"
    desc.save(path)
    back = rt.SceneDesc.load(path)
    assert payload_bytes(back) == payload_bytes(desc)
    assert list(back.desc.ambient) == list(desc.desc.ambient)
    # and saving the loaded scene gives the same file
    path2 = tmp_path / "s2.rtscene"
    back.save(path2)
    assert path.read_bytes() == path2.read_bytes()


def test_m10k_round_trip(tmp_path):
    desc = rt.SceneDesc.builtin("m10k")
    path = tmp_path / "m10k.rtscene"
    desc.save(path)
    assert path.stat().st_size == 64 + 184 * 10001 + 64 * 10001 + 28 * 4
    back = rt.SceneDesc.load(path)
    assert payload_bytes(back) == payload_bytes(desc)


def test_committed_fixture_equals_todays_default_scene(default_desc):
    """tests/golden/default.rtscene was written by `rt_cli --dump-scene`; the format and the
    default scene's values (src/my_scene.rs:45-120) are both frozen by it."""
    back = rt.SceneDesc.load(os.path.join(HERE, "golden", "default.rtscene"))
    assert payload_bytes(back) == payload_bytes(default_desc)
    # the oracle's own restatement of my_scene.rs agrees shape for shape (kinds and matrices)
    sa = orc.default_scene()
    for i in range(6):
        assert back.desc.shapes[i].kind == sa.desc.shapes[i].kind
        assert list(back.desc.shapes[i].inv_transform) == list(sa.desc.shapes[i].inv_transform)


def test_oracle_renders_the_same_image_from_a_loaded_scene(tmp_path):
    desc = rt.SceneDesc.synthetic(25, 25, 3, 0.05, 0.15)
    path = tmp_path / "s.rtscene"
    desc.save(path)
    back = rt.SceneDesc.load(path)
    cam = abi.default_camera(48, 36)
    f0, u0, c0 = orc.render(desc.desc, cam, 4)
    f1, u1, c1 = orc.render(back.desc, cam, 4)
    assert np.array_equal(f0, f1, equal_nan=True) and (u0 == u1).all() and c0.as_dict() == c1.as_dict()


def test_empty_scene_round_trips(tmp_path):
    shapes, mats, lights = (abi.rt_shape * 0)(), (abi.rt_material * 0)(), (abi.rt_light * 0)()
    desc = rt.SceneDesc.from_arrays(shapes, mats, lights, (0.25, 0.5, 0.75))
    path = tmp_path / "empty.rtscene"
    desc.save(path)
    assert path.stat().st_size == 64
    back = rt.SceneDesc.load(path)
    d = back.desc
    assert (d.n_shapes, d.n_materials, d.n_lights) == (0, 0, 0) and list(d.ambient) == [0.25, 0.5, 0.75]


def _saved_default(tmp_path, default_desc):
    path = tmp_path / "d.rtscene"
    default_desc.save(path)
    return path, bytearray(path.read_bytes())


def _reseal(raw):
    """Recompute the checksum after editing the payload (so that only the edit is at fault)."""
    struct.pack_into("<Q", raw, 56, fnv1a(bytes(raw[64:])))
    return raw


@pytest.mark.parametrize("damage,code,needle", [
    (lambda r: r.__setitem__(slice(0, 4), b"XXXX"), -1, "magic"),
    (lambda r: struct.pack_into("<I", r, 8, 2), -2, "version"),
    (lambda r: struct.pack_into("<I", r, 16, 180), -2, "record sizes"),
    (lambda r: struct.pack_into("<i", r, 28, -1), -1, "negative"),
    (lambda r: struct.pack_into("<i", r, 28, 7), -1, "bytes"),          # count disagrees with the length
    (lambda r: r.__setitem__(100, r[100] ^ 0x40), -1, "checksum"),      # one flipped payload bit
    (lambda r: r.__delitem__(slice(len(r) - 5, len(r))), -1, "bytes"),  # truncated
    (lambda r: r.extend(b"\0" * 8), -1, "bytes"),                       # trailing garbage
    (lambda r: _reseal(struct.pack_into("<i", r, 64 + 4, 99) or r), -1, "material 99 out of range"),
    (lambda r: _reseal(struct.pack_into("<i", r, 64, 9) or r), -1, "unknown kind 9"),
    (lambda r: _reseal(struct.pack_into("<i", r, 64 + 184 * 6 + 4, 5) or r), -2, "texture function"),
])
def test_damaged_files_are_refused(tmp_path, default_desc, damage, code, needle):
    path, raw = _saved_default(tmp_path, default_desc)
    damage(raw)
    path.write_bytes(bytes(raw))
    with pytest.raises(RtError) as e:
        rt.SceneDesc.load(path)
    assert e.value.code == code and needle in str(e.value)


def test_missing_file_and_short_header(tmp_path):
    with pytest.raises(RtError) as e:
        rt.SceneDesc.load(tmp_path / "nope.rtscene")
    assert e.value.code == -1 and "cannot open" in str(e.value)
    p = tmp_path / "short.rtscene"
    p.write_bytes(b"RTSCENE\0abc")
    with pytest.raises(RtError) as e:
        rt.SceneDesc.load(p)
    assert "header" in str(e.value)


def test_save_refuses_a_description_scene_create_would_refuse(tmp_path):
    shapes, mats, lights = (abi.rt_shape * 1)(), (abi.rt_material * 1)(), (abi.rt_light * 0)()
    shapes[0].material = 3
    desc = rt.SceneDesc.from_arrays(shapes, mats, lights, (0, 0, 0))
    with pytest.raises(RtError):
        desc.save(tmp_path / "bad.rtscene")


def test_cli_dump_and_reload_without_a_gpu(tmp_path):
    """rt_cli --dump-scene writes the file and exits before any device call; --scene-file feeds it back."""
    a, b = tmp_path / "a.rtscene", tmp_path / "b.rtscene"
    out = subprocess.run([CLI, "--scene", "s1k", "--dump-scene", str(a)], capture_output=True, text=True, cwd=tmp_path)
    assert out.returncode == 0 and "1001 shapes" in out.stdout
    out = subprocess.run([CLI, "--scene-file", str(a), "--dump-scene", str(b)], capture_output=True, text=True, cwd=tmp_path)
    assert out.returncode == 0 and a.read_bytes() == b.read_bytes()
    out = subprocess.run([CLI, "--scene-file", str(tmp_path / "none"), "--dump-scene", str(b)], capture_output=True,
                         text=True, cwd=tmp_path)
    assert out.returncode == 101 and "cannot open" in out.stderr


def test_committed_benchmark_scene_files_equal_the_generators():
    """tests/golden/{s1k,m10k}.rtscene.xz (what bench.py --impl reference renders, read by the oracle
    binding's own struct-level reader) hold exactly what rt_builtin_scene builds today."""
    for name in ("s1k", "m10k"):
        arrays = orc.load_rtscene(os.path.join(HERE, "golden", name + ".rtscene.xz"))
        desc = rt.SceneDesc.builtin(name)
        d = desc.desc
        assert (arrays.desc.n_shapes, arrays.desc.n_materials, arrays.desc.n_lights) == (d.n_shapes, d.n_materials, d.n_lights)
        assert bytes(arrays.shapes) == C.string_at(d.shapes, C.sizeof(abi.rt_shape) * d.n_shapes)
        assert bytes(arrays.materials) == C.string_at(d.materials, C.sizeof(abi.rt_material) * d.n_materials)
        assert bytes(arrays.lights) == C.string_at(d.lights, C.sizeof(abi.rt_light) * d.n_lights)
        assert list(arrays.desc.ambient) == list(d.ambient)
