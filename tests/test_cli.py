"""rt_cli: the host driver with the reference's command line (src/cli.rs:34-175) and stage
printouts (src/main.rs:27-291).  CPU part: argument handling and the loud failure without a
GPU; GPU part: the PNG it writes holds the oracle's pixels."""
import os
import subprocess
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
CLI = os.path.join(os.path.dirname(HERE), "rust_tracer_b200", "rt_cli")


def run(*args, **kw):
    return subprocess.run([CLI, *args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, **kw)


def read_png_rgb8(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n = int.from_bytes(data[pos:pos + 4], "big")
        kind, body = data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n]
        assert zlib.crc32(kind + body) == int.from_bytes(data[pos + 8 + n:pos + 12 + n], "big")
        if kind == b"IHDR":
            w, h = int.from_bytes(body[:4], "big"), int.from_bytes(body[4:8], "big")
            assert body[8:] == bytes([8, 2, 0, 0, 0])  # 8-bit RGB, no interlace
        elif kind == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3)


def test_cli_rejects_bad_arguments_like_the_reference():
    assert run("--method", "fancy").returncode == 101      # panic!("Unexpected value provided for `--method`")
    assert "Unexpected value" in run("--method", "fancy").stderr
    assert "Expected integer for width" in run("-w", "abc").stderr  # cli.rs:123
    assert run("--gui").returncode == 101


def test_cli_prints_the_config_line_and_parses_h_as_height():
    out = run("-w", "40", "-h", "30", "-d", "3", "--method", "RayForest", "--gui").stdout  # method is case-insensitive
    assert "Config { width: 40, height: 30, depth: 3," in out and "method: RayForest" in out


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="only meaningful without a GPU")
def test_cli_fails_loudly_without_a_gpu():
    r = run("-w", "8", "-h", "8")
    assert r.returncode == 101 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_cli_basic_and_rayforest_images_equal_the_oracle(tmp_path):
    import orc
    import rust_tracer_b200 as rt
    from rust_tracer_b200 import abi
    desc = rt.SceneDesc.builtin("default")
    w, h, d = 96, 64, 5
    for method, mode in (("basic", 0), ("rayforest", 1)):
        out = str(tmp_path / (method + ".png"))
        r = run("-w", str(w), "-h", str(h), "-d", str(d), "--method", method, "--stats", "--out", out)
        assert r.returncode == 0, r.stderr
        assert ("render_basic_to_file:" if mode == 0 else "render_forest_to_file:") in r.stdout
        _, ou, _ = orc.render(desc.desc, abi.default_camera(w, h), d, mode=mode)
        assert (read_png_rgb8(out) == ou).all()
        if mode:
            assert "Number of Trees: %d" % (w * h) in r.stdout and "Number of Shapes: 6" in r.stdout


@pytest.mark.gpu
def test_cli_bench_filter_reports_trees_with_the_blue_sphere():
    r = run("-w", "64", "-h", "48", "-d", "4", "--method", "rayforest", "bench", "-n", "3", "-f")
    assert r.returncode == 0, r.stderr
    assert "Forest Size: 3072" in r.stdout and "Trees Evaluated:" in r.stdout and "Avg Per Op:" in r.stdout
    r = run("-w", "64", "-h", "48", "-d", "4", "bench", "-n", "2")
    assert r.returncode == 0 and r.stdout.count("render_scene:") == 2


@pytest.mark.gpu
def test_cli_renders_a_scene_file(tmp_path):
    """--scene-file: the committed fixture of the default scene gives the default scene's image."""
    import orc
    from rust_tracer_b200 import abi
    w, h, d = 80, 60, 4
    out = str(tmp_path / "f.png")
    r = run("-w", str(w), "-h", str(h), "-d", str(d), "--scene-file", os.path.join(HERE, "golden", "default.rtscene"),
            "--out", out)
    assert r.returncode == 0, r.stderr
    _, ou, _ = orc.render(orc.default_scene().desc, abi.default_camera(w, h), d)
    assert (read_png_rgb8(out) == ou).all()


@pytest.mark.gpu
def test_cli_gpus_flag_renders_the_same_image(tmp_path):
    """--gpus N: Basic mode through one rt_multi_render call per frame on the first N devices;
    the PNG equals the one-device PNG.  With one device in the box only the refusals are checked."""
    import rust_tracer_b200 as rt
    n = rt.device_count()
    assert "only %d CUDA devices" % n in run("--gpus", str(n + 1), "-w", "8", "-h", "8").stderr
    assert "--method basic" in run("--gpus", "2", "--method", "rayforest", "-w", "8", "-h", "8").stderr
    if n < 2:
        return
    w, h, d = 200, 131, 4
    one, many = str(tmp_path / "one.png"), str(tmp_path / "many.png")
    assert run("-w", str(w), "-h", str(h), "-d", str(d), "--out", one).returncode == 0
    r = run("-w", str(w), "-h", str(h), "-d", str(d), "--gpus", str(min(n, 8)), "--out", many)
    assert r.returncode == 0, r.stderr
    assert (read_png_rgb8(one) == read_png_rgb8(many)).all()
