"""Regenerates tests/golden/*.npz from the CPU oracle (oracle/liborc.so).

The reference ships no golden images and cannot be built here (Rust), so these fixtures
are oracle outputs frozen at the commit that pinned the oracle against the reference's
known-answer tests (tests/test_oracle_kat.py).  They guard the oracle against regressions
and give the GPU tests a committed target that does not depend on the oracle build.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import orc  # noqa: E402
import rust_tracer_b200 as rt  # noqa: E402
from rust_tracer_b200 import abi  # noqa: E402

CASES = {
    # name: (scene builder, width, height, depth, mode, spp, seed)
    "default_basic_96x72_d8": (lambda: rt.SceneDesc.builtin("default"), 96, 72, 8, 0, 1, 0),
    "default_forest_96x72_d8": (lambda: rt.SceneDesc.builtin("default"), 96, 72, 8, 1, 1, 0),
    "default_basic_160x90_d3": (lambda: rt.SceneDesc.builtin("default"), 160, 90, 3, 0, 1, 0),
    "default_spp4_64x48_d5": (lambda: rt.SceneDesc.builtin("default"), 64, 48, 5, 0, 4, 42),
    "spheres200_96x54_d5": (lambda: rt.SceneDesc.synthetic(200, 0, 1000, 0.05, 0.2), 96, 54, 5, 0, 1, 0),
    "mixed300_96x54_d5": (lambda: rt.SceneDesc.synthetic(150, 150, 10000, 0.03, 0.1), 96, 54, 5, 0, 1, 0),
}


def scene_files():
    """The benchmark scenes (SURVEY 8d generators, host side of the library) as scene description
    files: bench.py --impl reference renders them with the oracle and must not load the product."""
    import lzma
    import tempfile
    for name in ("s1k", "m10k"):
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, name + ".rtscene")
            rt.SceneDesc.builtin(name).save(path)
            with open(path, "rb") as fh:
                raw = fh.read()
        with open(os.path.join(HERE, name + ".rtscene.xz"), "wb") as fh:
            fh.write(lzma.compress(raw, preset=9))
        print(name, len(raw), "bytes")


def main():
    scene_files()
    for name, (mk, w, h, depth, mode, spp, seed) in CASES.items():
        desc = mk()
        f, u, c = orc.render(desc.desc, abi.default_camera(w, h), depth, mode=mode, spp=spp, seed=seed)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), rgb_u8=u, rgb_f32=f,
                            counters=np.array([c.rays, c.shadow_rays, c.nodes], dtype=np.uint64),
                            params=np.array([w, h, depth, mode, spp, seed], dtype=np.int64))
        print(name, u.shape, c.rays, c.nodes)


if __name__ == "__main__":
    main()
