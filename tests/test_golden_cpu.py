"""CPU: the oracle still reproduces the committed golden fixtures bit for bit."""
import glob
import os

import numpy as np
import pytest

import orc
from rust_tracer_b200 import abi
from golden.make_golden import CASES

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_golden(name):
    mk, w, h, depth, mode, spp, seed = CASES[name]
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    desc = mk()
    f, u, c = orc.render(desc.desc, abi.default_camera(w, h), depth, mode=mode, spp=spp, seed=seed)
    assert (u == g["rgb_u8"]).all()
    assert np.array_equal(f, g["rgb_f32"], equal_nan=True)
    assert [c.rays, c.shadow_rays, c.nodes] == list(g["counters"])


def test_every_fixture_has_a_case():
    names = {os.path.basename(p)[:-4] for p in glob.glob(os.path.join(HERE, "golden", "*.npz"))}
    assert names == set(CASES)


def test_oracle_thread_count_does_not_change_pixels():
    sa = orc.default_scene()
    cam = abi.default_camera(64, 48)
    f1, u1, c1 = orc.render(sa.desc, cam, 6, threads=1)
    f8, u8, c8 = orc.render(sa.desc, cam, 6, threads=8)
    assert np.array_equal(f1, f8, equal_nan=True) and (u1 == u8).all() and c1.rays == c8.rays


def test_oracle_row_subset_equals_full_frame_rows():
    sa = orc.default_scene()
    cam = abi.default_camera(80, 60)
    _, full, _ = orc.render(sa.desc, cam, 5)
    rows = np.array([3, 7, 30, 59], dtype=np.int32)
    _, sub, _ = orc.render(sa.desc, cam, 5, rows=rows)
    assert (sub == full[rows]).all()
