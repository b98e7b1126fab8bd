"""CPU: the sweep's work-item plan and the kernel's item decoding (csrc/sweep.cu: plan_items, decode_item -- the latter
is the same host/device function the kernel calls), through the test hooks of include/rt_b200_debug.h.  No GPU.

Invariant: the items 0 .. items-1 of a plan decode to every (ray batch, table slice) pair exactly once, the slices of a
batch cover the whole resident table, and every item number past the plan is refused -- so no ray misses a slot and none
is swept twice against one, whatever the launch size."""
import ctypes as C

import pytest

import rust_tracer_b200 as rt

SEG_PAD = 512


def _plan(r, total, heavy, per, warps, chunks=1, jobs=0):
    out = (C.c_int64 * 10)()
    assert rt.lib().rt_debug_plan_items(r, total, heavy, per, warps, chunks, jobs, out) == 0
    keys = ["items", "heavy_batches", "heavy_slices", "heavy_slice_slots", "reg_batches", "n_slices", "slice_slots",
            "tail_slices", "tail_slice_slots", "merge"]
    return dict(zip(keys, out)), out


def _decode(item, total, batch_rays, raw):
    out = (C.c_int64 * 3)()
    ok = rt.lib().rt_debug_decode_item(item, total, batch_rays, raw, out)
    assert ok in (0, 1)
    return ok, tuple(out)


@pytest.mark.parametrize("r", [1, 2, 4])
@pytest.mark.parametrize("total,heavy,per,warps,jobs", [
    (1, 0, 512, 2368, 0), (127, 0, 1024, 2368, 0), (128, 128, 10240, 2368, 0), (5000, 700, 10240, 2368, 0),
    (400_000, 30_000, 10240, 2368, 0), (400_000, 0, 10240, 592, 1), (959_010, 19_840, 10240, 592, 1),
    (3_000_000, 250_000, 10240, 2368, 0), (100_000, 100_000, 5120, 64, 0), (70_000, 10, 512, 2368, 0),
])
def test_every_batch_meets_every_slot_exactly_once(r, total, heavy, per, warps, jobs):
    p, raw = _plan(r, total, heavy, per, warps, 1, jobs)
    batch_rays = 32 * r
    batches = (total + batch_rays - 1) // batch_rays
    tail_batches = batches - p["heavy_batches"] - p["reg_batches"]
    assert 0 <= p["heavy_batches"] <= batches and p["reg_batches"] >= 0 and tail_batches >= 0
    assert p["items"] == (p["heavy_batches"] * p["heavy_slices"] + p["reg_batches"] * p["n_slices"] +
                          tail_batches * p["tail_slices"])
    used = {"heavy_slices": p["heavy_batches"], "n_slices": p["reg_batches"], "tail_slices": tail_batches}
    for slices, slots in (("heavy_slices", "heavy_slice_slots"), ("n_slices", "slice_slots"), ("tail_slices", "tail_slice_slots")):
        assert p[slots] % SEG_PAD == 0 and p[slots] >= SEG_PAD
        if not used[slices]:
            continue  # a run without batches: its shape is never decoded
        assert (p[slices] - 1) * p[slots] < per <= p[slices] * p[slots]  # the slices of a batch cover the table, none is empty
    sliced = p["n_slices"] > 1 or (p["heavy_batches"] and p["heavy_slices"] > 1) or (tail_batches and p["tail_slices"] > 1)
    assert p["merge"] == (1 if sliced else 0)  # (one table chunk here)
    seen = {}
    step = 1 if p["items"] <= 200_000 else 7  # a stride coprime to nothing special: a sample of the large plans
    for item in range(0, p["items"], step):
        ok, (batch, sl, ss) = _decode(item, total, batch_rays, raw)
        assert ok == 1, item
        want = ((p["heavy_slices"], p["heavy_slice_slots"]) if batch < p["heavy_batches"] else
                (p["n_slices"], p["slice_slots"]) if batch < p["heavy_batches"] + p["reg_batches"] else
                (p["tail_slices"], p["tail_slice_slots"]))
        assert 0 <= batch < batches and 0 <= sl < want[0] and ss == want[1]
        assert (batch, sl) not in seen, (item, seen.get((batch, sl)))
        seen[(batch, sl)] = item
    if step == 1:
        assert len(seen) == p["items"]  # with the bounds above: a bijection onto batches x their slices
    for item in (p["items"], p["items"] + 1, p["items"] + 12345, 2 ** 31):
        assert _decode(item, total, batch_rays, raw)[0] == 0


def test_small_launches_are_sliced_and_heavy_batches_cut_finer():
    p, _ = _plan(4, 400_000, 0, 10240, 2368)
    assert p["n_slices"] >= 2 and p["merge"] == 1  # 3,125 batches for 2,368 warps: the table is cut
    q, _ = _plan(4, 8_294_400, 0, 10240, 2368)
    assert q["n_slices"] == 1 and q["tail_slices"] > 1  # a whole 4K level: only the tail is cut
    h, _ = _plan(4, 3_000_000, 150_000, 10240, 2368)
    assert h["heavy_batches"] == (150_000 + 127) // 128 and h["heavy_slices"] >= h["n_slices"]
    assert rt.lib().rt_debug_plan_items(3, 1000, 0, 10240, 2368, 1, 0, (C.c_int64 * 10)()) == -1
    assert rt.lib().rt_debug_plan_items(4, 1000, 0, 1000, 2368, 1, 0, (C.c_int64 * 10)()) == -1  # not a multiple of SEG_PAD
