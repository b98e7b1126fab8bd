import sys, numpy as np
sys.path.insert(0, "/root/repo")
import rust_tracer_b200 as rt
desc = rt.SceneDesc.synthetic(300, 300, 5, 0.03, 0.1)
outs = []
for v in (0, 1, 2):
    with rt.tuned(primary_tiles=v):
        sc = rt.Scene(desc)
        f, u, st = sc.render(203, 77, 4, spp=4, seed=9)
        outs.append((f, u, st["rays"], st["nodes"]))
        sc.close()
for o in outs[1:]:
    assert np.array_equal(outs[0][0], o[0], equal_nan=True) and (outs[0][1] == o[1]).all() and outs[0][2:] == o[2:]
print("tiles 0/1/2 at 4 spp: same image, rays", outs[0][2])
