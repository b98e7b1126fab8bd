import os, sys
sys.path.insert(0, "/root/repo")
import rust_tracer_b200 as rt
world = int(sys.argv[1])
sc = rt.Scene(rt.SceneDesc.builtin("m10k"))
for rep in range(3):
    st = sc.render_device(3840, 2160, 5, rank=0, world=world)
print(st["total_ms"], st["sweep_ms"], st["kernel_launches"])
