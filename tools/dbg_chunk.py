import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import rust_tracer_b200 as rt
desc = rt.SceneDesc.synthetic(400, 400, 4242, 0.03, 0.1)
sc = rt.Scene(desc)
f0,u0,s0 = sc.render(192,108,5)
os.environ["RT_B200_TABLE_CAP"]="128"
f1,u1,s1 = sc.render(192,108,5)
print({k:(s0[k],s1[k]) for k in ("rays","exact_tests","kernel_launches","filter_flops")})
