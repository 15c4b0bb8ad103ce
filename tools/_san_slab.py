import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from particle_simulator_b200 import SlabGroup, scenes
scene = scenes.Scene(count=2000, width=160.0, height=120.0)
pos0, vel0 = scenes.spawn(scene, seed=1)
with SlabGroup(scene, 2, precision="f64", contacts_per_particle=32) as g:
    g.upload_global(pos0, vel0)
    g.step_substeps(3)
    print([len(g.download_owned(r)[0]) for r in range(2)])
print("done")
