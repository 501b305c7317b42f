"""A few steps of the fork's Realman stock dict at 2^20 envs (one dual launch per step): target for ncu (dev tool)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench as BN
from maniskill_b200 import presets

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
gs = [BN.PresetGroup(presets.realman_pd_ee_delta_pos(), presets.REALMAN_URDF, ("r_link6", "l_link6"), B, i) for i in range(2)]
for i in range(6):
    gs[i % 2].step()
torch.cuda.synchronize()
print("done")
