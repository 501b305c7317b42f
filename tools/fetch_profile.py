"""A few steps of the stock Fetch dict (arm + gripper + body + base velocity) at 2^20 envs, one fused launch per step: target for
ncu (dev tool)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench as BN
from maniskill_b200 import presets

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
gs = [BN.PresetGroup(presets.fetch_pd_ee_delta_pose(), presets.FETCH_URDF, ("gripper_link",), B, i) for i in range(2)]
for i in range(6):
    gs[i % 2].step()
torch.cuda.synchronize()
print("done")
