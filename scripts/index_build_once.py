import os, sys
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
from open3d_slam_private_b200 import capi as B, synth
d = synth.make_c2(M=5_000_000, N=1024, n_poses=1)
ref4 = synth.feat4(d["map_xyz"])
idx = B.Index(ref4, d["map_normals"])
idx.update(ref4, d["map_normals"])
print(idx.info())
