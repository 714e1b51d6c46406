import os, sys
sys.path.insert(0, os.getcwd())
from open3d_slam_private_b200 import capi as B, synth
d = synth.make_c2(M=5_000_000, N=1024, n_poses=1)
idx = B.Index(synth.feat4(d["map_xyz"]))
for _ in range(2):
    idx.self_normals(knn=20)
    print("device ms", idx.last_device_ms())
