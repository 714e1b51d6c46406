"""GPU box: feature-space 1-NN (b200_lc_feature_nn) device time, TMA-staged tiles (B200_LC_TMA=1, default) against
ordinary loads (B200_LC_TMA=0); run once per setting (the switch is read once per process)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B
lc = B.LoopClosure()
rng = np.random.default_rng(0)
for n in (6000, 20000, 60000):
    a, b = rng.uniform(0, 100, (n, 33)), rng.uniform(0, 100, (n, 33))
    ms = []
    for rep in range(6):
        ids, d2 = lc.feature_nn(a, b)
        ms.append(lc.last_device_ms())
    flops = 3.0 * 33 * n * n
    print("TMA=%s n=%d: %.3f ms (min of 5 after warm-up), %.2f TFLOP/s fp64, checksum %d" % (
        os.environ.get("B200_LC_TMA", "1"), n, min(ms[1:]), flops / (min(ms[1:]) * 1e-3) / 1e12, int(ids.sum())), flush=True)
