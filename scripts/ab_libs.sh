#!/bin/bash
# GPU box: A/B of differently compiled builds of the library on the C2 pose sweep (mean / median device ms over poses)
for lib in "$@"; do
  echo "== $lib"
  B200REG_LIB=$PWD/open3d_slam_private_b200/lib/$lib POSES=${POSES:-16} timeout 600 python scripts/pose_sweep.py 2>&1 | python -c "
import sys, numpy as np
v=[float(l.split()[2]) for l in sys.stdin if l.strip() and l.split()[0].isdigit()]
v=np.array(v[1:]); print('poses %d mean %.4f median %.4f p95 %.4f' % (len(v), v.mean(), np.median(v), np.quantile(v,.95)))"
done
