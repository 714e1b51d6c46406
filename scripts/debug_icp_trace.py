"""Debug helper (GPU box): per-iteration |T_iter(device) - T_iter(oracle)| free-running and one-step."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle_py as O
from open3d_slam_private_b200 import capi as B


def to_b(p):
    names = ("knn maxDist use_trimmed trimmed_ratio use_maxdist_filter maxdist_filter use_surface_normal_filter "
             "surface_normal_max_angle minimizer max_iterations use_differential min_diff_rot min_diff_trans "
             "smooth_length").split()
    return B.default_icp_params(**{n: getattr(p, n) for n in names})


g = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
c400 = np.load(g + "/car_cloud400.npz"); c401 = np.load(g + "/car_cloud401.npz")
ref4, refn, read4 = O.feat4(c400["xyz"]), c400["normals"], O.feat4(c401["xyz"])
for kw in (dict(minimizer=1, trimmed_ratio=0.8), dict(), dict(use_differential=0, max_iterations=7)):
    p = O.default_params(**kw)
    o = O.icp(p, ref4, refn, read4)
    icp = B.ICP(to_b(p)); icp.init_reference(ref4, refn)
    r = icp.compute(read4)
    lg = icp.log()
    of = O.icp(p, ref4, refn, read4, forced_T_iters=lg["T_iters"])
    print(kw, "iters", r["iterations"], o["iterations"], "final dT", np.abs(r["T"] - o["T"]).max())
    for i in range(len(lg["P"])):
        print("  it %2d free %.2e step %.2e P %d %d %d" % (i, np.abs(lg["T_iters"][i] - o["T_iters"][i]).max() if i < o["iterations"] else -1,
              np.abs(lg["T_iters"][i] - of["T_iters"][i]).max(), lg["P"][i], o["P"][i] if i < o["iterations"] else -1, of["P"][i]))
