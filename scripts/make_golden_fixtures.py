#!/usr/bin/env python
"""Generate tests/golden/ fixtures from the reference's bundled test data (run in the build container only;
/root/reference does not exist on the GPU box).

Inputs (data files, not sources), all under /root/reference/libpointmatcher/examples/data/:
  cloud.00000.vtk / cloud.00001.vtk      ASCII VTK, xyz           (utest.cpp:83-84, BASELINE config C1)
  car_cloud400.csv / car_cloud401.csv    CSV x,y,z,nx,ny,nz       (utest.cpp:347-348)
  icp_data/default{Identity,RemoveNaN,MaxDist,BoundingBox,PointToPlaneMinDist,DistanceLimit,MaxQuantileOnAxis,
           FixStepSampling,ObservationDirection,OrientNormals,SimpleSensorNoise}DataPointsFilter.ref_trans
                                                                  (utest.cpp:81-161 golden transforms)
Reference goldens hard-coded in the reference's tests are restated in tests/golden/reference_goldens.json:
  validT3d (utest.cpp:355-359), icpSingular expected T (utest.cpp:195-197), icpIdentity (utest.cpp:203-220).

Outputs: float32 .npz (compressed) so the parity tests can run without /root/reference.
"""
import json
import os
import sys

import numpy as np

REF = "/root/reference/libpointmatcher/examples/data"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def load_vtk_points(path):
    with open(path) as f:
        lines = f.read().split("\n")
    for i, ln in enumerate(lines):
        if ln.startswith("POINTS"):
            n = int(ln.split()[1])
            vals = " ".join(lines[i + 1:]).split()
            pts = np.array(vals[: 3 * n], dtype=np.float32).reshape(n, 3)
            return pts
    raise RuntimeError("no POINTS block in " + path)


def load_csv(path):
    # delimiters " \t,;" and optional header line, as IO.cpp:558-640
    import re
    with open(path) as f:
        lines = [ln for ln in f.read().split("\n") if ln.strip()]
    first = [t for t in re.split(r"[ \t,;]+", lines[0].strip()) if t]
    try:
        [float(t) for t in first]
        header = ["x", "y", "z"]
    except ValueError:
        header = first
        lines = lines[1:]
    data = np.array([[t for t in re.split(r"[ \t,;]+", ln.strip()) if t] for ln in lines], dtype=np.float32)
    return header, data


def main():
    if not os.path.isdir(REF):
        print("reference data not present; nothing to do", file=sys.stderr)
        return 1
    os.makedirs(OUT, exist_ok=True)
    for name in ("cloud.00000", "cloud.00001"):
        pts = load_vtk_points(os.path.join(REF, name + ".vtk"))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), xyz=pts)
        print(name, pts.shape)
    for name in ("car_cloud400", "car_cloud401"):
        header, d = load_csv(os.path.join(REF, name + ".csv"))
        assert header[:3] == ["x", "y", "z"], header
        if d.shape[1] >= 6:
            assert header[3:6] == ["nx", "ny", "nz"], header
            np.savez_compressed(os.path.join(OUT, name + ".npz"), xyz=d[:, :3], normals=d[:, 3:6])
        else:
            np.savez_compressed(os.path.join(OUT, name + ".npz"), xyz=d[:, :3])
        print(name, d.shape)
    ref_trans = np.loadtxt(os.path.join(REF, "icp_data", "defaultIdentityDataPointsFilter.ref_trans"))
    goldens = {
        "validT3d": [[0.982304, 0.166685, -0.0854066, 0.0446816],
                     [-0.150189, 0.973488, 0.172524, 0.191998],
                     [0.111899, -0.156644, 0.981296, -0.0356313],
                     [0, 0, 0, 1]],
        "validT3d_tolerance": {"translation_norm": 0.1, "angle": 0.1},
        "defaultIdentityDataPointsFilter.ref_trans": ref_trans.tolist(),
        # more deterministic configs of the same test (reading-side data-point filters that are plain masks)
        **{n + ".ref_trans": np.loadtxt(os.path.join(REF, "icp_data", n + ".ref_trans")).tolist()
           for n in ("defaultRemoveNaNDataPointsFilter", "defaultMaxDistDataPointsFilter",
                     "defaultBoundingBoxDataPointsFilter", "defaultPointToPlaneMinDistDataPointsFilter",
                     "defaultDistanceLimitDataPointsFilter", "defaultMaxQuantileOnAxisDataPointsFilter",
                     "defaultFixStepSamplingDataPointsFilter", "defaultObservationDirectionDataPointsFilter",
                     "defaultOrientNormalsDataPointsFilter", "defaultSimpleSensorNoiseDataPointsFilter")},
        "ref_trans_criterion": "median|curT*data - refT*data| / median|curT*data| < 0.05 (utest.cpp:147-159)",
        "icpSingular_expected": [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 1], [0, 0, 0, 1]],
        "icpIdentity_epsilon": 1e-4,
    }
    with open(os.path.join(OUT, "reference_goldens.json"), "w") as f:
        json.dump(goldens, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
