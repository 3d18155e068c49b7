"""The reference's loop compiled against the C++ DROP-IN headers (tests/cpp/dropin_loop.cpp is the
wrap.h:237-342 / hough3dlines.cpp:200-317 loop verbatim in structure, using Hough / PointCloud /
Vector3d / initHoughSpace from pointcloud_segmentation_b200/include) must print the oracle's lines."""
import os
import re
import subprocess

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "pointcloud_segmentation_b200", "lib", "dropin_loop")


@pytest.mark.parametrize("key", ["testdata", "synthetic_c", "tower800"])
def test_dropin_classes_run_the_reference_loop(tmp_path, key):
    if not os.path.exists(EXE):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "pointcloud_segmentation_b200"), "dropin"], check=True)
    gold = G.load()
    pts = gold[f"{key}_points"]
    g, dx, minvotes, nlines = G.case_params(gold, key)
    dat = tmp_path / "cloud.dat"
    with open(dat, "w") as f:
        f.write("# test cloud\n")
        for p in pts:
            f.write(",".join(repr(float(v)) for v in p) + "\n")
    out = subprocess.run([EXE, str(dat), str(g), repr(dx), str(minvotes), str(nlines)], capture_output=True,
                         text=True, check=True).stdout.strip().splitlines()
    t = f"{key}_trace_"
    n_lines = int(gold[t + "n_lines"])
    assert len(out) == n_lines
    for k, line in enumerate(out):
        m = re.match(r"votes=(\d+) npoints=(\d+) a=(\S+) b=(\S+)", line)
        votes, npts = int(m.group(1)), int(m.group(2))
        a = np.array([float(v) for v in m.group(3).split(",")])
        b = np.array([float(v) for v in m.group(4).split(",")])
        assert votes == int(gold[f"{t}it{k}_votes"])
        assert npts == len(gold[f"{t}it{k}_inliers2"])
        assert np.array_equal(a, gold[f"{t}it{k}_a2"]) and np.array_equal(b, gold[f"{t}it{k}_b2"])
