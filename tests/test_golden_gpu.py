"""GPU path against the committed golden vectors recorded from the unmodified reference
(the reference's own fixtures testdata.dat / synthetic_*.dat / attpc_*.dat + one synthetic)."""
import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["fused", "multikernel"])
def loop_form(request, h3d):
    """Every test runs with the fused persistent loop kernel and with the general multi-kernel loop."""
    old = h3d.set_fused(request.param == "fused")
    yield request.param
    h3d.set_fused(old)


@pytest.fixture(scope="module")
def gold():
    return G.load()


@pytest.mark.parametrize("key", G.CASES)
def test_accumulator_matches_reference_golden(h3d, gold, key):
    pts = gold[f"{key}_points"]
    g, dx, _, _ = G.case_params(gold, key)
    x, _ = h3d.shift_to_origin(pts)
    mn, mx = h3d.min_max(x)
    h3d.init_space(g)
    h = h3d.Hough(mn, mx, dx)
    h.add(x)
    votes = h.votes()
    v, a, b, idx = h.get_line()
    num_x, num_b, hdx, max_x, total, gv, gidx = gold[f"{key}_hough"]
    assert (h.num_x, h.num_b, h.dx, h.max_x) == (int(num_x), int(num_b), hdx, max_x)
    assert int(votes.sum()) == int(total) and (v, idx) == (int(gv), int(gidx))
    assert G.sha(votes) == str(gold[f"{key}_votes_sha256"])
    assert np.array_equal(np.concatenate([a, b]), gold[f"{key}_line0"])
    h.close()


@pytest.mark.parametrize("key", G.CASES)
def test_loop_matches_reference_golden(h3d, gold, key):
    pts = gold[f"{key}_points"]
    g, dx, minvotes, nlines = G.case_params(gold, key)
    tr = h3d.hough3dlines(pts, g, dx, minvotes, nlines, mode=1)
    header = {f: getattr(tr, f) for f in ("rc", "n_lines", "n_input", "num_x", "num_b", "d", "dx", "max_x", "shift")}
    G.check_trace(gold, key, header, tr.iters)
