"""ls1g_cell_index_of_points_noalloc (include/ls1g.h): the entry point that allocates nothing - its device staging is a fixed
file-scope __device__ array - against the ordinary ls1g_cell_index_of_points, through the C ABI.  No counterpart in the
reference.  Collected last on purpose.

Bar: cell assignment is integer work, so the two entry points must agree bit for bit (the tolerance every cell-index test here
uses), on two calls in a row that share the scratch array - the second one shorter, so stale contents of the first would show."""
import numpy as np
import pytest

from tests.helpers import golden_cases, load
from tests.gpu_helpers import ls1g

pytestmark = pytest.mark.gpu


def _points(rng, box, n):
    """points inside the box, in the halo layer, outside it, and exactly on cell borders"""
    box = np.asarray(box, dtype=np.float64)
    p = rng.uniform(-0.3, 1.3, size=(n, 3)) * box
    k = n // 8
    p[:k] = np.round(p[:k] / 2.5) * 2.5              # multiples of the cutoff: candidates for cell borders
    p[k:2 * k] = rng.integers(0, 2, size=(k, 3)) * box  # box corners
    return p


def test_noalloc_entry_point_equals_the_allocating_one_twice_in_a_row():
    rng = np.random.default_rng(20261105)
    box = [25.3, 31.7, 20.9]
    with ls1g.Container(box, 2.5) as c:
        first = _points(rng, box, ls1g.Container.NOALLOC_MAX_POINTS)      # the maximum size
        second = _points(rng, box, 1237)                                   # shorter: the rest of the scratch keeps call one's data
        for pts in (first, second):
            ref = c.cell_index_of_points(pts)
            got = c.cell_index_of_points_noalloc(pts)
            assert got.dtype == np.uint32 and np.array_equal(got, ref)
            assert np.array_equal(got, c.cell_index_of_points(pts, host=True))
        assert len(c.cell_index_of_points_noalloc(np.zeros((0, 3)))) == 0
        # a second context on the same device shares the array (one address per device)
        with ls1g.Container([12.0, 12.0, 12.0], 3.0) as c2:
            pts = _points(rng, [12.0, 12.0, 12.0], 4096)
            assert np.array_equal(c2.cell_index_of_points_noalloc(pts), c2.cell_index_of_points(pts))
        assert np.array_equal(c.cell_index_of_points_noalloc(second), c.cell_index_of_points(second))


@pytest.mark.parametrize("name", golden_cases("cellindex_"))
def test_noalloc_entry_point_matches_the_reference_fixtures(name):
    """the adversarial border points of the reference-generated fixtures, bit-exact like test_cell_index_bit_exact"""
    d = load(name)
    with ls1g.Container(d["box"], float(d["rc"][0]), grid_mode=int(d["mode"][0])) as c:
        for _ in range(2):
            assert np.array_equal(c.cell_index_of_points_noalloc(d["points"]), d["cell_index"])


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
    with ls1g.Container([10.0, 10.0, 10.0], 2.5) as c:
        too_many = np.zeros((ls1g.Container.NOALLOC_MAX_POINTS + 1, 3))
        with pytest.raises(ls1g.Ls1gError, match="LS1G_NOALLOC_MAX_POINTS"):
            c.cell_index_of_points_noalloc(too_many)
        ok = np.full((8, 3), 5.0)
        assert np.array_equal(c.cell_index_of_points_noalloc(ok), c.cell_index_of_points(ok))      # still usable afterwards
