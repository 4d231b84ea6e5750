"""oracle/init_oracle.py against the golden vectors made with the reference's own depth_to_pts3d / RGB2SH /
inverse_sigmoid (tests/golden/make_init_golden.py): the mask and the compaction order exactly, floats to fp32 rounding
of a differently ordered 3-term dot product."""
import glob
import os

import numpy as np
import pytest

import _harness as Hn
from oracle import init_oracle as io

GOLDEN = sorted(glob.glob(os.path.join(Hn.ROOT, "tests", "golden", "init_*.npz")))


def check_against_golden(got, z, rtol=2e-6):
    assert np.array_equal(np.asarray(got["aggregated_mask"]).astype(bool), z["out_aggregated_mask"])
    assert np.array_equal(np.asarray(got["conf_index"]), np.nonzero(z["out_aggregated_mask"])[0])
    np.testing.assert_allclose(got["xyz"], z["out_xyz"], rtol=rtol, atol=2e-6)
    np.testing.assert_allclose(got["features_dc"], z["out_features_dc"], rtol=rtol, atol=1e-6)
    assert np.array_equal(got["features_rest"], z["out_features_rest"])
    assert np.array_equal(got["rotation"], z["out_rotation"])
    np.testing.assert_allclose(got["opacity"], z["out_opacity"], rtol=rtol)  # inf == inf for a single view
    np.testing.assert_allclose(got["conf_static"], z["out_conf_static"], rtol=0, atol=0)
    np.testing.assert_allclose(got["dist2"], z["out_dist2"], rtol=1e-4, atol=1e-12)  # xyz differs in the last bit
    np.testing.assert_allclose(got["scaling"], z["out_scaling"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_against_reference_golden(path):
    z = np.load(path)
    got = io.create_from_pointmaps(z["in_K"], z["in_poses"], z["in_depth"], z["in_confidence"], z["in_dyna_avg"], z["in_rgb"],
                                   float(z["in_conf_thre"]))
    assert len(GOLDEN) >= 3
    check_against_golden(got, z)
