"""The numpy restatement of the reference's Python observation post-processing is bit-identical to the golden
output produced by the reference's own classes (tests/golden/make_policy_obs_golden.py)."""
import os

import numpy as np

from golden_util import GOLDEN
from policy_obs_ref import policy_obs


def test_restatement_matches_reference_python():
    g = np.load(os.path.join(GOLDEN, "policy_obs.npz"))
    for norm, key in ((True, "obs_norm"), (False, "obs_raw")):
        mine = policy_obs(g["self_observation"], g["partner_observations"], g["agent_roadmap"], norm)
        assert mine.shape == g[key].shape == (2, 16, 6 + 127 * 6 + 200 * 13)
        assert np.array_equal(mine.view(np.uint32), g[key].view(np.uint32)), key
