"""The numpy restatement of get_rewards / get_expert_actions (tests/env_post_ref.py) against golden vectors
produced by the reference's own Python (tests/golden/make_env_post_golden.py)."""
import os

import numpy as np

import env_post_ref as R

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_post.npz"))


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_shaped_reward_restatement_matches_reference_python():
    assert int(G["info_sum"][0]) > 0, "golden has no Info flag set: the weighted terms would be untested"
    for mode in ("sparse_on_goal_achieved", "weighted_combination", "reward_conditioned", "distance_to_logs"):
        r = R.shaped_reward(G["info"], G["reward"], G["expert_trajectory"], G["absolute_self_observation"], mode,
                            collision_weight=-0.75, goal_achieved_weight=1.25, off_road_weight=-0.5,
                            log_distance_weight=0.02, agent_weights=G["agent_weights"],
                            world_time_steps=G["world_time_steps"])
        want = G["reward_" + mode]
        if mode == "distance_to_logs":       # exp() differs by an ulp between libms
            assert np.allclose(r, want, rtol=1e-6, atol=1e-9), mode
        else:
            assert np.array_equal(_bits(r), _bits(want)), mode


def test_expert_actions_restatement_matches_reference_python():
    for model in range(4):
        a = R.expert_actions(G["expert_trajectory"], model)
        assert np.array_equal(_bits(a), _bits(G["expert_actions_%d" % model])), model
