"""Golden vectors for the device-side policy observation (gd_policy_obs): the reference's own Python
post-processing (gpudrive/datatypes/observation.py, roadgraph.py, as assembled by GPUDriveTorchEnv.get_obs with
its defaults) applied on the CPU to tensors of the unmodified reference Manager. Run in the build container:

    python tests/golden/make_policy_obs_golden.py

Writes policy_obs.npz: the three input tensors (cropped to the live agent range) and the expected output."""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference")

import gpudrive_lab_b200  # noqa: E402

gpudrive_lab_b200.install()            # the reference's datatypes `import madrona_gpudrive` (for vehicleScale / EntityType)
from gpudrive.datatypes.observation import LocalEgoState, PartnerObs  # noqa: E402
from gpudrive.datatypes.roadgraph import LocalRoadGraphPoints  # noqa: E402

from golden_util import unpack  # noqa: E402
from helpers import param_dict, random_actions, ref_params  # noqa: E402


class _T:      # what from_tensor expects: an object with to_torch()
    def __init__(self, a):
        self.a = a

    def to_torch(self):
        return torch.from_numpy(np.array(self.a, copy=True))


def reference_get_obs(self_obs, partner, roadmap, norm):
    """GPUDriveTorchEnv._get_ego_state / _get_partner_obs / _get_road_map_obs / get_obs with mask=None."""
    ego = LocalEgoState.from_tensor(_T(self_obs), device="cpu")
    if norm:
        ego.normalize()
    ego_t = torch.stack([ego.speed, ego.vehicle_length, ego.vehicle_width, ego.rel_goal_x, ego.rel_goal_y,
                         ego.is_collided], dim=-1)
    po = PartnerObs.from_tensor(_T(partner), device="cpu")
    if norm:
        po.normalize()
    po_t = torch.concat([po.speed, po.rel_pos_x, po.rel_pos_y, po.orientation, po.vehicle_length,
                         po.vehicle_width], dim=-1).flatten(start_dim=2)
    rg = LocalRoadGraphPoints.from_tensor(_T(roadmap), device="cpu")
    rg.one_hot_encode_road_point_types()
    if norm:
        rg.normalize()
    rg_t = torch.cat([rg.x.unsqueeze(-1), rg.y.unsqueeze(-1), rg.segment_length.unsqueeze(-1),
                      rg.segment_width.unsqueeze(-1), rg.segment_height.unsqueeze(-1), rg.orientation.unsqueeze(-1),
                      rg.type], dim=-1).flatten(start_dim=2)
    return torch.cat((ego_t, po_t, rg_t), dim=-1).numpy()


def main():
    from oracle.refmgr import RefManager
    tmp = tempfile.mkdtemp()
    scenes = unpack(["synth_a", "synth_b"], tmp)
    pd = param_dict(dynamicsModel=2, maxNumControlledAgents=12)
    ref = RefManager(scenes, ref_params(pd))
    for t in range(7):
        ref.tensor("action")[...] = random_actions(t, 2, 128, 2, seed=17)
        ref.step()
    A = 16     # live agents (8 / 12) plus a few padding slots
    so = ref.tensor("self_observation")[:, :A].copy()
    po = ref.tensor("partner_observations")[:, :A].copy()
    rm = ref.tensor("agent_roadmap")[:, :A].copy()
    out = {"self_observation": so, "partner_observations": po, "agent_roadmap": rm,
           "obs_norm": reference_get_obs(so, po, rm, True), "obs_raw": reference_get_obs(so, po, rm, False)}
    np.savez_compressed(os.path.join(HERE, "policy_obs.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
