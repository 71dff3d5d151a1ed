"""numpy restatement of the reference's Python observation post-processing (the checker for gd_policy_obs):
gpudrive/env/env_torch.py:756-891,1172-1216, gpudrive/datatypes/observation.py:29-92,170-258,
gpudrive/datatypes/roadgraph.py:281-362, gpudrive/env/constants.py. float32 throughout, torch's op order."""
import numpy as np

F = np.float32


def _mm(v):          # normalize_min_max(v, -1000, 1000) = 2 * ((v - min) / (max - min)) - 1
    return F(2) * ((v - F(-1000)) / F(2000)) - F(1)


def policy_obs(self_obs, partner, roadmap, normalize=True):
    so, po, rm = (np.asarray(x, F) for x in (self_obs, partner, roadmap))
    s7 = F(0.7)
    ego = [so[..., 0], so[..., 1] * s7, so[..., 2] * s7, so[..., 4], so[..., 5], so[..., 6]]
    par = [po[..., 0], po[..., 1], po[..., 2], po[..., 3], po[..., 4] * s7, po[..., 5] * s7]
    road = [rm[..., c] for c in range(6)]
    if normalize:
        ego = [ego[0] / F(100), ego[1] / F(30), ego[2] / F(15), _mm(ego[3]), _mm(ego[4]), ego[5]]
        two_pi = F(2 * np.pi)
        par = [par[0] / F(100), _mm(par[1]), _mm(par[2]), par[3] / two_pi, par[4] / F(30), par[5] / F(15)]
        road = [_mm(road[0]), _mm(road[1]), road[2] / F(100), road[3] / F(100), road[4] / F(100), road[5] / two_pi]
    onehot = (rm[..., 6].astype(np.int64)[..., None] == np.arange(7)).astype(F)
    lead = so.shape[:-1]
    e = np.stack(ego, -1)
    p = np.stack(par, -1).reshape(*lead, -1)
    r = np.concatenate([np.stack(road, -1), onehot], -1).reshape(*lead, -1)
    return np.concatenate([e, p, r], -1).astype(F)
