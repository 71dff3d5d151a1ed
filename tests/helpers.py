"""Shared helpers of the parity tests: one parameter set applied to both sides, seeded actions,
and the comparison rules (bit-exact for flags / ids / index sets, 1e-5 relative for floats)."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REF_TESTS = "/root/reference/tests"
REF_EXAMPLES = "/root/reference/data/processed/examples"

# floating-point tolerance of BASELINE.json: 1e-5 relative. "Relative" is taken against
# max(|reference|, scale) where scale is the natural magnitude of the quantity (positions live in
# a +-150 m world, so scale = 100 m; angles and unit quantities use 1).
RTOL = 1e-5
TIE_RTOL = 1e-5        # a kNN / radius boundary difference is forgiven only inside the float tolerance
POS_SCALE = 100.0


def param_dict(**kw):
    d = dict(polylineReductionThreshold=0.1, observationRadius=50.0, rewardType=1,
             distanceToGoalThreshold=2.0, distanceToExpertThreshold=0.0, collisionBehaviour=2,
             maxNumControlledAgents=64, IgnoreNonVehicles=False, roadObservationAlgorithm=0,
             initOnlyValidAgentsAtFirstStep=True, isStaticAgentControlled=True, enableLidar=False,
             disableClassicalObs=False, dynamicsModel=1, readFromTracksToPredict=False)
    d.update(kw)
    return d


def ref_params(d):
    from oracle.refmgr import RefParams
    return RefParams(**d)


from gpudrive_lab_b200.workload import engine_params, random_actions  # noqa: E402,F401  (shared with bench.py)


def close(a, b, scale=1.0, rtol=RTOL):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.abs(a - b) <= rtol * np.maximum(np.abs(b), scale)


class Mismatch(AssertionError):
    pass


# One line per parity run, printed by conftest.pytest_terminal_summary so the exemption counters are visible in
# the tail of a `pytest -q` log (north_star: index sets bit-exact, floats 1e-5; every exemption is counted).
PARITY_REPORT: list[str] = []
EXEMPTIONS = ("knn_boundary_ties", "partner_radius_ties", "lidar_mismatch")


def report(tag, stats, allowed=None):
    """Record a parity run's counters and fail on any exemption that was not individually allowed.
    allowed: dict counter -> (max count, reason)."""
    allowed = allowed or {}
    line = f"parity[{tag}] " + " ".join(f"{k}={stats.get(k, 0)}" for k in
                                        ("knn_agents", "lidar_rays") + EXEMPTIONS)
    for k, (n, why) in allowed.items():
        line += f" (allowed {k}<={n}: {why})"
    PARITY_REPORT.append(line)
    print(line)
    for k in EXEMPTIONS:
        lim = allowed.get(k, (0, ""))[0]
        assert stats.get(k, 0) <= lim, f"{tag}: {k} = {stats.get(k, 0)} > {lim} (unexplained exemption)"


def compare_step(ref, eng, radius, lidar, tag="", stats=None):
    """ref / eng: dict name -> numpy array of the exported tensors after the same step.
    Raises Mismatch on the first violated rule. `stats` accumulates exemption counters."""
    stats = stats if stats is not None else {}
    W = ref["shape"].shape[0]

    def fail(msg):
        raise Mismatch(f"{tag}: {msg}")

    for name in ("shape", "done", "info", "steps_remaining", "controlled_state", "response_type"):
        if name not in ref:
            continue
        if not np.array_equal(ref[name], eng[name]):
            bad = np.argwhere(ref[name] != eng[name])[:5]
            fail(f"{name} differs (bit-exact rule) at {bad.tolist()} ref={ref[name][tuple(bad[0])]} eng={eng[name][tuple(bad[0])]}")

    for w in range(W):
        A = int(ref["shape"][w, 0])
        so_r, so_e = ref["self_observation"][w, :A], eng["self_observation"][w, :A]
        # collision flag and id: exact; speed / size / goal: 1e-5
        if not np.array_equal(so_r[:, 6:8], so_e[:, 6:8]):
            fail(f"w{w} self_observation collisionState/id differ: {np.argwhere(so_r[:, 6:8] != so_e[:, 6:8])[:4].tolist()}")
        ab_r, ab_e = ref["absolute_self_observation"][w, :A], eng["absolute_self_observation"][w, :A]
        live = ab_r[:, 2] < 1e30                    # not teleported to kPaddingPosition
        if not np.array_equal(ab_r[:, 2] < 1e30, ab_e[:, 2] < 1e30):
            fail(f"w{w} teleport state differs")
        ok = close(ab_e[:, 0:2], ab_r[:, 0:2], POS_SCALE).all(1) & close(ab_e[:, 3:8], ab_r[:, 3:8], 1.0).all(1)
        ok &= close(ab_e[:, 8:10], ab_r[:, 8:10], POS_SCALE).all(1) & close(ab_e[:, 10:14], ab_r[:, 10:14], 1.0).all(1)
        if not ok.all():
            i = int(np.argmin(ok))
            fail(f"w{w} agent{i} absolute obs beyond 1e-5: ref={ab_r[i]} eng={ab_e[i]}")
        ok = close(so_e[:, 0], so_r[:, 0], 1.0) & close(so_e[:, 1:4], so_r[:, 1:4], 1.0).all(1)
        # goal in the ego frame of a teleported agent is ~1e4 m away: relative to its own magnitude
        ok &= close(so_e[:, 4:6], so_r[:, 4:6], POS_SCALE).all(1)
        if not ok.all():
            i = int(np.argmin(ok))
            fail(f"w{w} agent{i} self obs beyond 1e-5: ref={so_r[i]} eng={so_e[i]}")
        if not close(eng["reward"][w, :A, 0], ref["reward"][w, :A, 0], POS_SCALE).all():
            fail(f"w{w} reward beyond 1e-5")

        # ---- partner observations: ordering + ids exact, values 1e-5
        if "partner_observations" in ref:
            pr, pe = ref["partner_observations"][w, :A], eng["partner_observations"][w, :A]
            same_id = pr[..., 8] == pe[..., 8]
            if not same_id.all():
                # a partner exactly at the observation radius may flip between row and zero() row
                bad = np.argwhere(~same_id)
                for (i, j) in bad:
                    row = pr[i, j] if pr[i, j, 8] >= 0 else pe[i, j]
                    dist = float(np.hypot(row[1], row[2]))
                    if abs(dist - radius) > TIE_RTOL * radius:
                        fail(f"w{w} agent{i} partner row {j} id differs ref={pr[i, j]} eng={pe[i, j]}")
                    stats["partner_radius_ties"] = stats.get("partner_radius_ties", 0) + 1
            m = same_id
            ok = close(pe[..., 0], pr[..., 0], 1.0) & close(pe[..., 1:3], pr[..., 1:3], POS_SCALE).all(-1)
            dh = np.abs(((pe[..., 3] - pr[..., 3]) + np.pi) % (2 * np.pi) - np.pi)
            ok &= dh <= RTOL * np.pi
            ok &= close(pe[..., 4:7], pr[..., 4:7], 1.0).all(-1) & (pe[..., 7] == pr[..., 7])
            if not (ok | ~m).all():
                i, j = np.argwhere(~(ok | ~m))[0]
                fail(f"w{w} agent{i} partner row {j} beyond 1e-5 ref={pr[i, j]} eng={pe[i, j]}")

        # ---- road map: same SET of rows per agent (row order is not a contract)
        if "agent_roadmap" in ref:
            rr, re_ = ref["agent_roadmap"][w, :A], eng["agent_roadmap"][w, :A]
            for i in range(A):
                r_rows = rr[i][rr[i][:, 6] != 0]
                e_rows = re_[i][re_[i][:, 6] != 0]
                # padding rows must be identical in content (id / mapType of fillZeros vs zero())
                r_pad = rr[i][rr[i][:, 6] == 0]
                e_pad = re_[i][re_[i][:, 6] == 0]
                if len(r_pad) and len(e_pad) and not np.array_equal(r_pad[0], e_pad[0]):
                    fail(f"w{w} agent{i} roadmap padding row differs ref={r_pad[0]} eng={e_pad[0]}")
                unmatched = _match_rows(r_rows, e_rows)
                if len(r_rows) != len(e_rows) or unmatched:
                    # tolerated only as a tie at the K boundary or at the radius
                    if not _boundary_tie(r_rows, e_rows, radius, unmatched):
                        fail(f"w{w} agent{i} roadmap set differs: ref {len(r_rows)} rows, eng {len(e_rows)} rows, "
                             f"unmatched {len(unmatched)}")
                    stats["knn_boundary_ties"] = stats.get("knn_boundary_ties", 0) + 1
                stats["knn_agents"] = stats.get("knn_agents", 0) + 1

        # ---- lidar: depth 1e-5, type exact; rays grazing a silhouette may flip (counted, bounded)
        if lidar:
            lr, le = ref["lidar"][w, :A], eng["lidar"][w, :A]
            ok = close(le[..., 0], lr[..., 0], 1.0) & (le[..., 1] == lr[..., 1])
            ok &= close(le[..., 2:4], lr[..., 2:4], 1.0).all(-1)
            stats["lidar_rays"] = stats.get("lidar_rays", 0) + ok.size
            stats["lidar_mismatch"] = stats.get("lidar_mismatch", 0) + int((~ok).sum())
            if (~ok).sum() > max(2, 2e-4 * ok.size):
                i, p, k = np.argwhere(~ok)[0]
                fail(f"w{w} agent{i} lidar plane{p} ray{k}: ref={lr[i, p, k]} eng={le[i, p, k]} "
                     f"({int((~ok).sum())}/{ok.size} rays differ)")
    return stats


def _match_rows(r_rows, e_rows):
    """Greedy bijection between two row sets on (x, y, heading, scale, type, id, mapType).
    Returns indices of engine rows with no reference partner within tolerance."""
    if len(e_rows) == 0:
        return []
    if len(r_rows) == 0:
        return list(range(len(e_rows)))
    d = np.abs(e_rows[:, None, 0:2].astype(np.float64) - r_rows[None, :, 0:2]).max(-1)
    tol = RTOL * POS_SCALE
    cand = d <= tol
    used = np.zeros(len(r_rows), bool)
    unmatched = []
    for i in range(len(e_rows)):
        js = np.nonzero(cand[i] & ~used)[0]
        hit = -1
        for j in js:
            dh = abs(((e_rows[i, 5] - r_rows[j, 5]) + np.pi) % (2 * np.pi) - np.pi)
            if dh <= RTOL * np.pi and np.array_equal(e_rows[i, 6:9], r_rows[j, 6:9]) and \
                    np.allclose(e_rows[i, 2:5], r_rows[j, 2:5], rtol=1e-6, atol=0):
                hit = j
                break
        if hit < 0:
            unmatched.append(i)
        else:
            used[hit] = True
    return unmatched


def _boundary_tie(r_rows, e_rows, radius, unmatched):
    """True if every difference between the two sets sits on the selection boundary: squared
    distance within 1e-5 relative of the largest selected one, or distance within 1e-5 of the radius
    (north_star's float tolerance: the two sides round the key of such a pair differently)."""
    if abs(len(r_rows) - len(e_rows)) > 1 or len(unmatched) > 1:
        return False
    def edge(rows, idxs):
        d2 = (rows[:, 0].astype(np.float64) ** 2 + rows[:, 1].astype(np.float64) ** 2)
        if len(d2) == 0:
            return True
        lim = d2.max()
        for i in idxs:
            if not (abs(d2[i] - lim) <= TIE_RTOL * lim or abs(np.sqrt(d2[i]) - radius) <= TIE_RTOL * radius):
                return False
        return True
    if not edge(e_rows, unmatched):
        return False
    if len(r_rows) != len(e_rows):
        # the extra row of the longer side must be at the radius
        longer = r_rows if len(r_rows) > len(e_rows) else e_rows
        d = np.sqrt(longer[:, 0].astype(np.float64) ** 2 + longer[:, 1] ** 2)
        return bool((np.abs(d - radius) <= TIE_RTOL * radius).any())
    return True


TENSORS_CMP = ["shape", "done", "info", "steps_remaining", "controlled_state", "response_type",
               "self_observation", "absolute_self_observation", "reward", "partner_observations",
               "agent_roadmap", "lidar"]


def snapshot_ref(mgr, names=TENSORS_CMP):
    return {n: mgr.tensor(n).copy() for n in names}


def snapshot_engine(sim, names=TENSORS_CMP):
    out = {}
    for n in names:
        getter = {"agent_roadmap": sim.agent_roadmap_tensor, "partner_observations": sim.partner_observations_tensor,
                  "self_observation": sim.self_observation_tensor,
                  "absolute_self_observation": sim.absolute_self_observation_tensor,
                  "controlled_state": sim.controlled_state_tensor, "response_type": sim.response_type_tensor,
                  "steps_remaining": sim.steps_remaining_tensor, "shape": sim.shape_tensor,
                  "done": sim.done_tensor, "info": sim.info_tensor, "reward": sim.reward_tensor,
                  "lidar": sim.lidar_tensor, "expert_trajectory": sim.expert_trajectory_tensor,
                  "map_observation": sim.map_observation_tensor, "world_means": sim.world_means_tensor,
                  "metadata": sim.metadata_tensor, "knn_index": sim.knn_index_tensor,
                  "action": sim.action_tensor}[n]
        out[n] = getter().to_torch().cpu().numpy()
    return out


class EngineAdapter:
    """The CUDA engine (SimManager stand-in over the C ABI) behind the surface tests/kat.py drives."""

    _GETTERS = {"action": "action_tensor", "reward": "reward_tensor", "done": "done_tensor", "info": "info_tensor",
                "self_observation": "self_observation_tensor", "partner_observations": "partner_observations_tensor",
                "agent_roadmap": "agent_roadmap_tensor", "lidar": "lidar_tensor", "shape": "shape_tensor",
                "absolute_self_observation": "absolute_self_observation_tensor",
                "steps_remaining": "steps_remaining_tensor", "controlled_state": "controlled_state_tensor",
                "response_type": "response_type_tensor", "knn_index": "knn_index_tensor",
                "map_observation": "map_observation_tensor", "expert_trajectory": "expert_trajectory_tensor",
                "bev_observation": "bev_observation_tensor"}

    def __init__(self, scenes, pd, S=128, enable_bev=False):
        from gpudrive_lab_b200 import madrona_gpudrive as mg
        self.sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(pd), max_agents=S,
                                 enable_bev=enable_bev)
        self._act = self.sim.action_tensor().to_torch()

    def tensor(self, name):
        return getattr(self.sim, self._GETTERS[name])().to_torch().cpu().numpy()

    def set_actions(self, a):
        import torch
        self._act.copy_(torch.from_numpy(np.ascontiguousarray(a)))
        torch.cuda.synchronize()

    def step(self):
        self.sim.step()

    def reset(self, worlds):
        self.sim.reset(list(worlds))


def ragged_scenes(out_dir):
    """Edge-case scenarios cut from a synthetic one: one agent with three short polylines; two agents with only
    a stop sign and one road edge; three agents and no road at all. Returns their paths."""
    import json
    from gpudrive_lab_b200 import synth
    base = synth.make_scenario(3, num_agents=6)
    none = {"sdc_track_index": 0, "objects_of_interest": [], "tracks_to_predict": []}
    a = dict(base, objects=base["objects"][:1], metadata=none,
             roads=[r for r in base["roads"] if r["type"] in ("lane", "road_edge")][:3])
    b = dict(base, objects=base["objects"][:2],
             roads=[r for r in base["roads"] if r["type"] == "stop_sign"][:1] +
                   [r for r in base["roads"] if r["type"] == "road_edge"][:1])
    c = dict(base, objects=base["objects"][:3], roads=[])
    paths = []
    for name, s in (("one_agent", a), ("two_roads", b), ("no_roads", c)):
        p = os.path.join(str(out_dir), f"ragged_{name}.json")
        with open(p, "w") as fh:
            json.dump(s, fh)
        paths.append(p)
    return paths
