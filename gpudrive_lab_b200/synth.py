"""Seeded synthetic Waymo-shaped scenarios in the reference's JSON schema
(src/json_serialization.hpp:18-414; data_utils' output format): the workload of the benchmark
and of the parity tests, because WOMD itself is not available offline.

World: `extent` x `extent` metres, a 2 x 2 grid of intersections. Every road is a gently curved
polyline (sum of two sines) sampled every `spacing` metres carrying: two lanes, a centre road
line, two road edges. Plus crosswalks, speed bumps and stop signs. Vehicles follow lanes with a
constant acceleration at 10 Hz for 91 samples; a few samples are marked invalid WOMD-style
(valid=false, position -10000). Everything derives from numpy's PCG64 seeded with
0x5EED0000 + index, so oracle and engine read the same bytes on any machine.
"""
from __future__ import annotations

import json
import math
import os

import numpy as np

SEED_BASE = 0x5EED0000


def _polyline(rng, start, direction, length, spacing, amp, wavelength):
    """Curved centre line: straight run along `direction` + lateral sine wobble."""
    n = int(length / spacing) + 1
    s = np.arange(n) * spacing
    ph1, ph2 = rng.uniform(0, 2 * math.pi, 2)
    lat = amp * np.sin(2 * math.pi * s / wavelength + ph1) + 0.35 * amp * np.sin(2 * math.pi * s / (0.37 * wavelength) + ph2)
    d = np.asarray(direction, float)
    nrm = np.array([-d[1], d[0]])
    pts = np.asarray(start, float)[None, :] + s[:, None] * d[None, :] + lat[:, None] * nrm[None, :]
    return pts


def _offset(pts, off):
    t = np.gradient(pts, axis=0)
    t /= np.linalg.norm(t, axis=1, keepdims=True) + 1e-12
    nrm = np.stack([-t[:, 1], t[:, 0]], axis=1)
    return pts + off * nrm


def _geom(pts):
    return [{"x": float(p[0]), "y": float(p[1])} for p in pts]


def make_scenario(index: int, num_agents: int = 64, extent: float = 300.0, spacing: float = 0.5,
                  amp: float = 4.0, wavelength: float = 90.0, origin=(2000.0, -1500.0),
                  invalid_frac: float = 0.05, stress: bool = False, edge_spacing: float | None = None) -> dict:
    """One scenario dict. `stress=True` packs the agents into a 120 m double intersection
    (collision-stress config); `edge_spacing` overrides the sampling of road edges only."""
    rng = np.random.Generator(np.random.PCG64(SEED_BASE + index))
    half = extent / 2
    ox, oy = origin
    roads = []
    lanes = []          # (points, road id) usable for placing vehicles
    rid = 0
    centres = [-extent / 4, extent / 4]
    for axis in (0, 1):
        for c in centres:
            if axis == 0:   # runs along x at y = c
                start, d = (ox - half, oy + c), (1.0, 0.0)
            else:
                start, d = (ox + c, oy - half), (0.0, 1.0)
            centre = _polyline(rng, start, d, extent, spacing, amp, wavelength)
            for off, typ, me in ((-1.75, "lane", 2), (1.75, "lane", 2), (0.0, "road_line", 12)):
                pts = _offset(centre, off)
                if off > 0:
                    pts = pts[::-1].copy()     # opposite driving direction
                roads.append({"geometry": _geom(pts), "type": typ, "map_element_id": me, "id": rid})
                if typ == "lane":
                    lanes.append(pts)
                rid += 1
            for off in (-5.5, 5.5):
                src = centre
                if edge_spacing is not None and edge_spacing != spacing:
                    src = _polyline(np.random.Generator(np.random.PCG64(SEED_BASE + index + 7919 * rid)), start, d,
                                    extent, edge_spacing, amp, wavelength)
                pts = _offset(src, off)
                # break the edge at the crossing roads so vehicles can pass through intersections
                keep = np.ones(len(pts), bool)
                other = pts[:, 1 - axis] - (oy if axis == 0 else ox)
                along = pts[:, axis] - (ox if axis == 0 else oy)
                for cc in centres:
                    keep &= np.abs(along - cc) > 9.0
                del other
                seg_start = 0
                for i in range(1, len(pts) + 1):
                    if i == len(pts) or keep[i] != keep[seg_start]:
                        if keep[seg_start] and i - seg_start >= 2:
                            roads.append({"geometry": _geom(pts[seg_start:i]), "type": "road_edge",
                                          "map_element_id": 15, "id": rid})
                            rid += 1
                        seg_start = i
    # crosswalks, speed bumps (4-point boxes) and stop signs (1 point) near the intersections
    for cx in centres:
        for cy in centres:
            x0, y0 = ox + cx, oy + cy
            w, l = 3.0, 9.0
            roads.append({"geometry": _geom(np.array([[x0 + 10, y0 - l / 2], [x0 + 10 + w, y0 - l / 2],
                                                      [x0 + 10 + w, y0 + l / 2], [x0 + 10, y0 + l / 2]])),
                          "type": "crosswalk", "map_element_id": 18, "id": rid}); rid += 1
            roads.append({"geometry": _geom(np.array([[x0 + 8.0, y0 - 8.0]])), "type": "stop_sign",
                          "map_element_id": 17, "id": rid}); rid += 1
    for k in range(2):
        x0, y0 = ox + centres[k] + 40.0, oy + centres[0] + float(rng.uniform(-1, 1))
        roads.append({"geometry": _geom(np.array([[x0, y0 - 4], [x0 + 1.2, y0 - 4], [x0 + 1.2, y0 + 4], [x0, y0 + 4]])),
                      "type": "speed_bump", "map_element_id": 19, "id": rid}); rid += 1

    # vehicles
    objects = []
    T = 91
    dt = 0.1
    per_lane = int(math.ceil(num_agents / len(lanes)))
    a_idx = 0
    for li, pts in enumerate(lanes):
        seglen = np.linalg.norm(np.diff(pts, axis=0), axis=1)
        cum = np.concatenate([[0.0], np.cumsum(seglen)])
        total = cum[-1]
        if stress:
            lo, hi = 0.30 * total, 0.70 * total
        else:
            lo, hi = 0.03 * total, 0.80 * total
        starts = np.sort(rng.uniform(lo, hi, per_lane))
        for s0 in starts:
            if a_idx >= num_agents:
                break
            v0 = float(rng.uniform(0.0, 15.0))
            acc = float(rng.uniform(-1.0, 1.5))
            t = np.arange(T) * dt
            v = np.clip(v0 + acc * t, 0.0, 20.0)
            s = s0 + np.concatenate([[0.0], np.cumsum(0.5 * (v[1:] + v[:-1]) * dt)])
            s = np.minimum(s, total - 1e-3)
            x = np.interp(s, cum, pts[:, 0])
            y = np.interp(s, cum, pts[:, 1])
            eps = 0.25
            xa = np.interp(np.minimum(s + eps, total), cum, pts[:, 0])
            ya = np.interp(np.minimum(s + eps, total), cum, pts[:, 1])
            xb = np.interp(np.maximum(s - eps, 0), cum, pts[:, 0])
            yb = np.interp(np.maximum(s - eps, 0), cum, pts[:, 1])
            heading = np.arctan2(ya - yb, xa - xb)
            vx, vy = v * np.cos(heading), v * np.sin(heading)
            valid = np.ones(T, bool)
            if rng.uniform() < 0.6 and invalid_frac > 0:
                span = max(1, int(round(invalid_frac * T / 0.6)))
                b = int(rng.integers(5, T - span - 1))
                valid[b:b + span] = False
            last = int(np.nonzero(valid)[0][-1])
            pos = [{"x": float(x[i]) if valid[i] else -10000.0, "y": float(y[i]) if valid[i] else -10000.0}
                   for i in range(T)]
            vel = [{"x": float(vx[i]) if valid[i] else 0.0, "y": float(vy[i]) if valid[i] else 0.0} for i in range(T)]
            objects.append({
                "position": pos, "width": float(rng.uniform(1.8, 2.2)), "length": float(rng.uniform(4.0, 5.5)),
                "height": 1.5, "heading": [float(h) for h in heading], "velocity": vel,
                "valid": [bool(b_) for b_ in valid],
                "goalPosition": {"x": float(x[last]), "y": float(y[last])},
                "type": "vehicle", "id": 100 + a_idx, "mark_as_expert": False,
            })
            a_idx += 1
    name = f"synthetic_{index:06d}.json"
    return {
        "name": name, "scenario_id": f"{(SEED_BASE + index) & 0xFFFFFFFF:08x}synth",
        "objects": objects, "roads": roads, "tl_states": {},
        "metadata": {"sdc_track_index": 0, "objects_of_interest": [],
                     "tracks_to_predict": [{"track_index": 1, "difficulty": 0}] if len(objects) > 1 else []},
    }


def _write_one(args):
    p, i, kw = args
    tmp = p + f".tmp{os.getpid()}"
    with open(tmp, "w") as fh:
        json.dump(make_scenario(i, **kw), fh, separators=(",", ":"))
    os.replace(tmp, p)
    return p


def write_scenarios(out_dir: str, count: int, start_index: int = 0, **kw) -> list[str]:
    """Write `count` scenarios (indices start_index..) and return their paths. Existing files are kept; many
    missing files are generated on all host cores (0.4 s each)."""
    os.makedirs(out_dir, exist_ok=True)
    paths, todo = [], []
    tag = "_".join(f"{k}{v}" for k, v in sorted(kw.items()))
    for i in range(start_index, start_index + count):
        p = os.path.join(out_dir, f"synthetic_{i:06d}{('_' + tag) if tag else ''}.json")
        if not os.path.exists(p):
            todo.append((p, i, kw))
        paths.append(p)
    if len(todo) > 16 and (os.cpu_count() or 1) > 1:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(min(len(os.sched_getaffinity(0)), 32)) as pool:
            pool.map(_write_one, todo, chunksize=4)
    else:
        for t in todo:
            _write_one(t)
    return paths
