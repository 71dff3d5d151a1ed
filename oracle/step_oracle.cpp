// TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's per-step simulation for ONE world:
// the checker that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg run beside the CUDA
// engine. Nothing in gpudrive_lab_b200/ includes, links or calls this file.
//
// What it restates (reference @ /root/reference, file:line given at every function):
//   movementSystem + the four dynamics models   src/sim.cpp:294-383, src/dynamics.hpp:11-115,186-194
//   collisionDetectionSystem + 2-D OBB SAT      src/sim.cpp:628-747, src/obb.hpp:12-92, src/sim.hpp:88-102
//   broadphase pair rules                       external/madrona/src/physics/broadphase.cpp:438-483,928-991
//   reward / stepTracker / done                 src/sim.cpp:560-626
//   self / absolute / partner observations      src/sim.cpp:168-240,769-783
//   agent road map (kNN heap and linear)        src/sim.cpp:242-280, src/knn.hpp:15-158
//   lidar                                       src/sim.cpp:394-460, broadphase.cpp:656-888, math.inl:1575-1601
//   bird's-eye-view grid                        src/sim.cpp:462-555, src/rasterizer.hpp:12-78
//   reset                                       src/level_gen.cpp:467-521, src/level_gen.hpp:16-38
// Scene construction (JSON -> entities, rows a16/a17 of SURVEY.md section 8) is NOT restated here: the
// oracle takes the entity tables as arrays. Those tables are pinned separately against the reference
// (tests/test_cpu_loader.py and tests/golden/).
//
// Style: straight scalar code with general 3-D vectors and full quaternions, the reference's own data
// structures (a binary max-heap for the K nearest, brute-force loops instead of the BVH), so that it
// shares no shortcut with the CUDA kernels (z-only quaternions, radix select, culling, work queues).
// Compile with -ffp-contract=off: the reference oracle build (oracle/ref/Makefile) does the same.
//
// Parity status: PINNED. tests/test_cpu_oracle.py checks every tensor of this restatement against the
// unmodified reference Manager (oracle/_ref) and against tests/golden/*.npz generated from it.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

constexpr int K_ROADMAP = 200;        // consts::kMaxAgentMapObservationsCount, src/consts.hpp:13
constexpr int EPISODE_LEN = 91;       // consts::episodeLen, src/consts.hpp:34
constexpr int TRAJ_LEN = 91;          // consts::kTrajectoryLength, src/consts.hpp:57
constexpr int N_LIDAR = 50;           // consts::numLidarSamples, src/consts.hpp:37
constexpr float PI_F = 3.14159265358979323846f;
constexpr float PI_M2 = PI_F * 2.f;
constexpr float LIDAR_ANGLE = PI_F / 3;     // consts::lidarAngle
constexpr float LIDAR_DIST = 200.f;         // consts::lidarDistance
constexpr float LIDAR_OFFSETS[3] = {0.5f, 0.1f, -0.1f};   // car, road edge, road line planes
constexpr int TRAJ_SIZE = 1456;
constexpr int T_POS = 0, T_VEL = 182, T_HEAD = 364, T_VALID = 455;

enum { ET_None, ET_RoadEdge, ET_RoadLine, ET_RoadLane, ET_CrossWalk, ET_SpeedBump, ET_StopSign, ET_Vehicle,
       ET_Pedestrian, ET_Cyclist, ET_Padding };
enum { RESP_Dynamic = 0, RESP_Kinematic = 1, RESP_Static = 2 };

struct V3 { float x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length(V3 a) { return sqrtf(dot(a, a)); }
inline float length2_xy(float x, float y) { return x * x + y * y; }

struct Quat { float w, x, y, z; };
// math.inl:659-691
inline Quat inv(Quat q) { return {q.w, -q.x, -q.y, -q.z}; }
inline V3 rotate(Quat q, V3 v) {
    V3 pure{q.x, q.y, q.z};
    V3 c1 = cross(pure, v);
    V3 c2 = cross(pure, c1);
    return v + 2.f * ((c1 * q.w) + c2);
}
inline Quat angle_axis_up(float angle) {
    float ch = cosf(angle / 2.f), sh = sinf(angle / 2.f);
    return {ch, 0.f * sh, 0.f * sh, 1.f * sh};
}
// math.inl:837-845
inline Quat mul(Quat a, Quat b) {
    return {(a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z), (a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y),
            (a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x), (a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w)};
}
// src/utils.hpp:11-25. The reference calls unqualified atan2 / fmod on floats: with <cmath> and no
// using-directive those are the C double functions, so the float arguments are promoted.
inline float quat_to_yaw(Quat q) {
    return (float)::atan2((double)(2.0f * (q.w * q.z + q.x * q.y)), (double)(1.0f - 2.0f * (q.y * q.y + q.z * q.z)));
}
inline float normalize_angle(float angle) {
    const float ret = (float)::fmod((double)angle, (double)PI_M2);
    return ret > PI_F ? ret - PI_M2 : (ret < -PI_F ? ret + PI_M2 : ret);
}
inline float angle_add(float a, float b) { return normalize_angle(a + b); }

struct Params {     // Parameters, src/init.hpp:111-127 (same field order as GdParams)
    float polylineReductionThreshold, observationRadius;
    int32_t rewardType;
    float distanceToGoalThreshold, distanceToExpertThreshold;
    int32_t collisionBehaviour;
    uint32_t maxNumControlledAgents;
    int32_t ignoreNonVehicles, roadObservationAlgorithm, initOnlyValidAgentsAtFirstStep, isStaticAgentControlled,
        enableLidar, disableClassicalObs, dynamicsModel, readFromTracksToPredict;
};

struct Road {
    V3 pos; Quat rot; V3 scale; float yaw; int type; float id; float mapType;
};
struct Agent {
    // static
    V3 vsize, scale; float gx, gy; int type; float id; int resp0; int controlled;
    const float *traj;
    // dynamic
    V3 pos; Quat rot; V3 vel; V3 ang;
    int collided;
};

struct World {
    Params prm;
    int S, A, R;
    std::vector<Road> roads;
    std::vector<Agent> agents;
    std::vector<float> traj;
    // exports, reference layouts for one world
    std::vector<float> action, reward, self_obs, partner, roadmap, lidar, abs_obs, bev;
    bool want_bev = false;
    std::vector<int32_t> done, steps, info, controlled, response, knn_index;
};

// ---------------------------------------------------------------- dynamics, src/dynamics.hpp
// forwardBicycleModel :52-81. `0.5 * a` is a double expression (double literal), the rest float.
void forward_bicycle(float *act, Agent &g) {
    const float maxSpeed = INFINITY; (void)maxSpeed;
    const float dt = 0.1f;
    float acc = fmaxf(-6.0f, fminf(act[0], 6.0f));
    float steer = fmaxf(-3.0f, fminf(act[1], 3.0f));
    act[0] = acc; act[1] = steer;
    float yaw = quat_to_yaw(g.rot);
    float speed = length(g.vel);
    g.pos.x = (float)((double)(g.pos.x + g.vel.x * dt) + 0.5 * acc * cosf(yaw) * dt * dt);
    g.pos.y = (float)((double)(g.pos.y + g.vel.y * dt) + 0.5 * acc * sinf(yaw) * dt * dt);
    float delta_yaw = (float)(steer * ((double)(speed * dt) + 0.5 * acc * dt * dt));
    float new_yaw = angle_add(yaw, delta_yaw);
    float new_speed = speed + acc * dt;
    g.vel = {new_speed * cosf(new_yaw), new_speed * sinf(new_yaw), 0.f};
    g.ang = {0.f, 0.f, delta_yaw / dt};
    g.rot = angle_axis_up(new_yaw);
}
// forwardDeltaModel :83-115
void forward_delta(const float *act, Agent &g) {
    const float dt = 0.1f;
    float yaw = quat_to_yaw(g.rot);
    float c = cosf(yaw), s = sinf(yaw);
    float dx = act[0] * c - act[1] * s;
    float dy = act[0] * s + act[1] * c;
    g.pos.x += dx; g.pos.y += dy;
    g.vel = {dx / dt, dy / dt, 0.f};
    g.ang = {0.f, 0.f, act[2] / dt};
    g.rot = angle_axis_up(angle_add(yaw, act[2]));
}
// forwardKinematics :11-50 (Nocturne bicycle, "Classic")
void forward_classic(const float *act, Agent &g) {
    const float dt = 0.1f;
    float speed = length(g.vel);
    float yaw = quat_to_yaw(g.rot);
    float v = speed + 0.5f * act[0] * dt;
    float tanDelta = tanf(act[1]);
    float beta = atanf(0.5f * tanDelta);
    float dxx = v * cosf(yaw + beta), dyy = v * sinf(yaw + beta);
    float w = v * cosf(beta) * tanDelta / g.vsize.x;
    float new_yaw = angle_add(yaw, w * dt);
    float new_speed = speed + act[0] * dt;
    g.pos.x += dxx * dt; g.pos.y += dyy * dt; g.pos.z = 1.f;
    g.rot = angle_axis_up(new_yaw);
    g.vel = {new_speed * cosf(new_yaw), new_speed * sinf(new_yaw), 0.f};
    g.ang = {0.f, 0.f, w};
}
// forwardStateModel :186-194
void forward_state(const float *act, Agent &g) {
    g.pos = {act[0], act[1], act[2]};
    g.rot = angle_axis_up(act[3]);
    g.vel = {act[4], act[5], act[6]};
    g.ang = {act[7], act[8], act[9]};
}

inline int current_step(int t) {    // getCurrentStep, src/sim.cpp (episodeLen - t), clamped to the row
    int idx = EPISODE_LEN - t;
    return idx < 0 ? 0 : (idx > TRAJ_LEN ? TRAJ_LEN : idx);
}

// ---------------------------------------------------------------- movementSystem, src/sim.cpp:294-383
void movement(World &w) {
    for (int a = 0; a < w.A; a++) {
        Agent &g = w.agents[a];
        if (g.collided) {
            switch (w.prm.collisionBehaviour) {
            case 0: w.done[a] = 1; g.vel = {0, 0, 0}; g.ang = {0, 0, 0}; break;
            case 1: w.done[a] = 1; g.pos = {-11000.f, -11000.f, FLT_MAX}; g.vel = {0, 0, 0}; g.ang = {0, 0, 0}; break;
            default:
                g.collided = 0;
                w.info[a * 5 + 0] = w.info[a * 5 + 1] = w.info[a * 5 + 2] = 0;
            }
        }
        if (w.response[a] == RESP_Static) continue;
        if (w.done[a]) {
            g.pos = {-11000.f, -11000.f, FLT_MAX}; g.vel = {0, 0, 0}; g.ang = {0, 0, 0};
            continue;
        }
        float *act = &w.action[(size_t)a * 10];
        if (w.controlled[a]) {
            switch (w.prm.dynamicsModel) {
            case 1: forward_bicycle(act, g); break;
            case 2: forward_delta(act, g); break;
            case 0: forward_classic(act, g); break;
            default: forward_state(act, g);
            }
        } else {
            const int i = current_step(w.steps[a]);
            g.pos = {g.traj[T_POS + 2 * i], g.traj[T_POS + 2 * i + 1], 1.f};
            g.vel = {g.traj[T_VEL + 2 * i], g.traj[T_VEL + 2 * i + 1], 0.f};
            g.ang = {0, 0, 0};
            g.rot = angle_axis_up(g.traj[T_HEAD + i]);
        }
    }
}

// ---------------------------------------------------------------- 2-D OBB, src/obb.hpp:12-92
struct Obb { float cx[4], cy[4], ax[2], ay[2], org[2]; };
Obb obb_from(V3 pos, Quat rot, V3 scale) {
    float theta = quat_to_yaw(rot);
    float Xx = cosf(theta), Xy = sinf(theta), Yx = -sinf(theta), Yy = cosf(theta);
    Xx *= scale.x; Xy *= scale.x; Yx *= scale.y; Yy *= scale.y;
    Obb o;
    o.cx[0] = pos.x - Xx - Yx; o.cy[0] = pos.y - Xy - Yy;
    o.cx[1] = pos.x + Xx - Yx; o.cy[1] = pos.y + Xy - Yy;
    o.cx[2] = pos.x + Xx + Yx; o.cy[2] = pos.y + Xy + Yy;
    o.cx[3] = pos.x - Xx + Yx; o.cy[3] = pos.y - Xy + Yy;
    float ex[2] = {o.cx[1] - o.cx[0], o.cx[3] - o.cx[0]}, ey[2] = {o.cy[1] - o.cy[0], o.cy[3] - o.cy[0]};
    for (int a = 0; a < 2; a++) {
        // Vector2::operator/=(float) multiplies by the reciprocal (madrona math.inl)
        float inv = 1.f / length2_xy(ex[a], ey[a]);
        o.ax[a] = ex[a] * inv; o.ay[a] = ey[a] * inv;
        o.org[a] = o.cx[0] * o.ax[a] + o.cy[0] * o.ay[a];
    }
    return o;
}
bool obb_overlaps(const Obb &self, const Obb &other) {
    for (int a = 0; a < 2; a++) {
        float t = other.cx[0] * self.ax[a] + other.cy[0] * self.ay[a];
        float tMin = t, tMax = t;
        for (int c = 1; c < 4; c++) {
            t = other.cx[c] * self.ax[a] + other.cy[c] * self.ay[a];
            if (t < tMin) tMin = t; else if (t > tMax) tMax = t;
        }
        if ((tMin > 1 + self.org[a]) || (tMax < self.org[a])) return false;
    }
    return true;
}
bool obb_collided(const Obb &a, const Obb &b) { return obb_overlaps(a, b) && obb_overlaps(b, a); }

// Sim::collisionPairs, src/sim.hpp:88-102 (type pairs whose contact is ignored)
bool pair_ignored(int a, int b) {
    auto one = [](int p, int q) {
        if (p == ET_Pedestrian || p == ET_Cyclist)
            return q == ET_RoadEdge || q == ET_RoadLine || q == ET_RoadLane || q == ET_CrossWalk || q == ET_SpeedBump;
        if (p == ET_Vehicle) return q == ET_CrossWalk || q == ET_SpeedBump || q == ET_RoadLine || q == ET_RoadLane;
        return false;
    };
    return one(a, b) || one(b, a);
}
void mark_info(int32_t *info, int other) {    // src/sim.cpp:713-724
    if (other > ET_None && other <= ET_StopSign) info[0] = 1;
    else if (other == ET_Vehicle) info[1] = 1;
    else if (other <= ET_Cyclist) info[2] = 1;
}

// Broadphase leaf z-range. Leaf AABB = mesh AABB (cube: z in [-1,1]; agent hull: z in [0,2]) scaled and
// translated, then expandAABBWithMotion (broadphase.cpp:438-462): per axis delta = 0.08 v -/+ 0.16.
// Only z is restated: the x/y ranges always contain the SAT box, so they cannot reject a SAT hit.
void leaf_z(float z, float d2, float mesh_lo, float mesh_hi, float vz, float &lo, float &hi) {
    lo = z + d2 * mesh_lo; hi = z + d2 * mesh_hi;
    float pos_delta = (2.f * 0.04f) * vz;
    float min_delta = pos_delta - 0.16f, max_delta = pos_delta + 0.16f;
    if (min_delta < 0.f) lo += min_delta;
    if (max_delta > 0.f) hi += max_delta;
}

// collisionDetectionSystem, src/sim.cpp:628-747, over every pair the broadphase can report
// (broadphase.cpp:928-991: not Static x Static, leaf AABBs overlap strictly, math.inl:1506-1514).
void collisions(World &w) {
    const int A = w.A;
    std::vector<Obb> box(A);
    std::vector<int> skip(A);
    std::vector<float> zlo(A), zhi(A);
    for (int a = 0; a < A; a++) {
        Agent &g = w.agents[a];
        box[a] = obb_from(g.pos, g.rot, g.scale);
        leaf_z(g.pos.z, g.scale.z, 0.f, 2.f, g.vel.z, zlo[a], zhi[a]);
        // isInvalidExpertOrDone :631-662, hasCollided as it stood before this pass
        if (!w.controlled[a]) skip[a] = !(g.traj[T_VALID + current_step(w.steps[a])] != 0.f);
        else skip[a] = w.done[a] && !g.collided;
    }
    std::vector<int> hit(A, 0);
    for (int a = 0; a < A; a++) {
        if (skip[a]) continue;
        const int ta = w.agents[a].type;
        const bool sa = w.response[a] == RESP_Static;
        for (int b = a + 1; b < A; b++) {
            if (skip[b]) continue;
            if (sa && w.response[b] == RESP_Static) continue;
            if (!(zlo[a] < zhi[b] && zlo[b] < zhi[a])) continue;
            if (!obb_collided(box[a], box[b])) continue;
            const int tb = w.agents[b].type;
            if (pair_ignored(ta, tb)) continue;
            hit[a] = hit[b] = 1;
            mark_info(&w.info[a * 5], tb);
            mark_info(&w.info[b * 5], ta);
        }
        if (sa) continue;            // road entities are Static: Static x Static is never a candidate
        for (int r = 0; r < w.R; r++) {
            const Road &rd = w.roads[r];
            // only RoadEdge and StopSign can survive collisionPairs for any agent type
            if (pair_ignored(ta, rd.type)) continue;
            float rlo, rhi;
            leaf_z(rd.pos.z, rd.scale.z, -1.f, 1.f, 0.f, rlo, rhi);
            if (!(zlo[a] < rhi && rlo < zhi[a])) continue;
            // cheap exact-safe reject: centres further apart than both half-diagonals (+ slack)
            float dx = w.agents[a].pos.x - rd.pos.x, dy = w.agents[a].pos.y - rd.pos.y;
            float ra = sqrtf(length2_xy(w.agents[a].scale.x, w.agents[a].scale.y));
            float rr = sqrtf(length2_xy(rd.scale.x, rd.scale.y));
            if (sqrtf(dx * dx + dy * dy) > (ra + rr) * 1.001f + 0.01f) continue;
            if (!obb_collided(box[a], obb_from(rd.pos, rd.rot, rd.scale))) continue;
            hit[a] = 1;
            mark_info(&w.info[a * 5], rd.type);
        }
    }
    for (int a = 0; a < A; a++) if (hit[a]) w.agents[a].collided = 1;
}

// rewardSystem, stepTrackerSystem, doneSystem: src/sim.cpp:560-626
void reward_done(World &w, bool decrement) {
    for (int a = 0; a < w.A; a++) {
        Agent &g = w.agents[a];
        const float dist = sqrtf(length2_xy(g.pos.x - g.gx, g.pos.y - g.gy));
        if (w.prm.rewardType == 0) w.reward[a] = -dist;
        else if (w.prm.rewardType == 1) w.reward[a] = dist < w.prm.distanceToGoalThreshold ? 1.f : 0.f;
        if (decrement) w.steps[a] = (int32_t)((uint32_t)w.steps[a] - 1u);
        const int t = w.steps[a];
        if (t == EPISODE_LEN && w.done[a] != 1) { w.done[a] = 0; continue; }
        else if (t == 0) w.done[a] = 1;
        if (w.done[a] != 1 || w.info[a * 5 + 3] != 1) {
            if (dist < w.prm.distanceToGoalThreshold) { w.done[a] = 1; w.info[a * 5 + 3] = 1; }
        }
    }
}

// ---------------------------------------------------------------- observations
// collectSelfObsSystem src/sim.cpp:168-186, collectAbsoluteObservationsSystem :769-783
void self_and_abs_obs(World &w) {
    for (int a = 0; a < w.A; a++) {
        const Agent &g = w.agents[a];
        float *so = &w.self_obs[(size_t)a * 8];
        so[0] = length(g.vel);
        so[1] = g.vsize.x; so[2] = g.vsize.y; so[3] = g.vsize.z;
        V3 rel = rotate(inv(g.rot), {g.gx - g.pos.x, g.gy - g.pos.y, 0.f});
        so[4] = rel.x; so[5] = rel.y;
        so[6] = g.collided ? 1.f : 0.f;
        so[7] = g.id;
        float *ab = &w.abs_obs[(size_t)a * 14];
        ab[0] = g.pos.x; ab[1] = g.pos.y; ab[2] = g.pos.z;
        ab[3] = g.rot.w; ab[4] = g.rot.x; ab[5] = g.rot.y; ab[6] = g.rot.z;   // Quat memory order w,x,y,z
        ab[7] = quat_to_yaw(g.rot);
        ab[8] = g.gx; ab[9] = g.gy;
        ab[10] = g.vsize.x; ab[11] = g.vsize.y; ab[12] = g.vsize.z;
        ab[13] = g.id;
    }
}

// collectPartnerObsSystem src/sim.cpp:188-240; OtherAgents order src/level_gen.cpp:450-464
void partner_obs(World &w) {
    if (w.prm.disableClassicalObs) return;
    const int P = w.S - 1;
    for (int a = 0; a < w.A; a++) {
        const Agent &g = w.agents[a];
        float *rows = &w.partner[(size_t)a * P * 9];
        int slot = 0;
        for (int j = 0; j < w.A; j++) {
            if (j == a) continue;
            const Agent &o = w.agents[j];
            V3 d = o.pos - g.pos;
            V3 rel = rotate(inv(g.rot), {d.x, d.y, 0.f});
            float *row = rows + (size_t)slot++ * 9;
            if (sqrtf(length2_xy(rel.x, rel.y)) > w.prm.observationRadius) {
                for (int k = 0; k < 8; k++) row[k] = 0.f;
                row[8] = -1.f;
            } else {
                row[0] = length(o.vel); row[1] = rel.x; row[2] = rel.y;
                row[3] = quat_to_yaw(mul(inv(g.rot), o.rot));
                row[4] = o.vsize.x; row[5] = o.vsize.y; row[6] = o.vsize.z;
                row[7] = (float)o.type; row[8] = o.id;
            }
        }
        for (; slot < P; slot++) {
            float *row = rows + (size_t)slot * 9;
            for (int k = 0; k < 8; k++) row[k] = 0.f;
            row[8] = -2.f;
        }
    }
}

// ReferenceFrame::observationOf, src/utils.hpp:27-65
struct MapRow { float v[9]; int idx; };
MapRow observation_of(const Agent &g, const Road &r, int idx) {
    V3 rel = rotate(inv(g.rot), {r.pos.x - g.pos.x, r.pos.y - g.pos.y, 0.f});
    MapRow m;
    m.v[0] = rel.x; m.v[1] = rel.y; m.v[2] = r.scale.x; m.v[3] = r.scale.y; m.v[4] = r.scale.z;
    m.v[5] = quat_to_yaw(mul(inv(g.rot), r.rot));
    m.v[6] = (float)r.type; m.v[7] = r.id; m.v[8] = r.mapType;
    m.idx = idx;
    return m;
}
inline bool closer(const MapRow &l, const MapRow &r) {       // cmp, src/knn.hpp:15-17
    return length2_xy(l.v[0], l.v[1]) < length2_xy(r.v[0], r.v[1]);
}

// collectMapObservationsSystem src/sim.cpp:242-280 and selectKNearestRoadEntities src/knn.hpp:83-158.
// The heap leaves its survivors in heap order (and a debug build of the reference sorts them); callers
// compare the rows of an agent as a SET, and knn_index names the road entity of every row.
void roadmap_obs(World &w) {
    if (w.prm.disableClassicalObs) return;
    const int K = K_ROADMAP;
    std::vector<MapRow> heap(K);
    for (int a = 0; a < w.A; a++) {
        const Agent &g = w.agents[a];
        float *rows = &w.roadmap[(size_t)a * K * 9];
        int32_t *kidx = &w.knn_index[(size_t)a * K];
        int n = 0;
        float zero_id = 0.f, zero_mt = 0.f;          // fillZeros: id = 0, mapType = 0 (knn.hpp:19-28)
        if (w.prm.roadObservationAlgorithm == 0) {
            const int first = std::min(w.R, K);
            for (int i = 0; i < first; i++) heap[i] = observation_of(g, w.roads[i], i);
            n = first;
            if (w.R >= K) {
                std::make_heap(heap.begin(), heap.end(), closer);
                for (int r = K; r < w.R; r++) {
                    MapRow cur = observation_of(g, w.roads[r], r);
                    if (!closer(cur, heap[0])) continue;
                    std::pop_heap(heap.begin(), heap.end(), closer);
                    heap[K - 1] = cur;
                    std::push_heap(heap.begin(), heap.end(), closer);
                }
            }
            // radiusFilter, knn.hpp:83-97
            int beyond = n, i = 0;
            while (i < beyond) {
                if (sqrtf(length2_xy(heap[i].v[0], heap[i].v[1])) <= w.prm.observationRadius) { ++i; continue; }
                heap[i] = heap[--beyond];
            }
            n = beyond;
        } else {
            // AllEntitiesWithRadiusFiltering: first K in index order inside the radius, MapObservation::zero() after
            for (int r = 0; r < w.R && n < K; r++) {
                MapRow cur = observation_of(g, w.roads[r], r);
                if (sqrtf(length2_xy(cur.v[0], cur.v[1])) > w.prm.observationRadius) continue;
                heap[n++] = cur;
            }
            zero_id = -1.f; zero_mt = -1.f;
        }
        for (int i = 0; i < K; i++) {
            float *row = rows + (size_t)i * 9;
            if (i < n) { memcpy(row, heap[i].v, sizeof(float) * 9); kidx[i] = heap[i].idx; }
            else {
                for (int k = 0; k < 7; k++) row[k] = 0.f;
                row[7] = zero_id; row[8] = zero_mt; kidx[i] = -1;
            }
        }
    }
}

// ---------------------------------------------------------------- lidar
// AABB::rayIntersects, math.inl:1575-1601 (fmaxf / fminf drop NaNs exactly as the reference's do)
bool aabb_ray(V3 lo, V3 hi, V3 o, V3 inv_d, float tmin, float tmax) {
    V3 tl{inv_d.x * (lo.x - o.x), inv_d.y * (lo.y - o.y), inv_d.z * (lo.z - o.z)};
    V3 tu{inv_d.x * (hi.x - o.x), inv_d.y * (hi.y - o.y), inv_d.z * (hi.z - o.z)};
    auto mn = [](float a, float b) { return a < b ? a : b; };   // Vector3::min / max are plain compares
    auto mx = [](float a, float b) { return a > b ? a : b; };
    float bmin = fmaxf(mn(tl.x, tu.x), fmaxf(mn(tl.y, tu.y), fmaxf(mn(tl.z, tu.z), tmin)));
    float bmax = fminf(mx(tl.x, tu.x), fminf(mx(tl.y, tu.y), fminf(mx(tl.z, tu.z), tmax)));
    return bmin <= bmax;
}

// traceRayIntoLeaf + traceRayIntoConvexPolyhedron for a box hull, broadphase.cpp:754-888.
// zlo/zhi: the hull's object-space z extent (cube -1..1, agent 0..2); x and y faces sit at +-1.
bool ray_leaf(V3 pos, Quat rot, V3 scale, float zlo, float zhi, V3 o, V3 d, float t_max, float *hit_t) {
    V3 lo = rotate(inv(rot), o - pos);
    lo.x /= scale.x; lo.y /= scale.y; lo.z /= scale.z;
    V3 ld = rotate(inv(rot), d);
    ld.x /= scale.x; ld.y /= scale.y; ld.z /= scale.z;
    V3 inv_d{1.f / ld.x, 1.f / ld.y, 1.f / ld.z};
    if (!aabb_ray({-1.f, -1.f, zlo}, {1.f, 1.f, zhi}, lo, inv_d, 0.f, t_max)) return false;
    const V3 normals[6] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    const float dist[6] = {1.f, 1.f, 1.f, 1.f, zhi, -zlo};
    float tfirst = 0.f, tlast = t_max;
    bool entered = false;
    for (int f = 0; f < 6; f++) {
        float denom = dot(normals[f], ld);
        float neg_dist = dist[f] - dot(normals[f], lo);
        if (denom == 0.0f) {
            if (neg_dist < 0.0f) return false;
        } else {
            float t = neg_dist / denom;
            if (denom < 0.0f) { if (t >= tfirst) { tfirst = t; entered = true; } }
            else { if (t <= tlast) tlast = t; }
            if (tfirst > tlast) return false;
        }
    }
    if (!entered) return false;      // only back faces hit: the origin is inside, treated as a miss
    *hit_t = tfirst;
    return true;
}

// lidarSystem, src/sim.cpp:394-460; BVH::traceRay broadphase.cpp:656-722 as a loop over every leaf
void lidar_obs(World &w) {
    if (!w.prm.enableLidar) return;
    for (int a = 0; a < w.A; a++) {
        const Agent &g = w.agents[a];
        V3 fwd = rotate(g.rot, {0.f, 1.f, 0.f});
        V3 right = rotate(g.rot, {1.f, 0.f, 0.f});
        const float head = w.controlled[a] ? w.action[(size_t)a * 10 + 2] : 0.f;
        for (int p = 0; p < 3; p++) {
            for (int i = 0; i < N_LIDAR; i++) {
                float theta = LIDAR_ANGLE * (2 * float(i) / float(N_LIDAR) - 1) + head;
                float x = cosf(theta), y = sinf(theta);
                V3 dir = x * right + y * fwd;
                float il = 1.f / length(dir);
                dir = {dir.x * il, dir.y * il, dir.z * il};
                V3 o = g.pos + LIDAR_OFFSETS[p] * V3{0.f, 0.f, 1.f};
                float t_max = LIDAR_DIST, t;
                int type = -1;
                for (int r = 0; r < w.R; r++) {
                    const Road &rd = w.roads[r];
                    if (ray_leaf(rd.pos, rd.rot, rd.scale, -1.f, 1.f, o, dir, t_max, &t)) { t_max = t; type = rd.type; }
                }
                for (int j = 0; j < w.A; j++) {
                    const Agent &ot = w.agents[j];
                    if (ray_leaf(ot.pos, ot.rot, ot.scale, 0.f, 2.f, o, dir, t_max, &t)) { t_max = t; type = ot.type; }
                }
                float *s = &w.lidar[(((size_t)a * 3 + p) * N_LIDAR + i) * 4];
                if (type < 0) { s[0] = 0.f; s[1] = 0.f; s[2] = 0.f; s[3] = 0.f; }
                else { s[0] = t_max; s[1] = (float)type; s[2] = t_max * x; s[3] = t_max * y; }
            }
        }
    }
}

// ---------------------------------------------------------------- bird's-eye-view grid
constexpr int BEV_RES = 200;      // consts::bev_rasterization_resolution, src/consts.hpp:49

// rasterizer::toGridCoords + rasterizeRotatedRectangle, src/rasterizer.hpp:12-78
void rasterize(float *grid, float cx, float cy, float yaw, float length, float width, size_t type, float radius) {
    const int resolution = BEV_RES;
    float half_w = width / 2.0f, half_l = length / 2.0f;
    float scale_px = (2 * radius) / resolution;
    float scale_m = resolution / (2 * radius);
    int gx = static_cast<int>((cx + radius) * scale_m), gy = static_cast<int>((cy + radius) * scale_m);
    gx = std::min(std::max(0, gx), resolution - 1); gy = std::min(std::max(0, gy), resolution - 1);
    float max_side = std::max(half_w, half_l);
    int box_radius = static_cast<int>(std::ceil(std::sqrt(2 * (max_side * max_side)) / scale_px));
    float cos_yaw = std::cos(-yaw), sin_yaw = std::sin(-yaw);
    for (int dy = -box_radius; dy <= box_radius; dy++) {
        for (int dx = -box_radius; dx <= box_radius; dx++) {
            int x = gx + dx, y = gy + dy;
            if (x < 0 || x >= resolution || y < 0 || y >= resolution) continue;
            float px = x * scale_px - radius, py = y * scale_px - radius;
            float ldx = px - cx, ldy = py - cy;
            float local_x = ldx * cos_yaw - ldy * sin_yaw;
            float local_y = ldx * sin_yaw + ldy * cos_yaw;
            constexpr float epsilon = 1e-3f;
            if (std::abs(local_x) <= half_l + epsilon && std::abs(local_y) <= half_w + epsilon)
                grid[(size_t)y * resolution + x] = (float)type;
        }
    }
}

// collectBevObservationsSystem, src/sim.cpp:462-555
void bev_obs(World &w) {
    if (w.prm.disableClassicalObs || !w.want_bev) return;
    const float radius = w.prm.observationRadius;
    for (int a = 0; a < w.A; a++) {
        const Agent &g = w.agents[a];
        float *grid = &w.bev[(size_t)a * BEV_RES * BEV_RES];
        std::fill(grid, grid + BEV_RES * BEV_RES, 0.f);
        int arr = 0;
        for (int r = 0; r < w.R && arr < K_ROADMAP; r++) {
            MapRow m = observation_of(g, w.roads[r], r);
            if (sqrtf(length2_xy(m.v[0], m.v[1])) > radius) continue;
            float d1 = std::max(m.v[3], (2 * radius / BEV_RES));        // minimum segment width
            rasterize(grid, m.v[0], m.v[1], m.v[5], m.v[2], d1, (size_t)m.v[6], radius);
            arr++;
        }
        for (int j = 0; j < w.A; j++) {
            if (j == a) continue;
            const Agent &o = w.agents[j];
            V3 d = o.pos - g.pos;
            V3 rel = rotate(inv(g.rot), {d.x, d.y, 0.f});
            float heading = quat_to_yaw(mul(inv(g.rot), o.rot));
            if (sqrtf(length2_xy(rel.x, rel.y)) > radius) continue;
            rasterize(grid, rel.x, rel.y, heading, o.vsize.x, o.vsize.y, (size_t)o.type, radius);
        }
    }
}

// setupRestOfTasks, src/sim.cpp:785-943
void rest_of_step(World &w, bool decrement) {
    collisions(w);
    reward_done(w, decrement);
    self_and_abs_obs(w);
    partner_obs(w);
    roadmap_obs(w);
    lidar_obs(w);
    bev_obs(w);
}

// resetAgent + resetAgentInterface + getZeroAction: src/level_gen.cpp:23-54, src/level_gen.hpp:16-38
void reset_world(World &w) {
    for (int a = 0; a < w.A; a++) {
        Agent &g = w.agents[a];
        g.pos = {g.traj[T_POS], g.traj[T_POS + 1], 1.f};
        g.rot = angle_axis_up(g.traj[T_HEAD]);
        if (g.resp0 == RESP_Static) g.vel = {0, 0, 0};
        else g.vel = {g.traj[T_VEL], g.traj[T_VEL + 1], 0.f};
        g.ang = {0, 0, 0};
        g.collided = 0;
        float *act = &w.action[(size_t)a * 10];
        for (int k = 0; k < 10; k++) act[k] = 0.f;
        if (w.prm.dynamicsModel == 3) act[2] = 1.f;
        w.steps[a] = EPISODE_LEN; w.done[a] = 0; w.reward[a] = 0.f;
        int32_t *info = &w.info[a * 5];
        info[0] = info[1] = info[2] = info[3] = 0; info[4] = g.type;
        w.response[a] = g.resp0;
    }
}

} // namespace

// ---------------------------------------------------------------- C interface for tests (ctypes)
extern "C" {

// agent_static [A,16] = {length,width,height, scale d0,d1,d2, goal x,y, type, id, responseType, controlled, meta[4]}
// traj [A,1456]; roads [R,12] = {x,y,z, qw,qz, d0,d1,d2, yaw, type, id, mapType}
void *gdo_create(const void *params, int S, int A, int R, const float *agent_static, const float *traj,
                 const float *roads) {
    World *w = new World();
    memcpy(&w->prm, params, sizeof(Params));
    w->S = S; w->A = A; w->R = R;
    w->traj.assign(traj, traj + (size_t)A * TRAJ_SIZE);
    w->agents.resize(A);
    w->roads.resize(R);
    for (int r = 0; r < R; r++) {
        const float *q = roads + (size_t)r * 12;
        Road &rd = w->roads[r];
        rd.pos = {q[0], q[1], q[2]}; rd.rot = {q[3], 0.f, 0.f, q[4]}; rd.scale = {q[5], q[6], q[7]};
        rd.yaw = q[8]; rd.type = (int)q[9]; rd.id = q[10]; rd.mapType = q[11];
    }
    const int P = S - 1;
    w->action.assign((size_t)S * 10, 0.f); w->reward.assign(S, 0.f); w->self_obs.assign((size_t)S * 8, 0.f);
    w->partner.assign((size_t)S * P * 9, 0.f); w->roadmap.assign((size_t)S * K_ROADMAP * 9, 0.f);
    w->lidar.assign((size_t)S * 3 * N_LIDAR * 4, 0.f); w->abs_obs.assign((size_t)S * 14, 0.f);
    w->done.assign(S, 1); w->steps.assign(S, 0); w->info.assign((size_t)S * 5, 0);
    w->controlled.assign(S, 0); w->response.assign(S, RESP_Static); w->knn_index.assign((size_t)S * K_ROADMAP, -1);
    for (int a = 0; a < A; a++) {
        const float *s = agent_static + (size_t)a * 16;
        Agent &g = w->agents[a];
        g.vsize = {s[0], s[1], s[2]}; g.scale = {s[3], s[4], s[5]}; g.gx = s[6]; g.gy = s[7];
        g.type = (int)s[8]; g.id = s[9]; g.resp0 = (int)s[10]; g.controlled = (int)s[11];
        g.traj = &w->traj[(size_t)a * TRAJ_SIZE];
        w->controlled[a] = g.controlled;
    }
    // padding slots (createPaddingEntities, src/level_gen.cpp:300-330): self id -1, every partner row zero()
    // (id -1), map rows MapObservation::zero()
    for (int a = A; a < S; a++) {
        w->self_obs[(size_t)a * 8 + 7] = -1.f;
        for (int j = 0; j < P; j++) w->partner[((size_t)a * P + j) * 9 + 8] = -1.f;
        for (int j = 0; j < K_ROADMAP; j++) {
            w->roadmap[((size_t)a * K_ROADMAP + j) * 9 + 7] = -1.f;
            w->roadmap[((size_t)a * K_ROADMAP + j) * 9 + 8] = -1.f;
        }
    }
    reset_world(*w);
    rest_of_step(*w, false);      // the constructor's Reset graph (src/mgr.cpp:565, src/sim.cpp:953-957)
    return w;
}
// switches the bird's-eye-view grid on (160 KB per agent slot) and computes it for the current state
void gdo_enable_bev(void *h) {
    World &w = *static_cast<World *>(h);
    w.want_bev = true;
    w.bev.assign((size_t)w.S * BEV_RES * BEV_RES, 0.f);
    bev_obs(w);
}
void gdo_destroy(void *h) { delete static_cast<World *>(h); }
void gdo_step(void *h) { World &w = *static_cast<World *>(h); movement(w); rest_of_step(w, true); }
// resets this world (do_reset != 0) and runs the Reset graph, which recomputes every world (sim.cpp:953-957)
void gdo_reset(void *h, int do_reset) {
    World &w = *static_cast<World *>(h);
    if (do_reset) reset_world(w);
    rest_of_step(w, false);
}
// tensor ids: 0 action 1 reward 2 self_obs 3 partner 4 roadmap 5 lidar 6 abs_obs 7 bev | 16 done 17 steps 18 info
// 19 controlled 20 response 21 knn_index
void *gdo_tensor(void *h, int id) {
    World &w = *static_cast<World *>(h);
    switch (id) {
    case 0: return w.action.data(); case 1: return w.reward.data(); case 2: return w.self_obs.data();
    case 3: return w.partner.data(); case 4: return w.roadmap.data(); case 5: return w.lidar.data();
    case 6: return w.abs_obs.data(); case 7: return w.bev.data(); case 16: return w.done.data(); case 17: return w.steps.data();
    case 18: return w.info.data(); case 19: return w.controlled.data(); case 20: return w.response.data();
    case 21: return w.knn_index.data();
    }
    return nullptr;
}

// Unit entry points for the reference's own known-answer tests (tests/CollisionDetectionTests.cpp,
// tests/EgocentricRoadObservationTests.cpp).
int gdo_obb_collided(const float *a /*x,y,yaw,d0,d1*/, const float *b) {
    Obb oa = obb_from({a[0], a[1], 0.f}, angle_axis_up(a[2]), {a[3], a[4], 1.f});
    Obb ob = obb_from({b[0], b[1], 0.f}, angle_axis_up(b[2]), {b[3], b[4], 1.f});
    return obb_collided(oa, ob) ? 1 : 0;
}
void gdo_observation_of(const float *ego /*x,y,yaw*/, const float *ent /*x,y,yaw*/, float *out /*x,y,heading*/) {
    Agent g{}; g.pos = {ego[0], ego[1], 0.f}; g.rot = angle_axis_up(ego[2]);
    Road r{}; r.pos = {ent[0], ent[1], 0.f}; r.rot = angle_axis_up(ent[2]); r.scale = {1, 1, 1};
    MapRow m = observation_of(g, r, 0);
    out[0] = m.v[0]; out[1] = m.v[1]; out[2] = m.v[5];
}
void gdo_forward_model(int model, const float *state /*x,y,z,yaw,vx,vy,vz,length*/, float *action /*10*/,
                       float *out /*x,y,z,yaw,vx,vy,vz,speed*/) {
    Agent g{}; g.pos = {state[0], state[1], state[2]}; g.rot = angle_axis_up(state[3]);
    g.vel = {state[4], state[5], state[6]}; g.vsize = {state[7], 0.f, 0.f};
    switch (model) {
    case 1: forward_bicycle(action, g); break;
    case 2: forward_delta(action, g); break;
    case 0: forward_classic(action, g); break;
    default: forward_state(action, g);
    }
    out[0] = g.pos.x; out[1] = g.pos.y; out[2] = g.pos.z; out[3] = quat_to_yaw(g.rot);
    out[4] = g.vel.x; out[5] = g.vel.y; out[6] = g.vel.z; out[7] = length(g.vel);
}

} // extern "C"
