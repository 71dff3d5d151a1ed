// k_agents: one CTA per world. Phase A movementSystem (src/sim.cpp:294-383, dynamics.hpp),
// phase B collisionDetectionSystem over agent x agent and agent x {road edge, stop sign}
// boxes (src/sim.cpp:628-747, obb.hpp, with the broadphase pair rules of
// external/madrona/src/physics/broadphase.cpp:928-991), phase C rewardSystem /
// stepTrackerSystem / doneSystem (src/sim.cpp:560-626). Agent boxes live in shared memory
// between the phases; a warp owns one agent in phase B and its lanes stride over the boxes.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace gd {

struct AgentBox {        // shared-memory record per agent slot, phase B
    Obb2 obb;            // 14 floats
    float cx, cy, rad;   // bounding circle for the cheap reject
    float zlo, zhi;      // broadphase leaf z-range (expanded), strict overlap test
    int32_t type;
    int32_t flags;       // bit0 skip (invalid expert / done), bit1 static response, bit2 live
};

constexpr int AB_SKIP = 1, AB_STATIC = 2, AB_LIVE = 4, AB_DONE0 = 8;   // DONE0: Done was set when the step began

// expandAABBWithMotion (broadphase.cpp:438-462) for one axis
__device__ __forceinline__ void expand_axis(float v, float &lo, float &hi) {
    float pos_delta = __fmul_rn(kLeafVelExpansion, v);
    float min_delta = __fsub_rn(pos_delta, 0.16f);
    float max_delta = __fadd_rn(pos_delta, 0.16f);
    if (min_delta < 0.f) lo = __fadd_rn(lo, min_delta);
    if (max_delta > 0.f) hi = __fadd_rn(hi, max_delta);
}

__device__ __forceinline__ int collide_info_bit(int otherType) {
    // src/sim.cpp:713-724
    if (otherType > ET_None && otherType <= ET_StopSign) return 1;   // collidedWithRoad
    if (otherType == ET_Vehicle) return 2;                             // collidedWithVehicle
    if (otherType <= ET_Cyclist) return 4;                             // collidedWithNonVehicle
    return 0;
}

// Sim::collisionPairs (src/sim.hpp:88-102): true if the pair is ignored
__device__ __forceinline__ bool pair_ignored(int a, int b) {
    auto one = [](int p, int q) {
        if (p == ET_Pedestrian || p == ET_Cyclist)
            return q == ET_RoadEdge || q == ET_RoadLine || q == ET_RoadLane || q == ET_CrossWalk || q == ET_SpeedBump;
        if (p == ET_Vehicle)
            return q == ET_CrossWalk || q == ET_SpeedBump || q == ET_RoadLine || q == ET_RoadLane;
        return false;
    };
    return one(a, b) || one(b, a);
}

template <int BLOCK>
__global__ void __launch_bounds__(BLOCK)
k_agents(const SceneDev *__restrict__ scenes, StateDev st, Exports ex, Params prm, int S,
         int decrement, int do_move, int only_flagged)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    AgentBox *box = reinterpret_cast<AgentBox *>(smem_raw);
    // episode statistics of this world's step (gd_stats): 0 -> 1 transitions of the Info / Done flags
    __shared__ unsigned s_cnt[ST_NUM];
    __shared__ unsigned s_cost[128];                   // (max_agents <= 128, engine.cu)
    if (threadIdx.x < ST_NUM) s_cnt[threadIdx.x] = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) st.retrace[0] = 0;     // k_obs -> k_lidar_fix hand-off list of this step

    const int w = blockIdx.x;
    if (only_flagged && !st.reset_flag[w]) return;      // auto-reset: only the worlds that were just re-bound
    const SceneDev &sc = scenes[st.world_scene[w]];
    const int A = sc.numAgents;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int NWARP = BLOCK / 32;

    // ------------------------------------------------------------------ phase A: movement
    for (int a = tid; a < A; a += BLOCK) {
        const size_t wa = (size_t)w * S + a;
        s_cost[a] = st.cost[wa];
        float4 p = st.pos[wa];
        float4 rv = st.rv[wa];
        int collided = st.collided[wa];
        int done = ex.done[wa];
        const int done0 = done;
        const int resp = ex.response[wa];
        const int controlled = ex.controlled[wa];
        const int t = ex.steps_remaining[wa];
        const float4 agb = sc.ag_b[a];

        if (do_move) {
            if (collided) {   // sim.cpp:302-323
                if (prm.collisionBehaviour == CB_AgentStop) {
                    done = 1; rv.z = 0.f; rv.w = 0.f; p.w = 0.f;
                } else if (prm.collisionBehaviour == CB_AgentRemoved) {
                    done = 1; p.x = kPadX; p.y = kPadY; p.z = kPadZ; rv.z = 0.f; rv.w = 0.f; p.w = 0.f;
                } else {
                    collided = 0;
                    int32_t *info = ex.info + wa * kInfoSize;
                    info[0] = 0; info[1] = 0; info[2] = 0;
                }
            }
            if (resp != RT_Static) {
                if (done) {   // sim.cpp:333-343
                    p.x = kPadX; p.y = kPadY; p.z = kPadZ; rv.z = 0.f; rv.w = 0.f; p.w = 0.f;
                } else if (controlled) {
                    float *act = ex.action + wa * kActionSize;
                    const float dt = 0.1f;
                    if (prm.dynamicsModel == DM_InvertibleBicycle) {        // dynamics.hpp:52-81
                        float acc = fmaxf(-6.0f, fminf(act[0], 6.0f));
                        float steer = fmaxf(-3.0f, fminf(act[1], 3.0f));
                        act[0] = acc; act[1] = steer;
                        float yaw = quat_to_yaw_hp(rv.x, rv.y);
                        float speed = len3(rv.z, rv.w, p.w);
                        double half_a = __dmul_rn(0.5, (double)acc);
                        double ddt = (double)dt;
                        double tx = __dmul_rn(__dmul_rn(__dmul_rn(half_a, (double)cos_hp(yaw)), ddt), ddt);
                        double ty = __dmul_rn(__dmul_rn(__dmul_rn(half_a, (double)sin_hp(yaw)), ddt), ddt);
                        p.x = (float)__dadd_rn((double)__fadd_rn(p.x, __fmul_rn(rv.z, dt)), tx);
                        p.y = (float)__dadd_rn((double)__fadd_rn(p.y, __fmul_rn(rv.w, dt)), ty);
                        double inner = __dadd_rn((double)__fmul_rn(speed, dt), __dmul_rn(__dmul_rn(half_a, ddt), ddt));
                        float delta_yaw = (float)__dmul_rn((double)steer, inner);
                        float new_yaw = angle_add(yaw, delta_yaw);
                        float new_speed = __fadd_rn(speed, __fmul_rn(acc, dt));
                        rv.z = __fmul_rn(new_speed, cos_hp(new_yaw));
                        rv.w = __fmul_rn(new_speed, sin_hp(new_yaw));
                        p.w = 0.f;
                        angle_axis_z_hp(new_yaw, rv.x, rv.y);
                    } else if (prm.dynamicsModel == DM_DeltaLocal) {         // dynamics.hpp:83-115
                        float yaw = quat_to_yaw_hp(rv.x, rv.y);
                        float c = cos_hp(yaw), s = sin_hp(yaw);
                        float dx = __fsub_rn(__fmul_rn(act[0], c), __fmul_rn(act[1], s));
                        float dy = __fadd_rn(__fmul_rn(act[0], s), __fmul_rn(act[1], c));
                        p.x = __fadd_rn(p.x, dx);
                        p.y = __fadd_rn(p.y, dy);
                        rv.z = __fdiv_rn(dx, dt);
                        rv.w = __fdiv_rn(dy, dt);
                        p.w = 0.f;
                        float new_yaw = angle_add(yaw, act[2]);
                        angle_axis_z_hp(new_yaw, rv.x, rv.y);
                    } else if (prm.dynamicsModel == DM_Classic) {            // dynamics.hpp:11-50
                        const float length = sc.ag_a[a].x;
                        float speed = len3(rv.z, rv.w, p.w);
                        float yaw = quat_to_yaw_hp(rv.x, rv.y);
                        float v = __fadd_rn(speed, __fmul_rn(__fmul_rn(0.5f, act[0]), dt));
                        float tanDelta = tan_hp(act[1]);
                        float beta = atan_hp(__fmul_rn(0.5f, tanDelta));
                        float ang = __fadd_rn(yaw, beta);
                        float dxx = __fmul_rn(v, cos_hp(ang)), dyy = __fmul_rn(v, sin_hp(ang));
                        float wz = __fdiv_rn(__fmul_rn(__fmul_rn(v, cos_hp(beta)), tanDelta), length);
                        float new_yaw = angle_add(yaw, __fmul_rn(wz, dt));
                        float new_speed = __fadd_rn(speed, __fmul_rn(act[0], dt));
                        p.x = __fadd_rn(p.x, __fmul_rn(dxx, dt));
                        p.y = __fadd_rn(p.y, __fmul_rn(dyy, dt));
                        p.z = 1.f;
                        angle_axis_z_hp(new_yaw, rv.x, rv.y);
                        rv.z = __fmul_rn(new_speed, cos_hp(new_yaw));
                        rv.w = __fmul_rn(new_speed, sin_hp(new_yaw));
                        p.w = 0.f;
                    } else {                                                 // dynamics.hpp:186-194
                        p.x = act[0]; p.y = act[1]; p.z = act[2];
                        rv.z = act[4]; rv.w = act[5]; p.w = act[6];
                        angle_axis_z_hp(act[3], rv.x, rv.y);
                    }
                } else {      // expert replay, sim.cpp:371-382 (index = episodeLen - t, pre-decrement)
                    int idx = kEpisodeLen - t;
                    idx = idx < 0 ? 0 : (idx > kTrajLen ? kTrajLen : idx);
                    const float *tr = ex.traj + wa * kTrajSize;
                    p.x = tr[kTrajPos + 2 * idx]; p.y = tr[kTrajPos + 2 * idx + 1]; p.z = 1.f;
                    rv.z = tr[kTrajVel + 2 * idx]; rv.w = tr[kTrajVel + 2 * idx + 1]; p.w = 0.f;
                    angle_axis_z_hp(tr[kTrajHead + idx], rv.x, rv.y);
                }
            }
            st.pos[wa] = p;
            st.rv[wa] = rv;
            ex.done[wa] = done;
        }

        // ---- box for phase B
        AgentBox &bx = box[a];
        const float yaw = quat_to_yaw_hp(rv.x, rv.y);
        st.yaw[wa] = yaw;                  // k_obs exports it (AbsoluteSelfObservation) and aims the lidar fan with it
        const float c = cos_hp(yaw), s = sin_hp(yaw);
        obb_from(p.x, p.y, c, s, agb.x, agb.y, bx.obb);
        bx.cx = p.x; bx.cy = p.y;
        bx.rad = sqrtf(agb.x * agb.x + agb.y * agb.y) * 1.0001f + 1e-3f;
        // leaf z-range: agent hull z in [0, 2] scaled by d2, expanded by motion (applyTRS + expandAABBWithMotion)
        float zlo = p.z, zhi = __fadd_rn(p.z, __fmul_rn(agb.z, 2.f));
        expand_axis(p.w, zlo, zhi);
        bx.zlo = zlo; bx.zhi = zhi;
        bx.type = (int)agb.w;
        // isInvalidExpertOrDone, sim.cpp:631-662 (hasCollided snapshot from before this pass)
        int skip = 0;
        if (!controlled) {
            int idx = kEpisodeLen - t;
            idx = idx < 0 ? 0 : (idx > kTrajLen ? kTrajLen : idx);
            const float valid = ex.traj[wa * kTrajSize + kTrajValid + idx];
            if (!(valid != 0.f)) skip = 1;
        } else if (done && !collided) {
            skip = 1;
        }
        bx.flags = AB_LIVE | (skip ? AB_SKIP : 0) | (resp == RT_Static ? AB_STATIC : 0) | (done0 ? AB_DONE0 : 0);
        st.collided[wa] = collided;
    }
    __syncthreads();

    // ------------------------------------------------------------------ phase B: collisions
    for (int a = warp; a < A; a += NWARP) {
        const AgentBox &me = box[a];
        int hit_bits = 0;
        unsigned nsat = 0;
        if (!(me.flags & AB_SKIP)) {
            // agent x agent
            for (int j = lane; j < A; j += 32) {
                if (j == a) continue;
                const AgentBox &ot = box[j];
                if (ot.flags & AB_SKIP) continue;
                if ((me.flags & AB_STATIC) && (ot.flags & AB_STATIC)) continue;   // broadphase.cpp:953-957
                if (!(me.zlo < ot.zhi && ot.zlo < me.zhi)) continue;              // AABB::overlaps, strict
                float dx = me.cx - ot.cx, dy = me.cy - ot.cy, rr = me.rad + ot.rad;
                if (dx * dx + dy * dy > rr * rr) continue;
                nsat++;
                if (!obb_collided(me.obb, ot.obb)) continue;
                if (pair_ignored(me.type, ot.type)) continue;
                hit_bits |= collide_info_bit(ot.type);
            }
            // agent x static road boxes (edges + stop signs); Static agents never pair with them
            if (!(me.flags & AB_STATIC)) {
                // one lane per bounding circle of 32 boxes, then the boxes of the circles the agent reaches
                const int CG = sc.numCollGroups;
                for (int gbase = 0; gbase < CG; gbase += 32) {
                  bool reach = false;
                  if (gbase + lane < CG) {
                      const float4 gc = sc.coll_grp[gbase + lane];
                      const float gx = me.cx - gc.x, gy = me.cy - gc.y, gr = me.rad + gc.z;
                      reach = gx * gx + gy * gy <= gr * gr;
                  }
                  for (unsigned gm = __ballot_sync(0xffffffffu, reach); gm; gm &= gm - 1) {
                    const int j = (gbase + __ffs(gm) - 1) * 32 + lane;
                    const float4 cu = sc.coll_cull[j];
                    float dx = me.cx - cu.x, dy = me.cy - cu.y, rr = me.rad + cu.z;
                    if (dx * dx + dy * dy > rr * rr) continue;
                    const int rtype = (int)cu.w;
                    if (pair_ignored(me.type, rtype)) continue;
                    const float4 q0 = sc.coll_obb[j * 4 + 0], q1 = sc.coll_obb[j * 4 + 1];
                    const float4 q2 = sc.coll_obb[j * 4 + 2], q3 = sc.coll_obb[j * 4 + 3];
                    // z: leaf range of the static box (q3.z +- q3.w, expanded by 0.16 each way)
                    float rzlo = __fsub_rn(__fsub_rn(q3.z, q3.w), 0.16f), rzhi = __fadd_rn(__fadd_rn(q3.z, q3.w), 0.16f);
                    if (!(me.zlo < rzhi && rzlo < me.zhi)) continue;
                    Obb2 rb;
                    rb.cx[0] = q0.x; rb.cy[0] = q0.y; rb.cx[1] = q0.z; rb.cy[1] = q0.w;
                    rb.cx[2] = q1.x; rb.cy[2] = q1.y; rb.cx[3] = q1.z; rb.cy[3] = q1.w;
                    rb.ax[0] = q2.x; rb.ay[0] = q2.y; rb.ax[1] = q2.z; rb.ay[1] = q2.w;
                    rb.org[0] = q3.x; rb.org[1] = q3.y;
                    nsat++;
                    if (!obb_collided(me.obb, rb)) continue;
                    hit_bits |= collide_info_bit(rtype);
                  }
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hit_bits |= __shfl_xor_sync(0xffffffffu, hit_bits, o);
        if (decrement) {
            nsat = __reduce_add_sync(0xffffffffu, nsat);
            if (lane == 0 && nsat) atomicAdd(&s_cnt[ST_SAT_TESTS], nsat);
        }
        if (lane == 0 && hit_bits) {
            const size_t wa = (size_t)w * S + a;
            st.collided[wa] = 1;
            int32_t *info = ex.info + wa * kInfoSize;
#pragma unroll
            for (int k = 0; k < 3; k++)
                if (hit_bits & (1 << k)) {
                    if (decrement && info[k] == 0) atomicAdd(&s_cnt[ST_COLLIDED_ROAD + k], 1u);
                    info[k] = 1;
                }
        }
    }
    __syncthreads();

    // ------------------------------------------------------------------ phase C: reward, --t, done
    if (decrement && tid == 0) s_cnt[ST_AGENT_STEPS] = (unsigned)A;
    for (int a = tid; a < A; a += BLOCK) {
        const size_t wa = (size_t)w * S + a;
        const float4 agc = sc.ag_c[a];
        if (decrement && ex.controlled[wa]) atomicAdd(&s_cnt[ST_CONTROLLED_STEPS], 1u);
        const float px = box[a].cx, py = box[a].cy;
        const float dxg = __fsub_rn(px, agc.x), dyg = __fsub_rn(py, agc.y);
        const float dist = __fsqrt_rn(len2_xy(dxg, dyg));
        float reward = 0.f;                                   // sim.cpp:560-587
        if (prm.rewardType == RW_DistanceBased) reward = -dist;
        else if (prm.rewardType == RW_OnGoalAchieved)
            reward = dist < prm.distanceToGoalThreshold ? 1.f : 0.f;
        else reward = ex.reward[wa];                          // Dense: unimplemented upstream (assert)
        ex.reward[wa] = reward;
        int t = ex.steps_remaining[wa];
        if (decrement) { t = (int)((uint32_t)t - 1u); ex.steps_remaining[wa] = t; }   // sim.cpp:589-592
        int done = ex.done[wa];                               // sim.cpp:597-626
        int32_t *info = ex.info + wa * kInfoSize;
        if (t == kEpisodeLen && done != 1) {
            ex.done[wa] = 0;
            continue;
        } else if (t == 0) {
            done = 1;
        }
        if (done != 1 || info[3] != 1) {
            if (dist < prm.distanceToGoalThreshold) {
                if (decrement && info[3] != 1) atomicAdd(&s_cnt[ST_GOALS_REACHED], 1u);
                done = 1; info[3] = 1;
            }
        }
        if (decrement && done == 1 && !(box[a].flags & AB_DONE0)) atomicAdd(&s_cnt[ST_DONES], 1u);
        ex.done[wa] = done;
    }
#ifndef GD_OBS_NO_COSTSORT
    // k_obs work order: the world's agents by decreasing cost of their last observation pass (ties: lower slot), so
    // that the warps of one k_obs CTA finish together; any permutation gives the same results
    for (int a = tid; a < A; a += BLOCK) {
        const unsigned c = s_cost[a];
        int r = 0;
        for (int j = 0; j < A; j++) { const unsigned cj = s_cost[j]; r += (cj > c || (cj == c && j < a)) ? 1 : 0; }
        st.order[(size_t)w * S + r] = (uint8_t)a;
    }
#endif
    if (decrement) {
        __syncthreads();
        if (tid < ST_NUM && s_cnt[tid]) atomicAdd(&st.stats[tid], (unsigned long long)s_cnt[tid]);
    }
}

} // namespace gd
