// Device-side restatement of two pieces of GPUDriveTorchEnv that sit right after the step:
//   k_shaped_reward  get_rewards(), gpudrive/env/env_torch.py:469-604: the weighted combination of the Info
//                    flags, optionally with per-agent weights ("reward_conditioned") or the distance-to-logs
//                    term; the reference clones info / expert_trajectory / absolute_self_observation to do it
//                    (the trajectory clone alone is 5.8 KB per agent slot per step) and loops over worlds in Python.
//   k_expert_actions get_expert_actions(), env_torch.py:1445-1509 with LogTrajectory
//                    (gpudrive/datatypes/trajectory.py:21-39): the clamped inferred actions, or the
//                    (x, y, 1, yaw, vx, vy, 0, 0, 0, 0) rows of the State model.
// Same float32 operations in torch's order; exp() of the distance term is the only transcendental (CUDA expf
// vs. torch's: within 2 ulp).
#pragma once
#include "engine.h"

namespace gd {

enum RewardShaping : int32_t { RS_Sparse = 0, RS_Weighted = 1, RS_Conditioned = 2, RS_DistanceToLogs = 3 };

struct RewardShapingArgs {
    int32_t mode;
    float collision_weight, goal_achieved_weight, off_road_weight, log_distance_weight;
    const float *agent_weights;         // [W,S,3] collision, goal_achieved, off_road (RS_Conditioned)
    const int32_t *world_time_steps;    // [W] or null: kEpisodeLen - steps_remaining of the world's first slot
};

// one thread per agent slot
__global__ void __launch_bounds__(256)
k_shaped_reward(Exports ex, RewardShapingArgs a, float *__restrict__ out, int W, int S)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * S) return;
    if (a.mode == RS_Sparse) { out[i] = ex.reward[i]; return; }           // reward_tensor().squeeze(2)
    const int32_t *info = ex.info + (size_t)i * kInfoSize;
    const float off_road = (float)info[0];
    const float collided = __fadd_rn((float)info[1], (float)info[2]);      // info[:, :, 1:3].float().sum(2)
    const float goal = (float)info[3];
    float cw = a.collision_weight, gw = a.goal_achieved_weight, ow = a.off_road_weight;
    if (a.mode == RS_Conditioned) {
        const float *w = a.agent_weights + (size_t)i * 3;
        cw = w[0]; gw = w[1]; ow = w[2];
    }
    float r = __fadd_rn(__fadd_rn(__fmul_rn(cw, collided), __fmul_rn(gw, goal)), __fmul_rn(ow, off_road));
    if (a.mode == RS_DistanceToLogs) {
        const int w = i / S;
        int t = a.world_time_steps ? a.world_time_steps[w] : kEpisodeLen - ex.steps_remaining[(size_t)w * S];
        t = t < 0 ? t + kTrajLen : t;                                       // torch indexing wraps negatives
        t = min(max(t, 0), kTrajLen - 1);
        const float *traj = ex.traj + (size_t)i * kTrajSize + kTrajPos;
        const float *ab = ex.abs_obs + (size_t)i * kAbsObsSize;
        const float dx = __fsub_rn(traj[2 * t], ab[0]), dy = __fsub_rn(traj[2 * t + 1], ab[1]);
        const float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));   // torch.norm(dim=-1)
        r = __fadd_rn(r, __fmul_rn(a.log_distance_weight, expf(-dist)));
    }
    out[i] = r;
}

// one thread per (agent slot, time step); cols = 3 (Classic / InvertibleBicycle / DeltaLocal) or 10 (State)
__global__ void __launch_bounds__(256)
k_expert_actions(Exports ex, float *__restrict__ out, size_t n, int model)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t slot = i / kTrajLen;
    const int t = (int)(i - slot * kTrajLen);
    const float *traj = ex.traj + slot * kTrajSize;
    if (model == DM_State) {
        float *o = out + i * 10;
        o[0] = traj[kTrajPos + 2 * t]; o[1] = traj[kTrajPos + 2 * t + 1]; o[2] = 1.f;
        o[3] = traj[kTrajHead + t];
        o[4] = traj[kTrajVel + 2 * t]; o[5] = traj[kTrajVel + 2 * t + 1];
        o[6] = o[7] = o[8] = o[9] = 0.f;
        return;
    }
    const float *act = traj + kTrajInv + (size_t)t * kActionSize;
    float a0 = act[0], a1 = act[1], a2 = act[2];
    auto clampf = [](float v, float lo, float hi) { return v != v ? v : fminf(fmaxf(v, lo), hi); };   // torch.clamp keeps NaN
    a0 = clampf(a0, -6.f, 6.f);
    if (model == DM_DeltaLocal) { a1 = clampf(a1, -6.f, 6.f); a2 = clampf(a2, -3.14159274f, 3.14159274f); }
    else a1 = clampf(a1, -0.3f, 0.3f);
    float *o = out + i * 3;
    o[0] = a0; o[1] = a1; o[2] = a2;
}

} // namespace gd
