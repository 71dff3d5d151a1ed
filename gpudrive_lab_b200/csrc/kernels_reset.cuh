// Scenario binding and world reset kernels.
//   k_bind  : copy a pooled scenario's static tensors into a world's export rows and write the
//             padding-slot defaults (createPersistentEntities / createPaddingEntities,
//             src/level_gen.cpp:308-336,396-465).
//   k_reset : re-initialise agents of flagged worlds from Trajectory[0]
//             (resetAgent / resetAgentInterface, src/level_gen.cpp:23-54).
#pragma once
#include "device_math.cuh"
#include "engine.h"
#include "kernels_policy.cuh"

namespace gd {

// grid = (num_listed_worlds, BIND_SPLIT); block = 256
constexpr int BIND_SPLIT = 8;

__global__ void __launch_bounds__(256)
k_bind(const int32_t *__restrict__ worlds, const SceneDev *__restrict__ scenes, StateDev st,
       Exports ex, int S, float *__restrict__ pol)
{
    // explicit work list, or (worlds == nullptr, auto-reset) every world whose reset flag is set
    const int w = worlds ? worlds[blockIdx.x] : (int)blockIdx.x;
    if (!worlds && !st.reset_flag[w]) return;
    const SceneDev &sc = scenes[st.world_scene[w]];
    const int A = sc.numAgents, R = sc.numRoads;
    const int tid = blockIdx.y * blockDim.x + threadIdx.x;
    const int nthr = gridDim.y * blockDim.x;

    if (!worlds)      // auto-reset binds pooled scenarios, which carry no deletions (set_maps clears the list too)
        for (int a = tid; a < S; a += nthr) ex.deleted[(size_t)w * S + a] = -1;
    // expert trajectories: live rows from the pool, padding rows Trajectory::zero (types.hpp:357-367)
    {
        float4 *dst = reinterpret_cast<float4 *>(ex.traj + (size_t)w * S * kTrajSize);
        const float4 *src = reinterpret_cast<const float4 *>(sc.traj);
        const int live4 = A * kTrajSize / 4, all4 = S * kTrajSize / 4;
        for (int i = tid; i < all4; i += nthr)
            dst[i] = i < live4 ? src[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // road-entity rows (setRoadEntitiesProps, level_gen.hpp:57-63); padding rows MapObservation::zero()
    {
        float *dst = ex.map_obs + (size_t)w * kMaxRoadEntityCount * kMapObsSize;
        for (int r = tid; r < kMaxRoadEntityCount; r += nthr) {
            float row[9];
            if (r < R) {
                float4 a = sc.road_a[r]; float2 b = sc.road_b[r]; float4 d = sc.road_d[r];
                uint32_t c = sc.road_c[r];
                row[0] = a.x; row[1] = a.y; row[2] = b.x; row[3] = b.y; row[4] = d.x;
                row[5] = d.y; row[6] = (float)(c >> 8); row[7] = d.z; row[8] = d.w;
            } else {
                row[0] = row[1] = row[2] = row[3] = row[4] = row[5] = row[6] = 0.f;
                row[7] = -1.f; row[8] = -1.f;
            }
#pragma unroll
            for (int k = 0; k < 9; k++) dst[(size_t)r * 9 + k] = row[k];
        }
    }
    // per-slot static exports + padding-slot defaults
    for (int a = tid; a < S; a += nthr) {
        const size_t wa = (size_t)w * S + a;
        if (a < A) {
            int4 m = sc.meta[a];
            reinterpret_cast<int4 *>(ex.metadata)[wa] = m;
            float4 c = sc.ag_c[a];
            ex.controlled[wa] = (int32_t)c.w;
            ex.response[wa] = (int32_t)c.z;
        } else {
            reinterpret_cast<int4 *>(ex.metadata)[wa] = make_int4(-1, -1, -1, -1);
            ex.controlled[wa] = 0;
            ex.response[wa] = RT_Static;
            ex.steps_remaining[wa] = 0;
            ex.done[wa] = 1;
            ex.reward[wa] = 0.f;
            int32_t *info = ex.info + wa * kInfoSize;
            info[0] = info[1] = info[2] = info[3] = 0; info[4] = ET_None;
            float *so = ex.self_obs + wa * kSelfObsSize;
            for (int k = 0; k < 7; k++) so[k] = 0.f;
            so[7] = -1.f;
            float *ab = ex.abs_obs + wa * kAbsObsSize;
            for (int k = 0; k < kAbsObsSize; k++) ab[k] = 0.f;
            float *ac = ex.action + wa * kActionSize;
            for (int k = 0; k < kActionSize; k++) ac[k] = 0.f;
            st.pos[wa] = make_float4(kPadX, kPadY, kPadZ, 0.f);
            st.rv[wa] = make_float4(1.f, 0.f, 0.f, 0.f);
            st.collided[wa] = 0;
        }
    }
    // padding slots: partner rows zero() (id -1), roadmap rows zero() (id -1, mapType -1), lidar 0
    {
        const int P = S - 1;
        const int padSlots = S - A;
        const long nP = (long)padSlots * P;
        float *pbase = ex.partner_obs + ((size_t)w * S + A) * P * kPartnerObsSize;
        for (long i = tid; i < nP; i += nthr) {
            float *row = pbase + i * kPartnerObsSize;
#pragma unroll
            for (int k = 0; k < 8; k++) row[k] = 0.f;
            row[8] = -1.f;
        }
        const long nM = (long)padSlots * kMaxAgentMapObs;
        float *mbase = ex.agent_roadmap + ((size_t)w * S + A) * kMaxAgentMapObs * kMapObsSize;
        int32_t *kbase = ex.knn_idx + ((size_t)w * S + A) * kMaxAgentMapObs;
        for (long i = tid; i < nM; i += nthr) {
            float *row = mbase + i * kMapObsSize;
#pragma unroll
            for (int k = 0; k < 7; k++) row[k] = 0.f;
            row[7] = -1.f; row[8] = -1.f;
            kbase[i] = -1;
        }
        const long nL = (long)padSlots * 3 * kNumLidarSamples * kLidarRowSize;
        float *lbase = ex.lidar + ((size_t)w * S + A) * 3 * kNumLidarSamples * kLidarRowSize;
        for (long i = tid; i < nL; i += nthr) lbase[i] = 0.f;
    }
    if (pol) {
        // fused policy observation (k_obs writes the live slots every step): the rows of the padding slots are the
        // image of their constant export rows (self obs zero, partner rows zero(), road-map rows zero() of type 0)
        const int PW = kPolicyEgo + (S - 1) * kPolicyPartner + kMaxAgentMapObs * kPolicyRoad;
        const float zrow[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, -1.f, -1.f};
        float oe[kPolicyEgo], op[kPolicyPartner], orow[kPolicyRoad];
        policy_ego_row<true>(0.f, 0.f, 0.f, 0.f, 0.f, 0.f, true, oe);
        policy_partner_row<true>(zrow, true, op);
        policy_road_row<true>(zrow, true, orow);
        const int road0 = kPolicyEgo + (S - 1) * kPolicyPartner;
        float *pb = pol + ((size_t)w * S + A) * PW;
        const long n = (long)(S - A) * PW;
        for (long i = tid; i < n; i += nthr) {
            const int k = (int)(i % PW);
            float v;
            if (k < kPolicyEgo) v = k < 3 ? oe[0] : (k < 5 ? oe[3] : oe[5]);            // (oe[0..2], oe[3..4] are equal)
            else if (k < road0) { const int q = (k - kPolicyEgo) % kPolicyPartner; v = (q == 1 || q == 2) ? op[1] : op[0]; }
            else { const int q = (k - road0) % kPolicyRoad; v = q < 2 ? orow[0] : (q == 6 ? orow[6] : orow[2]); }
            pb[i] = v;
        }
    }
    if (tid == 0) {
        ex.shape[w * 2 + 0] = A;
        ex.shape[w * 2 + 1] = R;
        ex.world_means[w * 3 + 0] = sc.mean[0];
        ex.world_means[w * 3 + 1] = sc.mean[1];
        ex.world_means[w * 3 + 2] = sc.mean[2];
    }
    if (tid < 32) {
        ex.map_name[w * 32 + tid] = sc.mapName[tid];
        ex.scenario_id[w * 32 + tid] = sc.scenarioId[tid];
    }
}

// grid = num_listed_worlds; block = S rounded up to a warp multiple
__global__ void k_reset(const int32_t *__restrict__ worlds, const SceneDev *__restrict__ scenes,
                        StateDev st, Exports ex, int S, int dynamicsModel)
{
    const int w = worlds ? worlds[blockIdx.x] : (int)blockIdx.x;
    if (!worlds && !st.reset_flag[w]) return;
    const SceneDev &sc = scenes[st.world_scene[w]];
    const int a = threadIdx.x;
    if (a >= sc.numAgents || a >= S) return;
    const size_t wa = (size_t)w * S + a;
    const float *tr = ex.traj + wa * kTrajSize;   // the Trajectory component, as resetAgent reads it
    const float x = tr[kTrajPos + 0], y = tr[kTrajPos + 1];
    float vx = tr[kTrajVel + 0], vy = tr[kTrajVel + 1];
    const float heading = tr[kTrajHead + 0];
    const float4 b = sc.ag_b[a];
    const float4 c = sc.ag_c[a];
    const int resp = (int)c.z;
    float qw, qz;
    angle_axis_z_hp(heading, qw, qz);
    if (resp == RT_Static) { vx = 0.f; vy = 0.f; }
    st.pos[wa] = make_float4(x, y, 1.f, 0.f);
    st.rv[wa] = make_float4(qw, qz, vx, vy);
    st.collided[wa] = 0;
    float *ac = ex.action + wa * kActionSize;     // getZeroAction, level_gen.hpp:16-38
#pragma unroll
    for (int k = 0; k < kActionSize; k++) ac[k] = 0.f;
    if (dynamicsModel == DM_State) ac[2] = 1.f;
    ex.steps_remaining[wa] = kEpisodeLen;
    ex.done[wa] = 0;
    ex.reward[wa] = 0.f;
    int32_t *info = ex.info + wa * kInfoSize;
    info[0] = info[1] = info[2] = info[3] = 0;
    info[4] = (int32_t)b.w;
    ex.response[wa] = resp;
}

// bind_worlds: the new world -> scene binding arrives with the work list
__global__ void k_set_scene(const int32_t *__restrict__ worlds, const int32_t *__restrict__ scene_ids, int n,
                            int32_t *__restrict__ world_scene)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) world_scene[worlds[i]] = scene_ids[i];
}

// Device-side auto-reset (gd_auto_reset): one warp per world. A world is finished when every controlled agent is
// done (what gpudrive/env/env_puffer.py tests before it resets a world) - or, with no controlled agent, when the
// episode has run out. A finished world moves on in the scenario pool: cursor' = (cursor + 1 + w % (n - 1)) % n,
// is re-bound to pool[cursor'] and flagged; k_bind / k_reset / the rest-of-step kernels then run over the flagged
// worlds only. No host round trip anywhere.
__global__ void __launch_bounds__(256)
k_autoreset_select(StateDev st, Exports ex, const SceneDev *__restrict__ scenes, const int32_t *__restrict__ pool,
                   int n_pool, int32_t *__restrict__ cursor, int W, int S)
{
    const int w = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (w >= W) return;
    const int A = scenes[st.world_scene[w]].numAgents;
    int ctrl = 0, pending = 0;
    for (int a = lane; a < A; a += 32) {
        const size_t wa = (size_t)w * S + a;
        const int c = ex.controlled[wa];
        ctrl += c ? 1 : 0;
        pending += (c && ex.done[wa] != 1) ? 1 : 0;
    }
    ctrl = __reduce_add_sync(0xffffffffu, ctrl);
    pending = __reduce_add_sync(0xffffffffu, pending);
    const bool finished = ctrl > 0 ? pending == 0 : (A == 0 || ex.steps_remaining[(size_t)w * S] <= 0);
    if (lane == 0) {
        int flag = 0;
        if (finished) {
            const int cur = cursor[w];
            const int nxt = n_pool > 1 ? (cur + 1 + w % (n_pool - 1)) % n_pool : 0;
            cursor[w] = nxt;
            st.world_scene[w] = pool[nxt];
            flag = 1;
            atomicAdd(&st.stats[ST_AUTO_RESETS], 1ull);
        }
        st.reset_flag[w] = flag;
    }
}

// stepHost: contiguous [W*S, cols] host actions -> the first `cols` columns of the [W*S, 10] action rows
__global__ void k_scatter_actions(const float *__restrict__ src, int cols, float *__restrict__ action, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t row = i / (size_t)cols;
    action[row * kActionSize + (i - row * (size_t)cols)] = src[i];
}

} // namespace gd
