// Host manager of the B200 step engine: owns the device scenario pool, the world state and the
// export buffers, and launches the step / reset kernels. Replaces Manager::Impl + the Madrona
// executors behind src/mgr.cpp:399-588 for ExecMode::CUDA.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "gd_types.h"
#include "scenario.h"

namespace gd {

constexpr int kMaxZClasses = 8;
constexpr unsigned kRetraceCap = 1u << 16;

// Device view of one pooled scenario (all pointers into the pool arena).
struct SceneDev {
    const float4 *road_a;     // [Rpad] x, y, qw, qz
    const float2 *road_b;     // [Rpad] d0, d1
    const uint32_t *road_c;   // [Rpad] zclass | type << 8
    const float4 *road_d;     // [Rpad] d2, yaw(heading), id, mapType   (row emission only)
    // The same road entities sorted by (z-class, Morton code of the position) and cut into groups of 32
    // (each class padded to a multiple of 32 with far-away dummies), for the observation kernel:
    const float4 *s_a;        // [G*32] x, y, qw, qz
    const float2 *s_b;        // [G*32] d0, d1
    const uint32_t *s_c;      // [G*32] zclass | type << 8 | original road index << 16   (dummy: zclass 8)
    const float4 *s_d;        // [G*32] d2, yaw(heading), id, mapType
    const float4 *grp;        // [G][2] bounds of the group's boxes: circle cx, cy, r | z-class bit mask, then the
                              //        rectangle around the same centre: axis ux, uy, half extents a, b
    // collidable road boxes (RoadEdge, StopSign) in spatial order, 32 per bounding circle:
    const float4 *coll_grp;   // [Cpad/32] cx, cy, r around the bounding circles of the 32 boxes
    const float4 *coll_cull;  // [Cpad] cx, cy, bounding radius, type
    const float4 *coll_obb;   // [Cpad*4] corners(8) | axes(4) | origin(2) z dz
    const float *traj;        // [A][1456]
    const float4 *ag_a;       // [A] length, width, height, id
    const float4 *ag_b;       // [A] d0, d1, d2, type
    const float4 *ag_c;       // [A] goal.x, goal.y, responseType, controlled
    const int4 *meta;         // [A]
    float zc_z[kMaxZClasses];   // distinct (z, d2) pairs of the road entities
    float zc_d2[kMaxZClasses];
    int32_t numZClasses;
    int32_t numRoads, numAgents, numColl, numGroups, numCollGroups;
    float mean[3];
    int32_t mapName[32];
    int32_t scenarioId[32];
};

// Export buffers: reference AoS layouts, [world][slot]... (src/mgr.cpp:656-902)
struct Exports {
    float *action;        // [W,S,10]
    float *reward;        // [W,S,1]
    int32_t *done;        // [W,S,1]
    int32_t *info;        // [W,S,5]
    float *self_obs;      // [W,S,8]
    float *abs_obs;       // [W,S,14]
    float *partner_obs;   // [W,S,S-1,9]
    float *agent_roadmap; // [W,S,200,9]
    float *map_obs;       // [W,10000,9]
    float *lidar;         // [W,S,3,50,4]
    int32_t *steps_remaining; // [W,S,1]
    int32_t *shape;       // [W,2]
    int32_t *controlled;  // [W,S,1]
    int32_t *response;    // [W,S,1]
    float *traj;          // [W,S,1456]
    float *world_means;   // [W,3]
    int32_t *metadata;    // [W,S,4]
    int32_t *deleted;     // [W,S]
    int32_t *map_name;    // [W,32]
    int32_t *scenario_id; // [W,32]
    int32_t *valid_state; // [W,S,1] (never written by the reference)
    float *bev;           // [W,S,200,200,1] only with GD_FLAG_BEV (else null)
    int32_t *knn_idx;     // [W,S,200] debug side buffer: road-entity index per roadmap row, -1 pad
};

// Internal SoA agent state
struct StateDev {
    float4 *pos;          // [W,S] x, y, z, vz
    float4 *rv;           // [W,S] qw, qz, vx, vy
    float *yaw;           // [W,S] utils::quatToYaw of the agent's rotation, written by k_agents (fp64 atan2, once per step)
    int32_t *collided;    // [W,S] CollisionDetectionEvent.hasCollided
    int32_t *world_scene; // [W]
    int32_t *reset_flag;  // [W]
    uint32_t *cost;       // [W,S] cycles k_obs spent on the agent in its last launch
    uint8_t *order;       // [W,S] the world's agents by decreasing cost (k_agents): k_obs takes agents in this order, so
                          //       the eight warps of a CTA carry agents of similar cost (results do not depend on it)
    uint32_t *retrace;         // [1 + kRetraceCap] lidar rays k_obs hands to k_lidar_fix: count, then slot * 150 + ray
    unsigned long long *stats; // [16] episode statistics accumulators (StatSlot), written by k_agents
};

// slots of StateDev::stats
enum StatSlot : int {
    ST_AGENT_STEPS = 0, ST_CONTROLLED_STEPS, ST_COLLIDED_ROAD, ST_COLLIDED_VEHICLE, ST_COLLIDED_NONVEHICLE,
    ST_GOALS_REACHED, ST_DONES, ST_AUTO_RESETS, ST_SAT_TESTS, ST_NUM
};

enum TensorId : int32_t {
    T_ACTION = 0, T_REWARD, T_DONE, T_SELF_OBS, T_MAP_OBS, T_PARTNER_OBS, T_AGENT_ROADMAP, T_LIDAR,
    T_BEV, T_STEPS_REMAINING, T_SHAPE, T_CONTROLLED, T_ABS_OBS, T_VALID_STATE, T_INFO, T_RESPONSE,
    T_TRAJ, T_WORLD_MEANS, T_METADATA, T_DELETED, T_MAP_NAME, T_SCENARIO_ID, T_KNN_IDX, T_NUM
};

struct TensorDesc {
    void *ptr;
    int32_t dtype;   // 0 = f32, 1 = i32
    int32_t ndim;
    int64_t shape[8];
    int32_t device;
};

struct StepStats {      // gd_stats; everything but `steps` is accumulated on the device by k_agents
    uint64_t steps;
    uint64_t agent_steps;       // live agents x steps
    uint64_t controlled_steps;  // live controlled agents x steps
    uint64_t episodes_reset;    // worlds reset (reset / set_maps / bind_worlds / auto_reset), the ctor's included
    // (agent, step) pairs at which Info.collidedWith* / Info.reachedGoal / Done went 0 -> 1 (src/sim.cpp:597-626,713-724)
    uint64_t collided_road, collided_vehicle, collided_nonvehicle, goals_reached, dones;
    uint64_t sat_tests;         // two-sided OBB separating-axis tests evaluated (candidates that survived the cheap rejects)
};

class Engine {
public:
    Engine(int gpu_id, int max_agents, const Params &params, const std::vector<std::string> &scenes, int flags = 0);
    ~Engine();

    void step();                 // launch + sync (reference semantics, src/mgr.cpp:569-580)
    void stepAsync();            // launch only, on stream()
    void sync();
    void reset(const int32_t *worlds, int n);
    void setMaps(const std::vector<std::string> &scenes);
    void deleteAgents(int world, const int32_t *ids, int n);
    TensorDesc tensor(int id) const;
    // e2e entry: actions from HOST memory, results to HOST memory, copies inside the call
    void stepHost(const float *actions, int action_cols, float *reward, int32_t *done, int32_t *info);
    // scenario pool (device auto-reset)
    int poolAdd(const std::string &path);       // returns scene id
    void bindWorlds(const int32_t *worlds, const int32_t *scene_ids, int n);   // + reset
    // device-side auto-reset: finished worlds move on to the next scenario of the pool, without the host
    void autoResetPool(const int32_t *scene_ids, int n);
    void autoReset();
    StepStats stats();
    // flat normalised policy observation (GPUDriveTorchEnv.get_obs) computed from the current exports
    TensorDesc policyObs(int normalize);
    // env post-processing on the device (kernels_env.cuh)
    TensorDesc shapedReward(int mode, float collision_w, float goal_w, float off_road_w, float log_dist_w,
                            const float *agent_weights, const int32_t *world_time_steps);
    TensorDesc expertActions();
    // options: 0 = write the knn index side buffer (default 1), 1 = per-kernel CUDA-event timing (default 0),
    // 2 = launch a step as one CUDA graph (default 1), 3 = policy_obs / shaped_reward / expert_actions return
    // without waiting for the device (default 0: they return when the tensor is ready), 4 = k_obs also writes the
    // flat policy observation (torch-CUDA arithmetic) as a second output (default 0)
    void setOption(int option, int value);
    // accumulated device time of the two step kernels since profiling was switched on (ms) and the step count
    void kernelTimes(double *agents_ms, double *obs_ms, int64_t *steps);
    cudaStream_t stream() const { return stream_; }
    int numWorlds() const { return W_; }
    int maxAgents() const { return S_; }
    int lastKernelCount() const { return kernels_per_step_; }
    uint64_t launches() const { return launches_; }

private:
    struct SceneHost {
        std::string key;
        SceneDev dev;
        int32_t numAgents, numRoads, numControlled;
        void *block = nullptr;       // arena block holding the scene's tables
        size_t block_bytes = 0;
        bool live = true;            // false: evicted, the id waits in free_ids_ for the next upload
        bool pinned = false;         // the caller holds the id (gd_pool_add, gd_bind_worlds, gd_auto_reset_pool): never evicted
        uint64_t last_use = 0;       // set_maps / deleteAgents generation that last bound a world to it
    };
    int sceneFor(const std::string &path, const std::vector<int32_t> &deleted);
    int uploadScene(const SceneData &s, const std::string &key);
    void *arenaAlloc(size_t bytes, size_t *block_bytes);
    void evictUnusedScenes();
    void syncSceneTable();
    void launchBindReset(const std::vector<int32_t> &worlds, bool rebind, bool scene_ids_uploaded = false);
    void launchRest(bool decrement, bool only_flagged = false);
    void allocExports();

    int gpu_, W_, S_;
    bool bev_enabled_ = false;
    Params params_;
    cudaStream_t stream_ = nullptr;
    std::vector<std::string> scene_paths_;            // per world
    std::vector<std::vector<int32_t>> deleted_;       // per world
    std::vector<int32_t> world_scene_;                // per world -> scene id
    std::vector<SceneHost> scenes_;
    std::map<std::string, int> scene_index_;
    std::map<std::string, std::shared_ptr<ParsedMap>> parsed_cache_;
    SceneDev *d_scenes_ = nullptr;
    size_t d_scenes_cap_ = 0;
    bool scene_table_dirty_ = true;
    std::vector<void *> arena_slabs_;
    size_t arena_off_ = 0, arena_cap_ = 0;
    std::vector<std::pair<void *, size_t>> arena_free_;   // blocks of evicted scenes, reused by later uploads
    std::vector<int> free_ids_;                           // scene ids of evicted scenes
    uint64_t bind_generation_ = 0;
    int max_roads_ = 0;        // over the pool
    int max_scene_agents_ = 0; // over the pool: k_obs launches CTAs for these slots only
    Exports ex_{};
    StateDev st_{};
    std::vector<void *> owned_;
    int32_t *d_worklist_ = nullptr;   // [W] scratch for reset lists
    int32_t *d_pool_ids_ = nullptr, *d_pool_cursor_ = nullptr;   // auto-reset pool and the per-world cursor into it
    int n_pool_ = 0;
    bool host_binding_stale_ = false;   // auto-reset re-bound worlds on the device: world_scene_ needs a read-back
    void refreshHostBinding();
    int32_t *h_worklist_ = nullptr;   // [W] pinned staging of the reset lists (stream-ordered upload)
    cudaEvent_t ev_worklist_ = nullptr;   // the last upload from h_worklist_ has been consumed
    float *h_pinned_actions_ = nullptr;
    float *d_shaped_reward_ = nullptr;    // [W,S], allocated on first use
    float *d_expert_actions_ = nullptr;   // [W,S,91,3 or 10], allocated on first use
    float *d_policy_obs_ = nullptr;       // [W,S,6 + (S-1)*6 + 200*13], allocated on first use
    float *d_act_stage_ = nullptr;        // [W*S*10] contiguous landing buffer of stepHost's actions
    cudaStream_t copy_stream_ = nullptr;  // stepHost: reward / done / info leave while k_obs still runs
    cudaEvent_t ev_agents_ = nullptr, ev_copied_ = nullptr;
    bool mark_agents_done_ = false;
    int kernels_per_step_ = 0;
    uint64_t launches_ = 0;
    uint64_t steps_ = 0, episodes_reset_ = 0;
    int opt_knn_idx_ = 1, opt_profile_ = 0, opt_graph_ = 1, opt_async_aux_ = 0, opt_fused_policy_ = 0;
    // one step = one graph launch: [ev] k_agents [ev] k_obs [ev] (k_bev); rebuilt when a launch parameter changes
    struct StepGraphKey {
        const void *scenes = nullptr; const void *pol = nullptr; int groups = -1, knn_idx = -1, profile = -1;
        bool operator==(const StepGraphKey &o) const {
            return scenes == o.scenes && pol == o.pol && groups == o.groups && knn_idx == o.knn_idx && profile == o.profile;
        }
    };
    StepGraphKey graph_key_;
    cudaGraph_t graph_ = nullptr;
    cudaGraphExec_t graph_exec_ = nullptr;
    cudaGraphNode_t graph_ev_nodes_[3] = {nullptr, nullptr, nullptr};
    void buildStepGraph(const StepGraphKey &key);
    void destroyStepGraph();
    int obsGroups() const;
    float *fusedPolicy() const { return opt_fused_policy_ ? d_policy_obs_ : nullptr; }
    int policyWidth() const;
    float key_thr_ = 0.f;
    static constexpr int kProfRing = 64;
    std::vector<cudaEvent_t> prof_ev_;      // 3 events per ring slot
    int prof_pending_ = 0;                  // slots recorded, not yet folded into the sums
    double prof_agents_ms_ = 0, prof_obs_ms_ = 0;
    int64_t prof_steps_ = 0;
    void foldProfile();
};

} // namespace gd
