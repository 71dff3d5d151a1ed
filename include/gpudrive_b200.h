/* C ABI of the B200-native GPUDrive step engine (libgpudrive_b200.so).
 *
 * Drop-in boundary for the path behind the reference's `madrona_gpudrive.SimManager`
 * (nanobind class, src/bindings.cpp:91-149) / `Manager` (src/mgr.hpp:28-100). The reference has
 * no C ABI; each entry point below names the Manager member it replaces. Plain pointers and
 * sizes only; no torch / C++ types cross this boundary. Every call returns 0 on success and a
 * non-zero code on failure, with the message available from gd_last_error() (the reference
 * aborts the process instead: FATAL / assert / REQ_CUDA).
 *
 * Threading: one caller thread per handle. gd_step / gd_reset / gd_set_maps / gd_delete_agents
 * return after the device is idle, as Manager::step does (external/madrona/src/mw/cuda_exec.cpp:1892).
 * Tensor pointers are device pointers on GdConfig.gpu_id and stay valid for the handle's lifetime.
 *
 * Stream contract: the engine launches on its own *blocking* stream (gd_stream()), as Madrona does
 * (external/madrona/include/madrona/cuda_utils.inl:62). Writes into the action tensor (or any other tensor) that
 * the caller enqueued on the legacy default stream - what torch does by default, e.g.
 * `sim.action_tensor().to_torch()[:, :, :3].copy_(a)` (gpudrive/env/env_torch.py:652) - are therefore ordered
 * before the next gd_step / gd_step_async / gd_reset / gd_policy_obs ..., and the caller's later reads on that
 * stream after them; no explicit synchronisation is needed. A caller that writes from a non-blocking stream of
 * its own must order that stream against gd_stream() itself.
 */
#ifndef GPUDRIVE_B200_H
#define GPUDRIVE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Parameters, src/init.hpp:111-127 (bool fields widened to int32). Enum values:
 * rewardType 0 DistanceBased | 1 OnGoalAchieved | 2 Dense (src/init.hpp:76-81)
 * collisionBehaviour 0 AgentStop | 1 AgentRemoved | 2 Ignore (:90-95)
 * dynamicsModel 0 Classic | 1 InvertibleBicycle | 2 DeltaLocal | 3 State (:97-103)
 * roadObservationAlgorithm 0 KNearestEntitiesWithRadiusFiltering | 1 AllEntitiesWithRadiusFiltering (:105-109) */
typedef struct GdParams {
    float polylineReductionThreshold;
    float observationRadius;
    int32_t rewardType;
    float distanceToGoalThreshold;
    float distanceToExpertThreshold;
    int32_t collisionBehaviour;
    uint32_t maxNumControlledAgents;
    int32_t ignoreNonVehicles;
    int32_t roadObservationAlgorithm;
    int32_t initOnlyValidAgentsAtFirstStep;
    int32_t isStaticAgentControlled;
    int32_t enableLidar;
    int32_t disableClassicalObs;
    int32_t dynamicsModel;
    int32_t readFromTracksToPredict;
} GdParams;

/* Manager::Config, src/mgr.hpp:30-44. max_agents is consts::kMaxAgentCount (src/consts.hpp:11),
 * a constructor constant here: 128 reproduces the reference's tensor shapes. */
typedef struct GdConfig {
    int32_t gpu_id;
    int32_t max_agents;
    GdParams params;
    int32_t flags;      /* GD_FLAG_* */
} GdConfig;

/* GD_FLAG_BEV: allocate and compute bev_observation_tensor every step (collectBevObservationsSystem,
 * src/sim.cpp:462-555). The reference always does; here it is opt-in because the export is 160 KB per agent
 * slot (W x S x 200 x 200 floats: 21 GB at 1024 worlds, more than one GPU holds at 8192). */
enum GdFlags { GD_FLAG_BEV = 1 };

typedef void *GdHandle;

/* Tensor ids: one per SimManager getter (src/bindings.cpp:111-149, src/mgr.cpp:656-902). */
enum GdTensorId {
    GD_T_ACTION = 0,            /* action_tensor                    f32 [W,S,10]      (input) */
    GD_T_REWARD,                /* reward_tensor                    f32 [W,S,1]       */
    GD_T_DONE,                  /* done_tensor                      i32 [W,S,1]       */
    GD_T_SELF_OBSERVATION,      /* self_observation_tensor          f32 [W,S,8]       */
    GD_T_MAP_OBSERVATION,       /* map_observation_tensor           f32 [W,10000,9]   */
    GD_T_PARTNER_OBSERVATIONS,  /* partner_observations_tensor      f32 [W,S,S-1,9]   */
    GD_T_AGENT_ROADMAP,         /* agent_roadmap_tensor             f32 [W,S,200,9]   */
    GD_T_LIDAR,                 /* lidar_tensor                     f32 [W,S,3,50,4]  */
    GD_T_BEV_OBSERVATION,       /* bev_observation_tensor           f32 [W,S,200,200,1] (needs GD_FLAG_BEV) */
    GD_T_STEPS_REMAINING,       /* steps_remaining_tensor           i32 [W,S,1]       */
    GD_T_SHAPE,                 /* shape_tensor                     i32 [W,2]         */
    GD_T_CONTROLLED_STATE,      /* controlled_state_tensor          i32 [W,S,1]       */
    GD_T_ABSOLUTE_SELF_OBSERVATION, /* absolute_self_observation_tensor f32 [W,S,14]  */
    GD_T_VALID_STATE,           /* valid_state_tensor               i32 [W,S,1] zeros */
    GD_T_INFO,                  /* info_tensor                      i32 [W,S,5]       */
    GD_T_RESPONSE_TYPE,         /* response_type_tensor             i32 [W,S,1]       */
    GD_T_EXPERT_TRAJECTORY,     /* expert_trajectory_tensor         f32 [W,S,1456]    */
    GD_T_WORLD_MEANS,           /* world_means_tensor               f32 [W,3]         */
    GD_T_METADATA,              /* metadata_tensor                  i32 [W,S,4]       */
    GD_T_DELETED_AGENTS,        /* deleted_agents_tensor            i32 [W,S]         */
    GD_T_MAP_NAME,              /* map_name_tensor                  i32 [W,32]        */
    GD_T_SCENARIO_ID,           /* scenario_id_tensor               i32 [W,32]        */
    GD_T_KNN_INDEX,             /* (new) road-entity index of each agent_roadmap row, -1 pad: i32 [W,S,200] */
    GD_T_COUNT
};

typedef struct GdTensorDesc {
    void *ptr;          /* device pointer */
    int32_t dtype;      /* 0 = float32, 1 = int32 */
    int32_t ndim;
    int64_t shape[8];   /* C-contiguous */
    int32_t device;     /* CUDA device ordinal */
} GdTensorDesc;

typedef struct GdStats {
    uint64_t steps;
    uint64_t agent_steps;        /* live agents x steps */
    uint64_t controlled_steps;   /* live controlled agents x steps */
    uint64_t episodes_reset;     /* worlds reset by any route (ctor, reset, set_maps, bind_worlds, auto_reset) */
    /* (agent, step) pairs at which Info.collidedWith{Road,Vehicle,NonVehicle} / Info.reachedGoal / Done went
     * 0 -> 1 (src/sim.cpp:597-626,713-724), accumulated on the device by the step kernel */
    uint64_t collided_road, collided_vehicle, collided_nonvehicle, goals_reached, dones;
    uint64_t sat_tests;          /* two-sided OBB SAT tests evaluated (pairs that survived the cheap rejects) */
} GdStats;

/* Manager::Manager(cfg) — src/mgr.cpp:565 (Impl::init :399-563, then reset({})). num_worlds = n_scenes. */
int gd_create(const GdConfig *cfg, const char *const *scenes, int32_t n_scenes, GdHandle *out);
/* Manager::~Manager — src/mgr.cpp:567 */
int gd_destroy(GdHandle h);
/* Manager::step — src/mgr.cpp:569-580 */
int gd_step(GdHandle h);
/* Launch a step on the engine stream without waiting; pair with gd_sync (no reference equivalent). */
int gd_step_async(GdHandle h);
int gd_sync(GdHandle h);
/* Manager::reset(worldsToReset) — src/mgr.cpp:582-588 (triggerReset :904-920 + Reset graph) */
int gd_reset(GdHandle h, const int32_t *worlds, int32_t n);
/* Manager::setMaps — src/mgr.cpp:590-654 */
int gd_set_maps(GdHandle h, const char *const *scenes, int32_t n_scenes);
/* Manager::deleteAgents — src/mgr.cpp:665-715 (one world per call) */
int gd_delete_agents(GdHandle h, int32_t world, const int32_t *agent_ids, int32_t n);
/* Manager::*Tensor() — src/mgr.cpp:656-902 */
int gd_tensor(GdHandle h, int32_t tensor_id, GdTensorDesc *out);
/* Host-buffer step: copies actions[W,S,action_cols] host->device into action_tensor[:, :, :action_cols],
 * steps, copies reward / done / info (any may be NULL) device->host, then returns.
 * Equivalent of the env's `action_tensor().copy_(a); sim.step(); reward.cpu()` round trip
 * (gpudrive/env/env_torch.py:606-660). */
int gd_step_host(GdHandle h, const float *actions, int32_t action_cols, float *reward, int32_t *done,
                 int32_t *info);
/* The action staging of gd_step_host without any allocation. This entry point has no counterpart in the reference;
 * it was added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). It copies actions[n_rows, action_cols] (host - pinned or
 * pageable - or device memory) into a staging array that lives in the library's device image (a file-scope
 * `static __device__` array of 4 Mi floats = 16 MiB, which every GPU that loads the library holds until exit) and
 * scatters them into the first `action_cols` columns of action_rows_out[n_rows, 10], a DEVICE buffer on the
 * current device that the caller owns (e.g. the GD_T_ACTION tensor); the other columns are left untouched. Both
 * operations are enqueued on `stream` (a cudaStream_t as an integer, 0 = the legacy default stream); the call does
 * not wait for them. It calls no cudaMalloc / cudaFree / allocating library routine. n_rows * action_cols beyond
 * 4 194 304 floats is rejected with an error (gd_last_error). The first call on a device resolves the array's
 * address (cudaGetSymbolAddress): make that one call outside graph capture. All calls share the one array, which
 * keeps the last call's contents: calls whose work overlaps on the GPU (different streams, no ordering between
 * them) race; calls on one stream, or ordered by events, are safe. */
int gd_scatter_actions_noalloc(const float *actions, int64_t n_rows, int32_t action_cols, float *action_rows_out,
                               uint64_t stream);
/* Flat policy observation of every agent slot, computed on the device from the current exports in one pass:
 * what GPUDriveTorchEnv.get_obs() builds in Python with the default ego_state + partner_obs + road_map_obs
 * (gpudrive/env/env_torch.py:756-891,1172-1216; normalisation constants gpudrive/env/constants.py).
 * out: f32 [W, S, 6 + (S-1)*6 + 200*13], valid until the next call. normalize: 0 = raw (config.norm_obs off),
 * 1 = normalised with torch's CPU arithmetic (`x / c` is an IEEE division), 2 = normalised with torch's CUDA
 * arithmetic (`x / c` is x times the float reciprocal of c, ATen BinaryDivTrueKernel.cu) - bit-identical to what
 * get_obs() returns with device="cpu" / device="cuda" respectively. */
int gd_policy_obs(GdHandle h, int32_t normalize, GdTensorDesc *out);
/* Shaped reward of every agent slot, computed on the device from info / reward / expert_trajectory /
 * absolute_self_observation in one pass: GPUDriveTorchEnv.get_rewards (gpudrive/env/env_torch.py:469-604).
 * mode: 0 "sparse_on_goal_achieved" (the reward tensor), 1 "weighted_combination", 2 "reward_conditioned"
 * (agent_weights: device f32 [W,S,3] = collision, goal_achieved, off_road), 3 "distance_to_logs"
 * (world_time_steps: device i32 [W], or NULL for episode_len - steps_remaining). out: f32 [W,S], valid until the
 * next call. */
typedef struct GdRewardShaping {
    int32_t mode;
    float collision_weight, goal_achieved_weight, off_road_weight, log_distance_weight;
    const float *agent_weights;
    const int32_t *world_time_steps;
} GdRewardShaping;
int gd_shaped_reward(GdHandle h, const GdRewardShaping *cfg, GdTensorDesc *out);
/* Expert actions of the logged trajectories: GPUDriveTorchEnv.get_expert_actions (env_torch.py:1445-1509).
 * out: f32 [W,S,91,3] clamped inferred actions (Classic / InvertibleBicycle: +-6, +-0.3, -; DeltaLocal: +-6, +-6,
 * +-pi) or [W,S,91,10] (x, y, 1, yaw, vx, vy, 0, 0, 0, 0) for the State model; valid until the next call. */
int gd_expert_actions(GdHandle h, GdTensorDesc *out);
/* Device scenario pool: load a scenario once, re-bind worlds to pooled scenarios + reset them. */
int gd_pool_add(GdHandle h, const char *scene_path, int32_t *scene_id_out);
int gd_bind_worlds(GdHandle h, const int32_t *worlds, const int32_t *scene_ids, int32_t n);
/* Device-side auto-reset from the scenario pool (BASELINE config 4; replaces the env's host loop "find finished
 * worlds -> sim.reset(list)" of gpudrive/env/env_puffer.py and the set_maps resampling of
 * gpudrive/env/env_torch.py:1351-1370). gd_auto_reset_pool registers the pooled scene ids once (world w starts at
 * entry w % n). gd_auto_reset, called after a step, finds on the device every world whose controlled agents are
 * all done, advances its cursor (cursor + 1 + w % (n - 1)) % n, re-binds it to that pooled scene, resets it and
 * recomputes its observations - stream-ordered, without a host round trip, and without waiting for the device. */
int gd_auto_reset_pool(GdHandle h, const int32_t *scene_ids, int32_t n);
int gd_auto_reset(GdHandle h);
/* Scenario pool files (replaces the per-world MapReader JSON parse of Manager::Impl::init /
 * setMaps, src/mgr.cpp:399-470,590-654, for pools that are loaded again and again). gd_pool_convert parses
 * n JSON scenarios on `threads` host threads (<= 0: all cores) and writes them into one binary file; the file
 * depends on polylineReductionThreshold only. Afterwards "<pool file>#<index>" is accepted wherever a
 * scenario path is (gd_create, gd_set_maps, gd_pool_add, gd_scene_*) and yields bit-identical tensors. */
int gd_pool_convert(const char *const *json_paths, int32_t n, float polylineReductionThreshold,
                    const char *out_path, int32_t threads);
int gd_pool_count(const char *pool_path, int32_t *count_out, float *threshold_out);
int gd_stats(GdHandle h, GdStats *out);
/* Engine options (no reference equivalent): GD_OPT_KNN_INDEX (default 1) writes the GD_T_KNN_INDEX side
 * buffer; GD_OPT_PROFILE (default 0) brackets the two step kernels with CUDA events on the engine stream;
 * GD_OPT_GRAPH (default 1) launches a step as one CUDA graph (the reference's GPU backend does,
 * external/madrona/src/mw/cuda_exec.cpp:1699,1891); GD_OPT_ASYNC_AUX (default 0) lets gd_policy_obs /
 * gd_shaped_reward / gd_expert_actions return without waiting for the device. */
enum GdOption { GD_OPT_KNN_INDEX = 0, GD_OPT_PROFILE = 1, GD_OPT_GRAPH = 2, GD_OPT_ASYNC_AUX = 3,
                /* (default 0) the step's observation kernel also writes the gd_policy_obs tensor (normalize = 2,
                 * torch-CUDA arithmetic) as a second output, from the rows it holds on chip: gd_policy_obs(h, 2, ..)
                 * then returns it without a launch, and the 11.8 KB per slot it would re-read are never moved */
                GD_OPT_FUSED_POLICY_OBS = 4 };
int gd_set_option(GdHandle h, int32_t option, int32_t value);
/* Accumulated device time (ms) of k_agents / k_obs over the steps taken since GD_OPT_PROFILE was enabled. */
int gd_kernel_times(GdHandle h, double *agents_ms, double *obs_ms, int64_t *steps);
int gd_num_worlds(GdHandle h);
int gd_max_agents(GdHandle h);
/* kernels launched since creation (bench.py's gpu_launches) */
uint64_t gd_launch_count(GdHandle h);
/* cudaStream_t the engine launches on (as an integer), for timing with CUDA events */
uint64_t gd_stream(GdHandle h);
const char *gd_last_error(void);

/* Host-only scenario loader check (no GPU needed): parses a scenario and returns entity counts.
 * Replaces MapReader::parseAndWriteOut + createPersistentEntities for inspection
 * (src/MapReader.cpp:55-62, src/level_gen.cpp:396-465). */
typedef struct GdSceneInfo {
    int32_t num_agents, num_roads, num_controlled;
    float mean[3];
} GdSceneInfo;
int gd_scene_info(const char *scene_path, const GdParams *params, int32_t max_agents, GdSceneInfo *out);
/* Fills caller buffers with the host-built static tensors of one scenario (any pointer may be NULL):
 * map_obs [num_roads,9], traj [num_agents,1456], agent_static [num_agents,16] =
 * {length,width,height, scale d0,d1,d2, goal x,y, type, id, responseType, controlled, meta[4]}. */
int gd_scene_tensors(const char *scene_path, const GdParams *params, int32_t max_agents,
                     float *map_obs, int32_t map_rows_cap, float *traj, float *agent_static,
                     int32_t agent_cap);
/* Position.z and Rotation (w, z; x = y = 0) of every road entity: pose [num_roads,4] = {z, qw, qz, 0}
 * (makeRoadEdge / makeCube / makeStopSign, src/level_gen.cpp:150-290). Host only. */
int gd_scene_road_pose(const char *scene_path, const GdParams *params, int32_t max_agents, float *pose,
                       int32_t rows_cap);

#ifdef __cplusplus
}
#endif
#endif /* GPUDRIVE_B200_H */
