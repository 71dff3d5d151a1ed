// extern "C" surface of libgpudrive_b200.so (include/gpudrive_b200.h) over gd::Engine.
#include "../../include/gpudrive_b200.h"

#include <cstring>
#include <exception>
#include <stdexcept>
#include <string>
#include <vector>

#include "engine.h"
#include "scenario.h"

static thread_local std::string g_last_error;

namespace gd {
// for entry points that live in the CUDA translation unit (engine.cu): same per-thread message as GD_TRY's
void setLastError(const std::string &msg) { g_last_error = msg; }
}

static_assert(sizeof(GdParams) == sizeof(gd::Params), "GdParams and gd::Params must match");
static_assert(sizeof(GdTensorDesc) == sizeof(gd::TensorDesc), "tensor desc mismatch");
static_assert((int)GD_T_COUNT == (int)gd::T_NUM, "tensor id tables out of sync");
static_assert(sizeof(GdStats) == sizeof(gd::StepStats), "stats mismatch");

#define GD_TRY(body)                                   \
    try {                                              \
        body;                                          \
        return 0;                                      \
    } catch (const std::exception &e) {                \
        g_last_error = e.what();                       \
        return 1;                                      \
    } catch (...) {                                    \
        g_last_error = "unknown error";                \
        return 2;                                      \
    }

static gd::Engine *E(GdHandle h) {
    if (!h) throw std::runtime_error("null handle");
    return static_cast<gd::Engine *>(h);
}

static gd::Params toParams(const GdParams *p) {
    gd::Params q;
    memcpy(&q, p, sizeof(q));
    return q;
}

extern "C" {

int gd_create(const GdConfig *cfg, const char *const *scenes, int32_t n_scenes, GdHandle *out) {
    GD_TRY({
        if (!cfg || !scenes || !out) throw std::runtime_error("gd_create: null argument");
        std::vector<std::string> v;
        for (int i = 0; i < n_scenes; i++) v.emplace_back(scenes[i]);
        *out = new gd::Engine(cfg->gpu_id, cfg->max_agents, toParams(&cfg->params), v, cfg->flags);
    })
}

int gd_destroy(GdHandle h) { GD_TRY({ delete E(h); }) }
int gd_step(GdHandle h) { GD_TRY({ E(h)->step(); }) }
int gd_step_async(GdHandle h) { GD_TRY({ E(h)->stepAsync(); }) }
int gd_sync(GdHandle h) { GD_TRY({ E(h)->sync(); }) }
int gd_reset(GdHandle h, const int32_t *worlds, int32_t n) { GD_TRY({ E(h)->reset(worlds, n); }) }

int gd_set_maps(GdHandle h, const char *const *scenes, int32_t n) {
    GD_TRY({
        std::vector<std::string> v;
        for (int i = 0; i < n; i++) v.emplace_back(scenes[i]);
        E(h)->setMaps(v);
    })
}

int gd_delete_agents(GdHandle h, int32_t world, const int32_t *ids, int32_t n) {
    GD_TRY({ E(h)->deleteAgents(world, ids, n); })
}

int gd_tensor(GdHandle h, int32_t id, GdTensorDesc *out) {
    GD_TRY({
        gd::TensorDesc d = E(h)->tensor(id);
        memcpy(out, &d, sizeof(d));
    })
}

int gd_step_host(GdHandle h, const float *actions, int32_t action_cols, float *reward, int32_t *done,
                 int32_t *info) {
    GD_TRY({ E(h)->stepHost(actions, action_cols, reward, done, info); })
}

int gd_pool_add(GdHandle h, const char *path, int32_t *id_out) {
    GD_TRY({ *id_out = E(h)->poolAdd(path); })
}

int gd_bind_worlds(GdHandle h, const int32_t *worlds, const int32_t *scene_ids, int32_t n) {
    GD_TRY({ E(h)->bindWorlds(worlds, scene_ids, n); })
}

int gd_auto_reset_pool(GdHandle h, const int32_t *scene_ids, int32_t n) { GD_TRY({ E(h)->autoResetPool(scene_ids, n); }) }
int gd_auto_reset(GdHandle h) { GD_TRY({ E(h)->autoReset(); }) }

int gd_shaped_reward(GdHandle h, const GdRewardShaping *cfg, GdTensorDesc *out) {
    GD_TRY({
        gd::TensorDesc d = E(h)->shapedReward(cfg->mode, cfg->collision_weight, cfg->goal_achieved_weight,
                                              cfg->off_road_weight, cfg->log_distance_weight, cfg->agent_weights,
                                              cfg->world_time_steps);
        memcpy(out, &d, sizeof(d));
    })
}

int gd_expert_actions(GdHandle h, GdTensorDesc *out) {
    GD_TRY({
        gd::TensorDesc d = E(h)->expertActions();
        memcpy(out, &d, sizeof(d));
    })
}

int gd_pool_convert(const char *const *json_paths, int32_t n, float threshold, const char *out_path, int32_t threads) {
    GD_TRY({
        std::vector<std::string> paths;
        for (int32_t i = 0; i < n; i++) paths.emplace_back(json_paths[i]);
        gd::writeScenarioPool(out_path, paths, threshold, threads);
    })
}

int gd_pool_count(const char *pool_path, int32_t *count_out, float *threshold_out) {
    GD_TRY({ *count_out = gd::scenarioPoolCount(pool_path, threshold_out); })
}

int gd_stats(GdHandle h, GdStats *out) {
    GD_TRY({
        gd::StepStats s = E(h)->stats();
        memcpy(out, &s, sizeof(s));
    })
}

int gd_set_option(GdHandle h, int32_t option, int32_t value) { GD_TRY({ E(h)->setOption(option, value); }) }
int gd_kernel_times(GdHandle h, double *agents_ms, double *obs_ms, int64_t *steps) {
    GD_TRY({ E(h)->kernelTimes(agents_ms, obs_ms, steps); })
}
int gd_num_worlds(GdHandle h) { return h ? static_cast<gd::Engine *>(h)->numWorlds() : -1; }
int gd_max_agents(GdHandle h) { return h ? static_cast<gd::Engine *>(h)->maxAgents() : -1; }
uint64_t gd_launch_count(GdHandle h) { return h ? static_cast<gd::Engine *>(h)->launches() : 0; }
uint64_t gd_stream(GdHandle h) { return h ? (uint64_t)(uintptr_t)static_cast<gd::Engine *>(h)->stream() : 0; }
const char *gd_last_error(void) { return g_last_error.c_str(); }

int gd_scene_info(const char *path, const GdParams *params, int32_t max_agents, GdSceneInfo *out) {
    GD_TRY({
        gd::Params p = toParams(params);
        gd::ParsedMap m = gd::loadParsedMap(path, p.polylineReductionThreshold);
        gd::SceneData s = gd::buildScene(m, p, max_agents, nullptr, 0);
        out->num_agents = s.numAgents;
        out->num_roads = s.numRoads;
        out->num_controlled = s.numControlled;
        out->mean[0] = s.mean[0]; out->mean[1] = s.mean[1]; out->mean[2] = s.mean[2];
    })
}

int gd_scene_tensors(const char *path, const GdParams *params, int32_t max_agents, float *map_obs,
                     int32_t map_rows_cap, float *traj, float *agent_static, int32_t agent_cap) {
    GD_TRY({
        gd::Params p = toParams(params);
        gd::ParsedMap m = gd::loadParsedMap(path, p.polylineReductionThreshold);
        gd::SceneData s = gd::buildScene(m, p, max_agents, nullptr, 0);
        if (map_obs) {
            for (int r = 0; r < s.numRoads && r < map_rows_cap; r++) {
                float *row = map_obs + (size_t)r * 9;
                row[0] = s.rx[r]; row[1] = s.ry[r]; row[2] = s.rd0[r]; row[3] = s.rd1[r]; row[4] = s.rd2[r];
                row[5] = s.ryaw[r]; row[6] = (float)s.rtype[r]; row[7] = s.rid[r]; row[8] = (float)s.rmaptype[r];
            }
        }
        const int na = s.numAgents < agent_cap ? s.numAgents : agent_cap;
        if (traj && na) memcpy(traj, s.traj.data(), sizeof(float) * (size_t)na * gd::kTrajSize);
        if (agent_static) {
            for (int a = 0; a < na; a++) {
                float *row = agent_static + (size_t)a * 16;
                row[0] = s.vsize[a * 3]; row[1] = s.vsize[a * 3 + 1]; row[2] = s.vsize[a * 3 + 2];
                row[3] = s.scale[a * 3]; row[4] = s.scale[a * 3 + 1]; row[5] = s.scale[a * 3 + 2];
                row[6] = s.goal[a * 2]; row[7] = s.goal[a * 2 + 1];
                row[8] = (float)s.type[a]; row[9] = (float)s.id[a];
                row[10] = (float)s.response[a]; row[11] = (float)s.controlled[a];
                for (int k = 0; k < 4; k++) row[12 + k] = (float)s.meta[a * 4 + k];
            }
        }
    })
}

int gd_policy_obs(GdHandle h, int32_t normalize, GdTensorDesc *out) {
    GD_TRY({
        gd::TensorDesc d = E(h)->policyObs(normalize);
        memcpy(out, &d, sizeof(d));
    })
}

int gd_scene_road_pose(const char *path, const GdParams *params, int32_t max_agents, float *pose,
                       int32_t rows_cap) {
    GD_TRY({
        gd::Params p = toParams(params);
        gd::ParsedMap m = gd::loadParsedMap(path, p.polylineReductionThreshold);
        gd::SceneData s = gd::buildScene(m, p, max_agents, nullptr, 0);
        for (int r = 0; r < s.numRoads && r < rows_cap; r++) {
            float *row = pose + (size_t)r * 4;
            row[0] = s.rz[r]; row[1] = s.rqw[r]; row[2] = s.rqz[r]; row[3] = 0.f;
        }
    })
}

} // extern "C"
