// C API over the UNMODIFIED reference Manager (src/mgr.hpp:28-100) in ExecMode::CPU.
// Test infrastructure: used by tests/ (golden-vector generation, oracle validation)
// and by bench.py's cpu_baseline / --impl reference legs. Never on the product path.
#include "mgr.hpp"
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <dlfcn.h>
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>

using namespace madrona_gpudrive;

// DATA_DIR hook (see data_dir.hpp): directory holding cube_collision.obj and
// agent_collision_simplified.obj.
std::string gd_ref_data_dir()
{
    if (const char *e = std::getenv("GD_REF_ASSETS")) return std::string(e);
    Dl_info info;
    if (dladdr((void *)&gd_ref_data_dir, &info) && info.dli_fname) {
        std::string p(info.dli_fname);
        auto s = p.rfind('/');
        p = (s == std::string::npos) ? std::string(".") : p.substr(0, s);
        return p + "/../ref/assets/";
    }
    return "oracle/ref/assets/";
}

static void gd_ref_segv(int sig)
{
    void *frames[64];
    int n = backtrace(frames, 64);
    const char msg[] = "gd_ref: fatal signal, backtrace:\n";
    (void)!write(2, msg, sizeof(msg) - 1);
    backtrace_symbols_fd(frames, n, 2);
    _exit(128 + sig);
}

extern "C" {

// Debug aid: GD_REF_BACKTRACE=1 prints a native backtrace on SIGSEGV/SIGABRT.
void gd_ref_install_backtrace(void)
{
    static char altstack[1 << 16];
    stack_t ss; ss.ss_sp = altstack; ss.ss_size = sizeof(altstack); ss.ss_flags = 0;
    sigaltstack(&ss, nullptr);
    struct sigaction sa; memset(&sa, 0, sizeof(sa));
    sa.sa_handler = gd_ref_segv; sa.sa_flags = SA_ONSTACK;
    sigaction(SIGSEGV, &sa, nullptr); sigaction(SIGABRT, &sa, nullptr);
}

// Mirrors Parameters (src/init.hpp:111-127) field by field.
struct GdRefParams {
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
};

struct GdRefTensor {
    void *ptr;
    int32_t dtype;   // 0 = f32, 1 = i32
    int32_t ndim;
    int64_t shape[8];
};

void *gd_ref_create(const char *const *scenes, int n_scenes, const GdRefParams *p)
{
    Parameters params;
    params.polylineReductionThreshold = p->polylineReductionThreshold;
    params.observationRadius = p->observationRadius;
    params.rewardParams.rewardType = (RewardType)p->rewardType;
    params.rewardParams.distanceToGoalThreshold = p->distanceToGoalThreshold;
    params.rewardParams.distanceToExpertThreshold = p->distanceToExpertThreshold;
    params.collisionBehaviour = (CollisionBehaviour)p->collisionBehaviour;
    params.maxNumControlledAgents = p->maxNumControlledAgents;
    params.IgnoreNonVehicles = p->ignoreNonVehicles != 0;
    params.roadObservationAlgorithm = (FindRoadObservationsWith)p->roadObservationAlgorithm;
    params.initOnlyValidAgentsAtFirstStep = p->initOnlyValidAgentsAtFirstStep != 0;
    params.isStaticAgentControlled = p->isStaticAgentControlled != 0;
    params.enableLidar = p->enableLidar != 0;
    params.disableClassicalObs = p->disableClassicalObs != 0;
    params.dynamicsModel = (DynamicsModel)p->dynamicsModel;
    params.readFromTracksToPredict = p->readFromTracksToPredict != 0;

    Manager::Config cfg;
    cfg.execMode = madrona::ExecMode::CPU;
    cfg.gpuID = 0;
    for (int i = 0; i < n_scenes; i++) cfg.scenes.emplace_back(scenes[i]);
    cfg.params = params;
    cfg.enableBatchRenderer = false;
    return new Manager(cfg);
}

void gd_ref_destroy(void *h) { delete (Manager *)h; }
void gd_ref_step(void *h) { ((Manager *)h)->step(); }
void gd_ref_reset(void *h, const int32_t *worlds, int n)
{
    ((Manager *)h)->reset(std::vector<int32_t>(worlds, worlds + n));
}
void gd_ref_set_maps(void *h, const char *const *scenes, int n)
{
    std::vector<std::string> v;
    for (int i = 0; i < n; i++) v.emplace_back(scenes[i]);
    ((Manager *)h)->setMaps(v);
}
void gd_ref_delete_agents(void *h, int32_t world, const int32_t *ids, int n)
{
    std::unordered_map<int32_t, std::vector<int32_t>> m;
    m[world] = std::vector<int32_t>(ids, ids + n);
    ((Manager *)h)->deleteAgents(m);
}

// Tensor ids follow the SimManager getter names (src/bindings.cpp:111-149).
int gd_ref_tensor(void *h, const char *name, GdRefTensor *out)
{
    Manager *m = (Manager *)h;
    std::string n(name);
    auto fill = [&](const madrona::py::Tensor &t) {
        out->ptr = t.devicePtr();
        out->dtype = t.type() == madrona::py::TensorElementType::Float32 ? 0 : 1;
        out->ndim = (int32_t)t.numDims();
        for (int i = 0; i < out->ndim; i++) out->shape[i] = t.dims()[i];
        return 0;
    };
    if (n == "action") return fill(m->actionTensor());
    if (n == "reward") return fill(m->rewardTensor());
    if (n == "done") return fill(m->doneTensor());
    if (n == "self_observation") return fill(m->selfObservationTensor());
    if (n == "map_observation") return fill(m->mapObservationTensor());
    if (n == "partner_observations") return fill(m->partnerObservationsTensor());
    if (n == "agent_roadmap") return fill(m->agentMapObservationsTensor());
    if (n == "lidar") return fill(m->lidarTensor());
    if (n == "bev_observation") return fill(m->bevObservationTensor());
    if (n == "steps_remaining") return fill(m->stepsRemainingTensor());
    if (n == "shape") return fill(m->shapeTensor());
    if (n == "controlled_state") return fill(m->controlledStateTensor());
    if (n == "absolute_self_observation") return fill(m->absoluteSelfObservationTensor());
    if (n == "info") return fill(m->infoTensor());
    if (n == "response_type") return fill(m->responseTypeTensor());
    if (n == "expert_trajectory") return fill(m->expertTrajectoryTensor());
    if (n == "world_means") return fill(m->worldMeansTensor());
    if (n == "metadata") return fill(m->metadataTensor());
    if (n == "deleted_agents") return fill(m->deletedAgentsTensor());
    if (n == "map_name") return fill(m->mapNameTensor());
    if (n == "scenario_id") return fill(m->scenarioIdTensor());
    return -1;
}

int gd_ref_consts(int32_t *out4)
{
    out4[0] = (int32_t)consts::kMaxAgentCount;
    out4[1] = (int32_t)consts::kMaxRoadEntityCount;
    out4[2] = (int32_t)consts::kMaxAgentMapObservationsCount;
    out4[3] = (int32_t)consts::numLidarSamples;
    return 0;
}

}
