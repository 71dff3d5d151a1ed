// Headless driver of the B200 step engine through its C ABI only (no Python, no torch): the counterpart of
// the reference's `headless` executable (src/headless.cpp:16-167). Same command line and the same two report
// lines, plus options to choose the scenes.
//
//   gd_headless CUDA NUM_STEPS [--rand-actions] [--worlds N] [--gpu ID] [--scenes a.json b.json ...]
//
// Without --scenes it looks for the reference's three example scenes relative to the working directory, as the
// reference does. "CPU" is refused: this engine has no CPU backend.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../include/gpudrive_b200.h"

#define CHECK(call)                                                                  \
    do {                                                                             \
        if ((call) != 0) {                                                           \
            fprintf(stderr, "%s failed: %s\n", #call, gd_last_error());              \
            return 1;                                                                \
        }                                                                            \
    } while (0)

int main(int argc, char *argv[])
{
    if (argc < 3) {
        fprintf(stderr, "%s TYPE NUM_STEPS [--rand-actions] [--worlds N] [--gpu ID] [--scenes a.json ...]\n", argv[0]);
        return -1;
    }
    const std::string type(argv[1]);
    if (type == "CPU") {
        fprintf(stderr, "ExecMode CPU: this engine has no CPU backend (the reference's own CPU build is oracle/_ref)\n");
        return -1;
    }
    if (type != "CUDA") {
        fprintf(stderr, "Invalid ExecMode\n");
        return -1;
    }
    const uint64_t num_steps = std::stoul(argv[2]);
    bool rand_actions = false;
    int worlds = 0, gpu = 0;
    std::vector<std::string> scenes;
    for (int i = 3; i < argc; i++) {
        const std::string a(argv[i]);
        if (a == "--rand-actions") rand_actions = true;
        else if (a == "--worlds" && i + 1 < argc) worlds = atoi(argv[++i]);
        else if (a == "--gpu" && i + 1 < argc) gpu = atoi(argv[++i]);
        else if (a == "--scenes") { while (i + 1 < argc && argv[i + 1][0] != '-') scenes.emplace_back(argv[++i]); }
        else { fprintf(stderr, "unknown argument %s\n", a.c_str()); return -1; }
    }
    if (scenes.empty())      // src/headless.cpp:37-41
        scenes = {"../data/processed/examples/tfrecord-00000-of-00150_78.json",
                  "../data/processed/examples/tfrecord-00043-of-00150_223.json",
                  "../data/processed/examples/tfrecord-00149-of-00150_111.json"};
    if (worlds <= 0) worlds = (int)scenes.size();
    std::vector<const char *> paths;
    for (int w = 0; w < worlds; w++) paths.push_back(scenes[w % scenes.size()].c_str());

    GdConfig cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gpu_id = gpu;
    cfg.max_agents = 128;
    // src/headless.cpp:54-65; the remaining fields are the defaults of Parameters (src/init.hpp:111-127)
    cfg.params.polylineReductionThreshold = 1.0f;
    cfg.params.observationRadius = 100.0f;
    cfg.params.rewardType = 0;
    cfg.params.distanceToGoalThreshold = 0.5f;
    cfg.params.distanceToExpertThreshold = 0.5f;
    cfg.params.collisionBehaviour = 0;
    cfg.params.maxNumControlledAgents = 0;
    cfg.params.initOnlyValidAgentsAtFirstStep = 1;
    GdHandle h = nullptr;
    CHECK(gd_create(&cfg, paths.data(), worlds, &h));

    GdTensorDesc shape, action;
    CHECK(gd_tensor(h, GD_T_SHAPE, &shape));
    CHECK(gd_tensor(h, GD_T_ACTION, &action));
    std::vector<int32_t> world_shape((size_t)worlds * 2);
    cudaSetDevice(gpu);
    if (cudaMemcpy(world_shape.data(), shape.ptr, world_shape.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess) {
        fprintf(stderr, "cannot read the shape tensor\n");
        return 1;
    }
    const int S = (int)action.shape[1], AC = (int)action.shape[2];
    std::vector<float> host_actions((size_t)worlds * S * AC, 0.f);

    std::random_device rd;
    std::mt19937 rand_gen(rd());
    std::uniform_real_distribution<float> acc_gen(-3.0, 2.0);
    std::uniform_real_distribution<float> steer_gen(-0.7, 0.7);

    const auto start = std::chrono::steady_clock::now();
    for (uint64_t i = 0; i < num_steps; i++) {
        if (rand_actions) {
            for (int j = 0; j < worlds; j++)
                for (int k = 0; k < world_shape[2 * j]; k++) {       // Manager::setAction, src/mgr.cpp:922-940
                    float *a = &host_actions[((size_t)j * S + k) * AC];
                    a[0] = acc_gen(rand_gen); a[1] = steer_gen(rand_gen); a[2] = 0.f;
                }
            CHECK(gd_step_host(h, host_actions.data(), AC, nullptr, nullptr, nullptr));
        } else {
            CHECK(gd_step(h));
        }
    }
    const auto end = std::chrono::steady_clock::now();
    const std::chrono::duration<double> elapsed = end - start;

    const double fps = (double)num_steps * (double)worlds / elapsed.count();
    printf("FPS %f\n", fps);
    uint64_t total_agents = 0;
    for (int j = 0; j < worlds; j++) total_agents += (uint64_t)world_shape[2 * j];
    // agent-steps per second (the reference multiplies its world-step FPS by the agent total instead,
    // src/headless.cpp:150-166, which counts the world count twice)
    printf("Agent-Normalized FPS %f\n", (double)num_steps * (double)total_agents / elapsed.count());
    GdStats st;
    CHECK(gd_stats(h, &st));
    printf("steps %llu agent_steps %llu launches %llu\n", (unsigned long long)st.steps,
           (unsigned long long)st.agent_steps, (unsigned long long)gd_launch_count(h));
    CHECK(gd_destroy(h));
    return 0;
}
