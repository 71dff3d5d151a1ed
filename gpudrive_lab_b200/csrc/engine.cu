// Host orchestration of the B200 step engine (see engine.h). All kernels live in this TU.
#include "engine.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>

#include "kernels_bev.cuh"
#include "kernels_env.cuh"
#include "kernels_lidar_fix.cuh"
#include "kernels_obs_main.cuh"
#include "kernels_policy.cuh"
#include "kernels_reset.cuh"
#include "kernels_step.cuh"

namespace gd {

#define GD_CUDA(expr)                                                                         \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess)                                                               \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e__) +  \
                                     " at " #expr);                                           \
    } while (0)

#ifndef GD_AGENTS_BLOCK
#define GD_AGENTS_BLOCK 256
#endif
static constexpr int AGENTS_BLOCK = GD_AGENTS_BLOCK;
static constexpr size_t ARENA_SLAB = 256ull << 20;

template <typename T>
static T *devAlloc(std::vector<void *> &owned, size_t count, int fill_byte = 0) {
    void *p = nullptr;
    size_t bytes = std::max<size_t>(count * sizeof(T), 16);
    GD_CUDA(cudaMalloc(&p, bytes));
    GD_CUDA(cudaMemset(p, fill_byte, bytes));
    owned.push_back(p);
    return static_cast<T *>(p);
}

Engine::Engine(int gpu_id, int max_agents, const Params &params, const std::vector<std::string> &scenes, int flags)
    : gpu_(gpu_id), W_((int)scenes.size()), S_(max_agents), bev_enabled_((flags & 1) != 0), params_(params)
{
    if (W_ <= 0) throw std::runtime_error("SimManager needs at least one scene (num_worlds = len(scenes))");
    if (S_ < 2 || S_ > 128) throw std::runtime_error("max_agents must be in [2, 128]");
    if (params_.rewardType == RW_Dense)
        throw std::runtime_error("RewardType.Dense is not implemented by the reference (src/sim.cpp:579-583)");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        throw std::runtime_error("gpudrive_b200: no CUDA device visible; this engine has no CPU fallback");
    if (gpu_ < 0 || gpu_ >= ndev) throw std::runtime_error("gpu_id out of range");
    GD_CUDA(cudaSetDevice(gpu_));
    // Blocking streams, like Madrona's (external/madrona/include/madrona/cuda_utils.inl:62): work the caller
    // enqueued on the legacy default stream - torch's `action_tensor().to_torch()[...] = a` - is ordered before
    // the step, and the caller's later reads on that stream after it. Callers on other non-blocking streams
    // order themselves against gd_stream().
    GD_CUDA(cudaStreamCreate(&stream_));
    GD_CUDA(cudaStreamCreate(&copy_stream_));
    GD_CUDA(cudaEventCreateWithFlags(&ev_agents_, cudaEventDisableTiming));
    GD_CUDA(cudaEventCreateWithFlags(&ev_copied_, cudaEventDisableTiming));
    GD_CUDA(cudaEventCreateWithFlags(&ev_worklist_, cudaEventDisableTiming));
    GD_CUDA(cudaMallocHost(&h_worklist_, sizeof(int32_t) * 2 * (size_t)W_));

    allocExports();

    // largest float key with sqrtf(key) <= observationRadius (exact stand-in for `length() <= radius`)
    {
        const float r = params_.observationRadius;
        float k = r * r;
        while (sqrtf(nextafterf(k, INFINITY)) <= r) k = nextafterf(k, INFINITY);
        while (k > 0.f && sqrtf(k) > r) k = nextafterf(k, -INFINITY);
        key_thr_ = (r >= 0.f) ? k : -1.f;
    }

    scene_paths_ = scenes;
    deleted_.assign(W_, {});
    world_scene_.assign(W_, 0);
    for (int w = 0; w < W_; w++) world_scene_[w] = sceneFor(scenes[w], deleted_[w]);
    GD_CUDA(cudaMemcpy(st_.world_scene, world_scene_.data(), W_ * sizeof(int32_t), cudaMemcpyHostToDevice));

    GD_CUDA(cudaFuncSetAttribute(k_obs<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ObsSmem)));
    GD_CUDA(cudaFuncSetAttribute(k_obs<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ObsSmem)));
    if (bev_enabled_)
        GD_CUDA(cudaFuncSetAttribute(k_bev, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BevSmem)));

    // Sim::Sim -> createPersistentEntities + initWorld for every world, then Manager's ctor reset({})
    std::vector<int32_t> all(W_);
    for (int w = 0; w < W_; w++) all[w] = w;
    launchBindReset(all, true);
    launchRest(false);
    sync();
}

Engine::~Engine() {
    cudaSetDevice(gpu_);
    cudaDeviceSynchronize();
    destroyStepGraph();
    for (auto &e : prof_ev_) cudaEventDestroy(e);
    if (h_worklist_) cudaFreeHost(h_worklist_);
    if (ev_worklist_) cudaEventDestroy(ev_worklist_);
    for (void *p : owned_) cudaFree(p);
    for (void *p : arena_slabs_) cudaFree(p);
    if (d_scenes_) cudaFree(d_scenes_);
    if (d_pool_ids_) cudaFree(d_pool_ids_);
    if (h_pinned_actions_) cudaFreeHost(h_pinned_actions_);
    if (ev_agents_) cudaEventDestroy(ev_agents_);
    if (ev_copied_) cudaEventDestroy(ev_copied_);
    if (copy_stream_) cudaStreamDestroy(copy_stream_);
    if (stream_) cudaStreamDestroy(stream_);
}

void Engine::allocExports() {
    const size_t WS = (size_t)W_ * S_;
    ex_.action = devAlloc<float>(owned_, WS * kActionSize);
    ex_.reward = devAlloc<float>(owned_, WS);
    ex_.done = devAlloc<int32_t>(owned_, WS);
    ex_.info = devAlloc<int32_t>(owned_, WS * kInfoSize);
    ex_.self_obs = devAlloc<float>(owned_, WS * kSelfObsSize);
    ex_.abs_obs = devAlloc<float>(owned_, WS * kAbsObsSize);
    ex_.partner_obs = devAlloc<float>(owned_, WS * (S_ - 1) * kPartnerObsSize);
    ex_.agent_roadmap = devAlloc<float>(owned_, WS * kMaxAgentMapObs * kMapObsSize);
    ex_.map_obs = devAlloc<float>(owned_, (size_t)W_ * kMaxRoadEntityCount * kMapObsSize);
    ex_.lidar = devAlloc<float>(owned_, WS * 3 * kNumLidarSamples * kLidarRowSize);
    ex_.steps_remaining = devAlloc<int32_t>(owned_, WS);
    ex_.shape = devAlloc<int32_t>(owned_, (size_t)W_ * 2);
    ex_.controlled = devAlloc<int32_t>(owned_, WS);
    ex_.response = devAlloc<int32_t>(owned_, WS);
    ex_.traj = devAlloc<float>(owned_, WS * kTrajSize);
    ex_.world_means = devAlloc<float>(owned_, (size_t)W_ * 3);
    ex_.metadata = devAlloc<int32_t>(owned_, WS * kMetaSize);
    ex_.deleted = devAlloc<int32_t>(owned_, WS, 0xff);    // Sim::Sim fills -1 (src/sim.cpp:1003-1006)
    ex_.map_name = devAlloc<int32_t>(owned_, (size_t)W_ * 32);
    ex_.scenario_id = devAlloc<int32_t>(owned_, (size_t)W_ * 32);
    ex_.valid_state = devAlloc<int32_t>(owned_, WS);
    ex_.bev = bev_enabled_ ? devAlloc<float>(owned_, WS * (size_t)(BEV_RES * BEV_RES)) : nullptr;
    ex_.knn_idx = devAlloc<int32_t>(owned_, WS * kMaxAgentMapObs, 0xff);
    st_.pos = devAlloc<float4>(owned_, WS);
    st_.rv = devAlloc<float4>(owned_, WS);
    st_.yaw = devAlloc<float>(owned_, WS);
    st_.cost = devAlloc<uint32_t>(owned_, WS);
    st_.order = devAlloc<uint8_t>(owned_, WS);
    {
        std::vector<uint8_t> ident(WS);
        for (size_t i = 0; i < WS; i++) ident[i] = (uint8_t)(i % (size_t)S_);
        GD_CUDA(cudaMemcpy(st_.order, ident.data(), WS, cudaMemcpyHostToDevice));
    }
    st_.collided = devAlloc<int32_t>(owned_, WS);
    st_.world_scene = devAlloc<int32_t>(owned_, W_);
    st_.reset_flag = devAlloc<int32_t>(owned_, W_);
    st_.stats = devAlloc<unsigned long long>(owned_, 16);
    st_.retrace = devAlloc<uint32_t>(owned_, 1 + (size_t)kRetraceCap);
    d_worklist_ = devAlloc<int32_t>(owned_, 2 * (size_t)W_);
    d_act_stage_ = devAlloc<float>(owned_, (size_t)W_ * S_ * kActionSize);
}

void *Engine::arenaAlloc(size_t bytes, size_t *block_bytes) {
    bytes = (bytes + 255) & ~size_t(255);
    // a block an evicted scene left behind, if one fits without wasting more than half of it (best fit)
    int best = -1;
    for (int i = 0; i < (int)arena_free_.size(); i++)
        if (arena_free_[i].second >= bytes && arena_free_[i].second <= 2 * bytes &&
            (best < 0 || arena_free_[i].second < arena_free_[best].second)) best = i;
    if (best >= 0) {
        void *out = arena_free_[best].first;
        *block_bytes = arena_free_[best].second;
        arena_free_.erase(arena_free_.begin() + best);
        return out;
    }
    *block_bytes = bytes;
    if (arena_slabs_.empty() || arena_off_ + bytes > arena_cap_) {
        size_t cap = std::max(ARENA_SLAB, bytes);
        void *p = nullptr;
        GD_CUDA(cudaMalloc(&p, cap));
        arena_slabs_.push_back(p);
        arena_off_ = 0;
        arena_cap_ = cap;
    }
    void *out = static_cast<char *>(arena_slabs_.back()) + arena_off_;
    arena_off_ += bytes;
    return out;
}

// Pack one SceneData into a contiguous blob, upload it into the pool arena, register its SceneDev.
int Engine::uploadScene(const SceneData &s, const std::string &key) {
    const int R = s.numRoads, A = s.numAgents;
    const int Rpad = (R + 31) & ~31;
    // collidable subset: road edges + stop signs (everything else is in Sim::collisionPairs' ignore list
    // for every agent type, src/sim.hpp:88-102)
    std::vector<int> coll;
    for (int r = 0; r < R; r++)
        if (s.rtype[r] == ET_RoadEdge || s.rtype[r] == ET_StopSign) coll.push_back(r);
    {   // spatial order (Morton code of the 1 m quantised position): k_agents tests 32 consecutive boxes per
        // bounding circle; collision results are flags, so the order of the table is free
        float minx = 3e38f, miny = 3e38f;
        for (int r : coll) { minx = std::min(minx, s.rx[r]); miny = std::min(miny, s.ry[r]); }
        auto spread = [](uint32_t v) {
            uint64_t x = v & 0xffffu;
            x = (x | (x << 8)) & 0x00ff00ffull; x = (x | (x << 4)) & 0x0f0f0f0full;
            x = (x | (x << 2)) & 0x33333333ull; x = (x | (x << 1)) & 0x55555555ull;
            return x;
        };
        std::vector<std::pair<uint64_t, int>> keys;
        for (int r : coll) {
            const uint32_t qx = (uint32_t)std::min(65535.f, std::max(0.f, s.rx[r] - minx));
            const uint32_t qy = (uint32_t)std::min(65535.f, std::max(0.f, s.ry[r] - miny));
            keys.push_back({spread(qx) | (spread(qy) << 1), r});
        }
        std::sort(keys.begin(), keys.end());
        for (size_t i = 0; i < keys.size(); i++) coll[i] = keys[i].second;
    }
    const int C = (int)coll.size(), Cpad = (C + 31) & ~31, CG = Cpad / 32;

    SceneDev dev;
    memset(&dev, 0, sizeof(dev));
    // z classes
    std::vector<uint8_t> zcls(R, 0);
    for (int r = 0; r < R; r++) {
        int c = -1;
        for (int k = 0; k < dev.numZClasses; k++)
            if (dev.zc_z[k] == s.rz[r] && dev.zc_d2[k] == s.rd2[r]) { c = k; break; }
        if (c < 0) {
            if (dev.numZClasses >= kMaxZClasses) throw std::runtime_error("too many distinct road z-classes");
            c = dev.numZClasses++;
            dev.zc_z[c] = s.rz[r];
            dev.zc_d2[c] = s.rd2[r];
        }
        zcls[r] = (uint8_t)c;
    }

    // spatial order for k_obs: by z-class, then Morton code of the (1 m quantised) position; classes are
    // padded to whole groups of 32 so a group never mixes lidar planes
    std::vector<int> sorted;       // sorted slot -> original road index, -1 = dummy
    {
        float minx = 3e38f, miny = 3e38f;
        for (int r = 0; r < R; r++) { minx = std::min(minx, s.rx[r]); miny = std::min(miny, s.ry[r]); }
        auto spread = [](uint32_t v) {
            uint64_t x = v & 0xffffu;
            x = (x | (x << 8)) & 0x00ff00ffull; x = (x | (x << 4)) & 0x0f0f0f0full;
            x = (x | (x << 2)) & 0x33333333ull; x = (x | (x << 1)) & 0x55555555ull;
            return x;
        };
        std::vector<std::pair<uint64_t, int>> keys(R);
        for (int r = 0; r < R; r++) {
            const uint32_t qx = (uint32_t)std::min(65535.f, std::max(0.f, s.rx[r] - minx));
            const uint32_t qy = (uint32_t)std::min(65535.f, std::max(0.f, s.ry[r] - miny));
            keys[r] = {((uint64_t)zcls[r] << 32) | spread(qx) | (spread(qy) << 1), r};
        }
        std::sort(keys.begin(), keys.end());
        for (int i = 0; i < R; i++) {
            if (i > 0 && (keys[i].first >> 32) != (keys[i - 1].first >> 32))
                while (sorted.size() % 32) sorted.push_back(-1);
            sorted.push_back(keys[i].second);
        }
        while (sorted.size() % 32) sorted.push_back(-1);
    }
    const int RS = (int)sorted.size(), G = RS / 32;

    size_t off = 0;
    auto section = [&](size_t bytes) { size_t o = off; off += (bytes + 15) & ~size_t(15); return o; };
    const size_t o_sa = section(sizeof(float4) * RS), o_sb = section(sizeof(float2) * RS),
                 o_sc = section(sizeof(uint32_t) * RS), o_sd = section(sizeof(float4) * RS),
                 o_gr = section(sizeof(float4) * 2 * std::max(G, 1));
    const size_t o_ra = section(sizeof(float4) * Rpad), o_rb = section(sizeof(float2) * Rpad),
                 o_rc = section(sizeof(uint32_t) * Rpad), o_rd = section(sizeof(float4) * Rpad),
                 o_cc = section(sizeof(float4) * Cpad), o_co = section(sizeof(float4) * 4 * Cpad),
                 o_cg = section(sizeof(float4) * std::max(CG, 1)),
                 o_tr = section(sizeof(float) * (size_t)A * kTrajSize), o_aa = section(sizeof(float4) * A),
                 o_ab = section(sizeof(float4) * A), o_ac = section(sizeof(float4) * A),
                 o_me = section(sizeof(int4) * A);
    std::vector<unsigned char> blob(off, 0);
    auto F4 = [&](size_t o) { return reinterpret_cast<float4 *>(blob.data() + o); };
    float4 *ra = F4(o_ra), *rd = F4(o_rd), *cc = F4(o_cc), *co = F4(o_co), *aa = F4(o_aa), *ab = F4(o_ab), *ac = F4(o_ac);
    float2 *rb = reinterpret_cast<float2 *>(blob.data() + o_rb);
    uint32_t *rc = reinterpret_cast<uint32_t *>(blob.data() + o_rc);
    int4 *me = reinterpret_cast<int4 *>(blob.data() + o_me);
    for (int r = 0; r < R; r++) {
        ra[r] = make_float4(s.rx[r], s.ry[r], s.rqw[r], s.rqz[r]);
        rb[r] = make_float2(s.rd0[r], s.rd1[r]);
        rc[r] = (uint32_t)zcls[r] | ((uint32_t)s.rtype[r] << 8);
        rd[r] = make_float4(s.rd2[r], s.ryaw[r], s.rid[r], (float)s.rmaptype[r]);
    }
    for (int r = R; r < Rpad; r++) {   // padding far away, never in any radius
        ra[r] = make_float4(1e30f, 1e30f, 1.f, 0.f);
        rb[r] = make_float2(1.f, 1.f);
    }
    {
        float4 *sa = F4(o_sa), *sd = F4(o_sd), *gr = F4(o_gr);
        float2 *sb = reinterpret_cast<float2 *>(blob.data() + o_sb);
        uint32_t *scc = reinterpret_cast<uint32_t *>(blob.data() + o_sc);
        for (int i = 0; i < RS; i++) {
            const int r = sorted[i];
            if (r < 0) {     // dummy: never in a radius, never in lidar range, z-class 8 (no plane)
                sa[i] = make_float4(1e30f, 1e30f, 1.f, 0.f); sb[i] = make_float2(1.f, 1.f);
                scc[i] = 8u; sd[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                sa[i] = ra[r]; sb[i] = rb[r]; sd[i] = rd[r];
                scc[i] = rc[r] | ((uint32_t)r << 16);
            }
        }
        for (int g = 0; g < G; g++) {
            // bounds of the group's boxes: an oriented rectangle along the principal axis of their corners
            // (polyline runs are long and thin) for distances, and the circle around its centre for angles
            std::vector<float> px, py;
            uint32_t cmask = 0;
            for (int k = 0; k < 32; k++) {
                const int r = sorted[g * 32 + k];
                if (r < 0) continue;
                cmask |= 1u << zcls[r];
                const float c = 1.f - 2.f * s.rqz[r] * s.rqz[r], sn = 2.f * s.rqw[r] * s.rqz[r];
                for (int q = 0; q < 4; q++) {
                    const float ex = (q & 1) ? s.rd0[r] : -s.rd0[r], ey = (q & 2) ? s.rd1[r] : -s.rd1[r];
                    px.push_back(s.rx[r] + ex * c - ey * sn);
                    py.push_back(s.ry[r] + ex * sn + ey * c);
                }
            }
            float4 circ = make_float4(1e30f, 1e30f, 0.f, 0.f), rect = make_float4(1.f, 0.f, 0.f, 0.f);
            if (!px.empty()) {
                double mx = 0, my = 0;
                for (size_t i = 0; i < px.size(); i++) { mx += px[i]; my += py[i]; }
                mx /= px.size(); my /= px.size();
                double sxx = 0, sxy = 0, syy = 0;
                for (size_t i = 0; i < px.size(); i++) {
                    const double dx = px[i] - mx, dy = py[i] - my;
                    sxx += dx * dx; sxy += dx * dy; syy += dy * dy;
                }
                const double th = 0.5 * atan2(2 * sxy, sxx - syy);
                const float ux = (float)cos(th), uy = (float)sin(th);
                float ulo = 3e38f, uhi = -3e38f, vlo = 3e38f, vhi = -3e38f;
                for (size_t i = 0; i < px.size(); i++) {
                    const float u = px[i] * ux + py[i] * uy, v = py[i] * ux - px[i] * uy;
                    ulo = std::min(ulo, u); uhi = std::max(uhi, u); vlo = std::min(vlo, v); vhi = std::max(vhi, v);
                }
                const float cu = 0.5f * (ulo + uhi), cv = 0.5f * (vlo + vhi);
                const float cx = cu * ux - cv * uy, cy = cu * uy + cv * ux;
                // margins: the rectangle is evaluated in float on coordinates of a few thousand metres
                const float ha = 0.5f * (uhi - ulo) * 1.0001f + 5e-3f, hb = 0.5f * (vhi - vlo) * 1.0001f + 5e-3f;
                float rad = 0.f;
                for (size_t i = 0; i < px.size(); i++) {
                    const float dx = px[i] - cx, dy = py[i] - cy;
                    rad = std::max(rad, sqrtf(dx * dx + dy * dy));
                }
                circ = make_float4(cx, cy, rad * 1.0001f + 5e-3f, 0.f);
                rect = make_float4(ux, uy, ha, hb);
            }
            memcpy(&circ.w, &cmask, 4);
            gr[2 * g] = circ;
            gr[2 * g + 1] = rect;
        }
    }
    for (int i = 0; i < C; i++) {
        const int r = coll[i];
        // OrientedBoundingBox2D::from on the host: same libm calls as the reference's CPU build
        const float theta = s.ryaw[r];
        const float cth = cosf(theta), sth = sinf(theta);
        const float Xx = cth * s.rd0[r], Xy = sth * s.rd0[r];
        const float Yx = -sth * s.rd1[r], Yy = cth * s.rd1[r];
        const float px = s.rx[r], py = s.ry[r];
        float cx[4], cy[4];
        cx[0] = (px - Xx) - Yx; cy[0] = (py - Xy) - Yy;
        cx[1] = (px + Xx) - Yx; cy[1] = (py + Xy) - Yy;
        cx[2] = (px + Xx) + Yx; cy[2] = (py + Xy) + Yy;
        cx[3] = (px - Xx) + Yx; cy[3] = (py - Xy) + Yy;
        float a0x = cx[1] - cx[0], a0y = cy[1] - cy[0], a1x = cx[3] - cx[0], a1y = cy[3] - cy[0];
        float inv0 = 1.f / (a0x * a0x + a0y * a0y), inv1 = 1.f / (a1x * a1x + a1y * a1y);
        a0x *= inv0; a0y *= inv0; a1x *= inv1; a1y *= inv1;
        const float org0 = cx[0] * a0x + cy[0] * a0y, org1 = cx[0] * a1x + cy[0] * a1y;
        co[i * 4 + 0] = make_float4(cx[0], cy[0], cx[1], cy[1]);
        co[i * 4 + 1] = make_float4(cx[2], cy[2], cx[3], cy[3]);
        co[i * 4 + 2] = make_float4(a0x, a0y, a1x, a1y);
        co[i * 4 + 3] = make_float4(org0, org1, s.rz[r], s.rd2[r]);
        const float rad = sqrtf(s.rd0[r] * s.rd0[r] + s.rd1[r] * s.rd1[r]) * 1.0001f + 1e-3f;
        cc[i] = make_float4(px, py, rad, (float)s.rtype[r]);
    }
    for (int i = C; i < Cpad; i++) cc[i] = make_float4(1e30f, 1e30f, 0.f, 0.f);
    {   // one circle per 32 consecutive collidable boxes, around their bounding circles
        float4 *cg = F4(o_cg);
        for (int g = 0; g < CG; g++) {
            float lox = 3e38f, loy = 3e38f, hix = -3e38f, hiy = -3e38f;
            const int n = std::min(32, C - g * 32);
            for (int k = 0; k < n; k++) {
                const float4 c = cc[g * 32 + k];
                lox = std::min(lox, c.x); hix = std::max(hix, c.x); loy = std::min(loy, c.y); hiy = std::max(hiy, c.y);
            }
            const float gx = 0.5f * (lox + hix), gy = 0.5f * (loy + hiy);
            float rad = 0.f;
            for (int k = 0; k < n; k++) {
                const float4 c = cc[g * 32 + k];
                rad = std::max(rad, sqrtf((c.x - gx) * (c.x - gx) + (c.y - gy) * (c.y - gy)) + c.z);
            }
            cg[g] = n > 0 ? make_float4(gx, gy, rad * 1.0001f + 1e-3f, 0.f) : make_float4(1e30f, 1e30f, 0.f, 0.f);
        }
    }
    if (A) memcpy(blob.data() + o_tr, s.traj.data(), sizeof(float) * (size_t)A * kTrajSize);
    for (int a = 0; a < A; a++) {
        aa[a] = make_float4(s.vsize[a * 3], s.vsize[a * 3 + 1], s.vsize[a * 3 + 2], (float)s.id[a]);
        ab[a] = make_float4(s.scale[a * 3], s.scale[a * 3 + 1], s.scale[a * 3 + 2], (float)s.type[a]);
        ac[a] = make_float4(s.goal[a * 2], s.goal[a * 2 + 1], (float)s.response[a], (float)s.controlled[a]);
        me[a] = make_int4(s.meta[a * 4], s.meta[a * 4 + 1], s.meta[a * 4 + 2], s.meta[a * 4 + 3]);
    }
    size_t block_bytes = 0;
    char *base = static_cast<char *>(arenaAlloc(blob.size(), &block_bytes));
    GD_CUDA(cudaMemcpy(base, blob.data(), blob.size(), cudaMemcpyHostToDevice));
    dev.road_a = reinterpret_cast<const float4 *>(base + o_ra);
    dev.road_b = reinterpret_cast<const float2 *>(base + o_rb);
    dev.road_c = reinterpret_cast<const uint32_t *>(base + o_rc);
    dev.road_d = reinterpret_cast<const float4 *>(base + o_rd);
    dev.s_a = reinterpret_cast<const float4 *>(base + o_sa);
    dev.s_b = reinterpret_cast<const float2 *>(base + o_sb);
    dev.s_c = reinterpret_cast<const uint32_t *>(base + o_sc);
    dev.s_d = reinterpret_cast<const float4 *>(base + o_sd);
    dev.grp = reinterpret_cast<const float4 *>(base + o_gr);
    dev.numGroups = G;
    dev.coll_cull = reinterpret_cast<const float4 *>(base + o_cc);
    dev.coll_obb = reinterpret_cast<const float4 *>(base + o_co);
    dev.coll_grp = reinterpret_cast<const float4 *>(base + o_cg);
    dev.numCollGroups = CG;
    dev.traj = reinterpret_cast<const float *>(base + o_tr);
    dev.ag_a = reinterpret_cast<const float4 *>(base + o_aa);
    dev.ag_b = reinterpret_cast<const float4 *>(base + o_ab);
    dev.ag_c = reinterpret_cast<const float4 *>(base + o_ac);
    dev.meta = reinterpret_cast<const int4 *>(base + o_me);
    dev.numRoads = R; dev.numAgents = A; dev.numColl = C;
    dev.mean[0] = s.mean[0]; dev.mean[1] = s.mean[1]; dev.mean[2] = s.mean[2];
    memcpy(dev.mapName, s.mapName, sizeof(dev.mapName));
    memcpy(dev.scenarioId, s.scenarioId, sizeof(dev.scenarioId));

    SceneHost h;
    h.key = key; h.dev = dev; h.numAgents = A; h.numRoads = R; h.numControlled = s.numControlled;
    h.block = base; h.block_bytes = block_bytes; h.last_use = bind_generation_;
    int id;
    if (!free_ids_.empty()) { id = free_ids_.back(); free_ids_.pop_back(); scenes_[id] = h; }
    else { scenes_.push_back(h); id = (int)scenes_.size() - 1; }
    scene_index_[key] = id;
    max_roads_ = std::max(max_roads_, R);
    max_scene_agents_ = std::max(max_scene_agents_, A);
    scene_table_dirty_ = true;
    return id;
}

// Bound the scenario cache (an env that resamples with set_maps over a large dataset would otherwise grow it by
// ~0.4 MB per scenario for ever): scenes no world is bound to and whose id no caller holds are kept up to a budget
// of max(W, 64), oldest use first out; their arena blocks and ids are reused by later uploads. Called with the
// stream idle (set_maps / deleteAgents are synchronous calls), so no kernel in flight reads an evicted table.
void Engine::evictUnusedScenes() {
    std::vector<char> bound(scenes_.size(), 0);
    for (int w = 0; w < W_; w++) bound[world_scene_[w]] = 1;
    std::vector<int> idle;
    for (int i = 0; i < (int)scenes_.size(); i++)
        if (scenes_[i].live && !scenes_[i].pinned && !bound[i]) idle.push_back(i);
    const size_t keep = (size_t)std::max(W_, 64);
    if (idle.size() <= keep) return;
    std::sort(idle.begin(), idle.end(), [&](int a, int b) { return scenes_[a].last_use < scenes_[b].last_use; });
    for (size_t k = 0; k + keep < idle.size(); k++) {
        SceneHost &h = scenes_[idle[k]];
        scene_index_.erase(h.key);
        arena_free_.emplace_back(h.block, h.block_bytes);
        h.live = false; h.block = nullptr; h.block_bytes = 0; h.key.clear();
        h.dev = SceneDev{};
        free_ids_.push_back(idle[k]);
    }
    scene_table_dirty_ = true;
}

int Engine::sceneFor(const std::string &path, const std::vector<int32_t> &deleted) {
    std::string key = path;
    if (!deleted.empty()) {
        key += "#del";
        for (int32_t d : deleted) key += ":" + std::to_string(d);
    }
    auto it = scene_index_.find(key);
    if (it != scene_index_.end()) { scenes_[it->second].last_use = bind_generation_; return it->second; }
    std::shared_ptr<ParsedMap> pm;
    auto pit = parsed_cache_.find(path);
    if (pit != parsed_cache_.end()) pm = pit->second;
    else {
        pm = std::make_shared<ParsedMap>(loadParsedMap(path, params_.polylineReductionThreshold));
        if (parsed_cache_.size() < 64) parsed_cache_[path] = pm;   // small LRU-less cache for deleteAgents
    }
    SceneData s = buildScene(*pm, params_, S_, deleted.data(), (int)deleted.size());
    return uploadScene(s, key);
}

void Engine::syncSceneTable() {
    if (!scene_table_dirty_) return;
    if (scenes_.size() > d_scenes_cap_) {
        GD_CUDA(cudaStreamSynchronize(stream_));
        if (d_scenes_) GD_CUDA(cudaFree(d_scenes_));
        d_scenes_cap_ = std::max<size_t>(scenes_.size() * 2, 64);
        GD_CUDA(cudaMalloc(&d_scenes_, d_scenes_cap_ * sizeof(SceneDev)));
    }
    std::vector<SceneDev> tmp(scenes_.size());
    for (size_t i = 0; i < scenes_.size(); i++) tmp[i] = scenes_[i].dev;
    GD_CUDA(cudaMemcpy(d_scenes_, tmp.data(), tmp.size() * sizeof(SceneDev), cudaMemcpyHostToDevice));
    scene_table_dirty_ = false;
}

void Engine::launchBindReset(const std::vector<int32_t> &worlds, bool rebind, bool scene_ids_uploaded) {
    if (worlds.empty()) return;
    syncSceneTable();
    const int n = (int)worlds.size();
    // pinned staging: wait only until the previous list has left the buffer, not for the device to drain
    GD_CUDA(cudaEventSynchronize(ev_worklist_));
    memcpy(h_worklist_, worlds.data(), n * sizeof(int32_t));
    GD_CUDA(cudaMemcpyAsync(d_worklist_, h_worklist_, n * sizeof(int32_t), cudaMemcpyHostToDevice, stream_));
    GD_CUDA(cudaEventRecord(ev_worklist_, stream_));
    if (scene_ids_uploaded) {
        k_set_scene<<<(n + 255) / 256, 256, 0, stream_>>>(d_worklist_, d_worklist_ + W_, n, st_.world_scene);
        launches_++;
    }
    if (rebind) {
        k_bind<<<dim3(n, BIND_SPLIT), 256, 0, stream_>>>(d_worklist_, d_scenes_, st_, ex_, S_, fusedPolicy());
        launches_++;
    }
    const int blk = ((S_ + 31) / 32) * 32;
    k_reset<<<n, blk, 0, stream_>>>(d_worklist_, d_scenes_, st_, ex_, S_, params_.dynamicsModel);
    launches_++;
    GD_CUDA(cudaGetLastError());
    episodes_reset_ += n;
}

int Engine::obsGroups() const {
    // CTAs only for slots some pooled scene populates (padding slots keep their bind-time rows)
    return std::max(1, (std::min(S_, max_scene_agents_) + OBS_WARPS - 1) / OBS_WARPS);
}

void Engine::destroyStepGraph() {
    if (graph_exec_) cudaGraphExecDestroy(graph_exec_);
    if (graph_) cudaGraphDestroy(graph_);
    graph_exec_ = nullptr; graph_ = nullptr;
    graph_key_ = StepGraphKey();
}

// One step as one graph launch (the reference's GPU backend launches its step graph the same way,
// external/madrona/src/mw/cuda_exec.cpp:1699,1891): [ev0] k_agents [ev1 | ev_agents] k_obs [ev2] (k_bev).
// Built node by node rather than by stream capture, so nothing the caller does on other streams meanwhile can
// invalidate it. The profiling events are re-pointed at the ring slot before every launch.
void Engine::buildStepGraph(const StepGraphKey &key) {
    destroyStepGraph();
    GD_CUDA(cudaGraphCreate(&graph_, 0));
    cudaGraphNode_t last = nullptr;
    auto dep = [&]() { return last ? &last : nullptr; };
    auto ndep = [&]() { return last ? (size_t)1 : (size_t)0; };
    auto add_event = [&](cudaEvent_t ev, cudaGraphNode_t *keep) {
        cudaGraphNode_t n;
        GD_CUDA(cudaGraphAddEventRecordNode(&n, graph_, dep(), ndep(), ev));
        if (keep) *keep = n;
        last = n;
    };
    auto add_kernel = [&](void *fn, dim3 grid, dim3 block, size_t smem, void **args) {
        cudaKernelNodeParams kp = {};
        kp.func = fn; kp.gridDim = grid; kp.blockDim = block; kp.sharedMemBytes = (unsigned)smem;
        kp.kernelParams = args; kp.extra = nullptr;
        cudaGraphNode_t n;
        GD_CUDA(cudaGraphAddKernelNode(&n, graph_, dep(), ndep(), &kp));
        last = n;
    };
    const SceneDev *scenes = d_scenes_;
    int S = S_, one = 1, zero = 0, groups = key.groups, knn = key.knn_idx;
    float key_thr = key_thr_;
    float *pol = fusedPolicy();
    if (key.profile) add_event(prof_ev_[0], &graph_ev_nodes_[0]);
    {
        void *args[] = {&scenes, &st_, &ex_, &params_, &S, &one, &one, &zero};
        add_kernel((void *)k_agents<AGENTS_BLOCK>, dim3(W_), dim3(AGENTS_BLOCK), sizeof(AgentBox) * (size_t)S_, args);
    }
    if (key.profile) add_event(prof_ev_[1], &graph_ev_nodes_[1]);
    add_event(ev_agents_, nullptr);         // reward / done / info are final here (stepHost's copy stream waits on it)
    {
        void *args[] = {&scenes, &st_, &ex_, &params_, &S, &groups, &knn, &key_thr, &zero, &pol};
        add_kernel(pol ? (void *)k_obs<true> : (void *)k_obs<false>, dim3(W_ * groups), dim3(OBS_BLOCK), sizeof(ObsSmem), args);
    }
    if (params_.enableLidar) {      // the few rays k_obs could not settle, re-traced exactly (counted with k_obs' time)
        void *args[] = {&scenes, &st_, &ex_, &S};
        add_kernel((void *)k_lidar_fix, dim3(FIX_GRID), dim3(FIX_BLOCK), 0, args);
    }
    if (key.profile) add_event(prof_ev_[2], &graph_ev_nodes_[2]);
    if (bev_enabled_) {
        void *args[] = {&scenes, &st_, &ex_, &params_, &S, &key_thr};
        add_kernel((void *)k_bev, dim3(W_ * S_), dim3(BEV_BLOCK), sizeof(BevSmem), args);
    }
    GD_CUDA(cudaGraphInstantiate(&graph_exec_, graph_, 0));
    graph_key_ = key;
}

// setupRestOfTasks (src/sim.cpp:785-943): collisions, reward, (--t), done, observation fan-out
void Engine::launchRest(bool decrement, bool only_flagged) {
    syncSceneTable();
    const int flagged = only_flagged ? 1 : 0;
    const size_t smem_agents = sizeof(AgentBox) * (size_t)S_;
    const bool prof = opt_profile_ && decrement;
    const int groups = obsGroups();
    kernels_per_step_ = 2 + (bev_enabled_ ? 1 : 0) + (params_.enableLidar ? 1 : 0);
    cudaEvent_t *ev = nullptr;
    if (prof) {
        if (prof_pending_ == kProfRing) foldProfile();
        ev = &prof_ev_[3 * prof_pending_];
    }
    if (decrement && opt_graph_) {
        StepGraphKey key;
        key.scenes = d_scenes_; key.groups = groups; key.knn_idx = opt_knn_idx_; key.profile = prof ? 1 : 0;
        key.pol = fusedPolicy();
        if (!graph_exec_ || !(key == graph_key_)) buildStepGraph(key);
        if (prof) {
            for (int i = 0; i < 3; i++)
                GD_CUDA(cudaGraphExecEventRecordNodeSetEvent(graph_exec_, graph_ev_nodes_[i], ev[i]));
            prof_pending_++;
        }
        GD_CUDA(cudaGraphLaunch(graph_exec_, stream_));
        launches_ += kernels_per_step_;
        return;
    }
    if (prof) GD_CUDA(cudaEventRecord(ev[0], stream_));
    k_agents<AGENTS_BLOCK><<<W_, AGENTS_BLOCK, smem_agents, stream_>>>(d_scenes_, st_, ex_, params_, S_,
                                                                    decrement ? 1 : 0, decrement ? 1 : 0, flagged);
    if (prof) GD_CUDA(cudaEventRecord(ev[1], stream_));
    if (mark_agents_done_) GD_CUDA(cudaEventRecord(ev_agents_, stream_));   // reward / done / info are final here
    if (fusedPolicy())
        k_obs<true><<<W_ * groups, OBS_BLOCK, sizeof(ObsSmem), stream_>>>(d_scenes_, st_, ex_, params_, S_, groups,
                                                                          opt_knn_idx_, key_thr_, flagged, fusedPolicy());
    else
        k_obs<false><<<W_ * groups, OBS_BLOCK, sizeof(ObsSmem), stream_>>>(d_scenes_, st_, ex_, params_, S_, groups,
                                                                           opt_knn_idx_, key_thr_, flagged, nullptr);
    if (params_.enableLidar) k_lidar_fix<<<FIX_GRID, FIX_BLOCK, 0, stream_>>>(d_scenes_, st_, ex_, S_);
    if (prof) { GD_CUDA(cudaEventRecord(ev[2], stream_)); prof_pending_++; }
    if (bev_enabled_)
        k_bev<<<W_ * S_, BEV_BLOCK, sizeof(BevSmem), stream_>>>(d_scenes_, st_, ex_, params_, S_, key_thr_);
    launches_ += kernels_per_step_;
    GD_CUDA(cudaGetLastError());
}

void Engine::stepAsync() {
    GD_CUDA(cudaSetDevice(gpu_));
    launchRest(true);
    steps_++;
}

void Engine::sync() { GD_CUDA(cudaStreamSynchronize(stream_)); }

void Engine::step() {
    stepAsync();
    sync();
}

void Engine::reset(const int32_t *worlds, int n) {
    GD_CUDA(cudaSetDevice(gpu_));
    // (the engine stream is a blocking one: tensor writes the caller made on the legacy default stream are
    //  already ordered before the kernels below, no device-wide synchronisation needed)
    std::vector<int32_t> list;
    for (int i = 0; i < n; i++) {
        if (worlds[i] < 0 || worlds[i] >= W_) throw std::runtime_error("reset: world index out of range");
        list.push_back(worlds[i]);
    }
    std::sort(list.begin(), list.end());
    list.erase(std::unique(list.begin(), list.end()), list.end());
    launchBindReset(list, false);
    launchRest(false);     // the Reset graph recomputes every world (src/sim.cpp:960-966)
    sync();
}

void Engine::setMaps(const std::vector<std::string> &scenes) {
    if ((int)scenes.size() != W_) throw std::runtime_error("set_maps: len(maps) must equal num_worlds");
    GD_CUDA(cudaSetDevice(gpu_));
    host_binding_stale_ = false;
    scene_paths_ = scenes;
    GD_CUDA(cudaStreamSynchronize(stream_));         // blocks of evicted scenes are reused below: nothing may still read them
    bind_generation_++;
    for (int w = 0; w < W_; w++) {
        deleted_[w].clear();                         // src/mgr.cpp:643-646
        world_scene_[w] = sceneFor(scenes[w], deleted_[w]);
    }
    evictUnusedScenes();
    GD_CUDA(cudaMemset(ex_.deleted, 0xff, (size_t)W_ * S_ * sizeof(int32_t)));
    GD_CUDA(cudaMemcpy(st_.world_scene, world_scene_.data(), W_ * sizeof(int32_t), cudaMemcpyHostToDevice));
    std::vector<int32_t> all(W_);
    for (int w = 0; w < W_; w++) all[w] = w;
    launchBindReset(all, true);
    launchRest(false);
    sync();
}

void Engine::deleteAgents(int world, const int32_t *ids, int n) {
    if (world < 0 || world >= W_) throw std::runtime_error("deleteAgents: world index out of range");
    if (n > S_) throw std::runtime_error("deleteAgents: more ids than agent slots");
    GD_CUDA(cudaSetDevice(gpu_));
    refreshHostBinding();
    // src/mgr.cpp:698-709: ids overwrite the head of the world's DeletedAgents list
    std::vector<int32_t> cur(S_);
    GD_CUDA(cudaMemcpy(cur.data(), ex_.deleted + (size_t)world * S_, S_ * sizeof(int32_t), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; i++) cur[i] = ids[i];
    GD_CUDA(cudaMemcpy(ex_.deleted + (size_t)world * S_, cur.data(), S_ * sizeof(int32_t), cudaMemcpyHostToDevice));
    deleted_[world].clear();
    for (int i = 0; i < S_; i++)
        if (cur[i] != -1) deleted_[world].push_back(cur[i]);
    world_scene_[world] = sceneFor(scene_paths_[world], deleted_[world]);
    GD_CUDA(cudaMemcpy(st_.world_scene + world, &world_scene_[world], sizeof(int32_t), cudaMemcpyHostToDevice));
    // the reference then resets every world; only `world` has ResetMap set (src/mgr.cpp:707-714)
    launchBindReset({world}, true);
    std::vector<int32_t> others;
    for (int w = 0; w < W_; w++)
        if (w != world) others.push_back(w);
    launchBindReset(others, false);
    launchRest(false);
    sync();
}

int Engine::poolAdd(const std::string &path) {
    GD_CUDA(cudaSetDevice(gpu_));
    GD_CUDA(cudaStreamSynchronize(stream_));         // (an upload may reuse the block of an evicted scene)
    const int id = sceneFor(path, {});
    scenes_[id].pinned = true;                       // the caller holds the id
    return id;
}

void Engine::bindWorlds(const int32_t *worlds, const int32_t *scene_ids, int n) {
    GD_CUDA(cudaSetDevice(gpu_));
    refreshHostBinding();
    std::vector<int32_t> list;
    for (int i = 0; i < n; i++) {
        if (worlds[i] < 0 || worlds[i] >= W_) throw std::runtime_error("bind: world index out of range");
        if (scene_ids[i] < 0 || scene_ids[i] >= (int)scenes_.size() || !scenes_[scene_ids[i]].live)
            throw std::runtime_error("bind: scene id out of range");
    }
    if (n > W_) throw std::runtime_error("bind: more entries than worlds");
    for (int i = 0; i < n; i++) {
        world_scene_[worlds[i]] = scene_ids[i];
        scenes_[scene_ids[i]].pinned = true;
        scene_paths_[worlds[i]] = scenes_[scene_ids[i]].key;
        list.push_back(worlds[i]);
    }
    if (n > 0) {
        // the new binding travels with the work list (pinned, stream-ordered): no synchronous copy, no host stall
        GD_CUDA(cudaEventSynchronize(ev_worklist_));
        memcpy(h_worklist_ + W_, scene_ids, n * sizeof(int32_t));
        GD_CUDA(cudaMemcpyAsync(d_worklist_ + W_, h_worklist_ + W_, n * sizeof(int32_t), cudaMemcpyHostToDevice, stream_));
        launchBindReset(list, true, /*scene_ids_uploaded=*/true);
    }
    launchRest(false);
    sync();
}

void Engine::autoResetPool(const int32_t *scene_ids, int n) {
    GD_CUDA(cudaSetDevice(gpu_));
    if (n <= 0) throw std::runtime_error("auto_reset_pool: empty pool");
    for (int i = 0; i < n; i++)
        if (scene_ids[i] < 0 || scene_ids[i] >= (int)scenes_.size() || !scenes_[scene_ids[i]].live)
            throw std::runtime_error("auto_reset_pool: scene id out of range");
    for (int i = 0; i < n; i++) scenes_[scene_ids[i]].pinned = true;
    GD_CUDA(cudaStreamSynchronize(stream_));
    if (d_pool_ids_) GD_CUDA(cudaFree(d_pool_ids_));
    GD_CUDA(cudaMalloc(&d_pool_ids_, sizeof(int32_t) * (size_t)n));
    GD_CUDA(cudaMemcpy(d_pool_ids_, scene_ids, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice));
    if (!d_pool_cursor_) d_pool_cursor_ = devAlloc<int32_t>(owned_, W_);
    std::vector<int32_t> cur(W_);
    for (int w = 0; w < W_; w++) cur[w] = w % n;        // world w starts at pool entry w % n
    GD_CUDA(cudaMemcpy(d_pool_cursor_, cur.data(), sizeof(int32_t) * (size_t)W_, cudaMemcpyHostToDevice));
    n_pool_ = n;
}

void Engine::autoReset() {
    GD_CUDA(cudaSetDevice(gpu_));
    if (!n_pool_) throw std::runtime_error("auto_reset: call auto_reset_pool first");
    syncSceneTable();
    k_autoreset_select<<<(W_ * 32 + 255) / 256, 256, 0, stream_>>>(st_, ex_, d_scenes_, d_pool_ids_, n_pool_,
                                                                    d_pool_cursor_, W_, S_);
    k_bind<<<dim3(W_, BIND_SPLIT), 256, 0, stream_>>>(nullptr, d_scenes_, st_, ex_, S_, fusedPolicy());
    const int blk = ((S_ + 31) / 32) * 32;
    k_reset<<<W_, blk, 0, stream_>>>(nullptr, d_scenes_, st_, ex_, S_, params_.dynamicsModel);
    launches_ += 3;
    launchRest(false, true);      // the Reset graph's rest-of-step, for the flagged worlds only
    GD_CUDA(cudaGetLastError());
    host_binding_stale_ = true;
}

void Engine::refreshHostBinding() {
    if (!host_binding_stale_) return;
    GD_CUDA(cudaMemcpyAsync(world_scene_.data(), st_.world_scene, W_ * sizeof(int32_t), cudaMemcpyDeviceToHost, stream_));
    GD_CUDA(cudaStreamSynchronize(stream_));
    for (int w = 0; w < W_; w++) {
        const std::string &key = scenes_[world_scene_[w]].key;
        scene_paths_[w] = key.substr(0, key.find("#del"));
        if (key.find("#del") == std::string::npos) deleted_[w].clear();
    }
    host_binding_stale_ = false;
}

void Engine::stepHost(const float *actions, int action_cols, float *reward, int32_t *done, int32_t *info) {
    GD_CUDA(cudaSetDevice(gpu_));
    const size_t WS = (size_t)W_ * S_;
    if (action_cols <= 0 || action_cols > kActionSize) throw std::runtime_error("stepHost: action_cols must be in [1, 10]");
    if (actions) {
        // one contiguous H2D into a landing buffer, then a device-side scatter into the [W,S,10] action rows
        const size_t n = WS * (size_t)action_cols;
        GD_CUDA(cudaMemcpyAsync(d_act_stage_, actions, n * sizeof(float), cudaMemcpyHostToDevice, stream_));
        k_scatter_actions<<<(unsigned)((n + 255) / 256), 256, 0, stream_>>>(d_act_stage_, action_cols, ex_.action, n);
        launches_++;
    }
    mark_agents_done_ = true;
    stepAsync();
    mark_agents_done_ = false;
    // k_agents has written reward / done / info; k_obs never touches them: copy them out on a second
    // stream while the observation kernel runs
    GD_CUDA(cudaStreamWaitEvent(copy_stream_, ev_agents_, 0));
    if (reward) GD_CUDA(cudaMemcpyAsync(reward, ex_.reward, WS * sizeof(float), cudaMemcpyDeviceToHost, copy_stream_));
    if (done) GD_CUDA(cudaMemcpyAsync(done, ex_.done, WS * sizeof(int32_t), cudaMemcpyDeviceToHost, copy_stream_));
    if (info) GD_CUDA(cudaMemcpyAsync(info, ex_.info, WS * kInfoSize * sizeof(int32_t), cudaMemcpyDeviceToHost, copy_stream_));
    GD_CUDA(cudaStreamSynchronize(copy_stream_));
    sync();
}

int Engine::policyWidth() const { return kPolicyEgo + (S_ - 1) * kPolicyPartner + kMaxAgentMapObs * kPolicyRoad; }

TensorDesc Engine::policyObs(int normalize) {
    GD_CUDA(cudaSetDevice(gpu_));
    const int width = policyWidth();
    if (!d_policy_obs_) d_policy_obs_ = devAlloc<float>(owned_, (size_t)W_ * S_ * width);
    if (opt_fused_policy_ && normalize == 2) {
        // k_obs has already written it (GD_OPT_FUSED_POLICY_OBS): nothing to launch
    } else {
        if (opt_fused_policy_)
            throw std::runtime_error("policy_obs: with GD_OPT_FUSED_POLICY_OBS the buffer holds the normalised observation in "
                                     "torch-CUDA arithmetic (normalize = 2); switch the option off for the other forms");
        if (normalize == 2) k_policy_obs<true><<<W_ * S_, 256, 0, stream_>>>(ex_, d_policy_obs_, S_, 1);
        else k_policy_obs<false><<<W_ * S_, 256, 0, stream_>>>(ex_, d_policy_obs_, S_, normalize ? 1 : 0);
        launches_++;
        GD_CUDA(cudaGetLastError());
    }
    if (!opt_async_aux_) GD_CUDA(cudaStreamSynchronize(stream_));
    TensorDesc d{};
    d.ptr = d_policy_obs_; d.dtype = 0; d.ndim = 3; d.shape[0] = W_; d.shape[1] = S_; d.shape[2] = width; d.device = gpu_;
    return d;
}

TensorDesc Engine::shapedReward(int mode, float collision_w, float goal_w, float off_road_w, float log_dist_w,
                                const float *agent_weights, const int32_t *world_time_steps) {
    GD_CUDA(cudaSetDevice(gpu_));
    if (mode < RS_Sparse || mode > RS_DistanceToLogs) throw std::runtime_error("unknown reward shaping mode");
    if (mode == RS_Conditioned && !agent_weights) throw std::runtime_error("reward_conditioned needs the [W,S,3] weight tensor");
    if (!d_shaped_reward_) d_shaped_reward_ = devAlloc<float>(owned_, (size_t)W_ * S_);
    RewardShapingArgs a{mode, collision_w, goal_w, off_road_w, log_dist_w, agent_weights, world_time_steps};
    const int n = W_ * S_;
    k_shaped_reward<<<(n + 255) / 256, 256, 0, stream_>>>(ex_, a, d_shaped_reward_, W_, S_);
    launches_++;
    GD_CUDA(cudaGetLastError());
    if (!opt_async_aux_) GD_CUDA(cudaStreamSynchronize(stream_));
    TensorDesc d{};
    d.ptr = d_shaped_reward_; d.dtype = 0; d.ndim = 2; d.shape[0] = W_; d.shape[1] = S_; d.device = gpu_;
    return d;
}

TensorDesc Engine::expertActions() {
    GD_CUDA(cudaSetDevice(gpu_));
    const int cols = params_.dynamicsModel == DM_State ? 10 : 3;
    const size_t n = (size_t)W_ * S_ * kTrajLen;
    if (!d_expert_actions_) d_expert_actions_ = devAlloc<float>(owned_, n * cols);
    k_expert_actions<<<(unsigned)((n + 255) / 256), 256, 0, stream_>>>(ex_, d_expert_actions_, n, params_.dynamicsModel);
    launches_++;
    GD_CUDA(cudaGetLastError());
    if (!opt_async_aux_) GD_CUDA(cudaStreamSynchronize(stream_));
    TensorDesc d{};
    d.ptr = d_expert_actions_; d.dtype = 0; d.ndim = 4; d.shape[0] = W_; d.shape[1] = S_; d.shape[2] = kTrajLen;
    d.shape[3] = cols; d.device = gpu_;
    return d;
}

void Engine::setOption(int option, int value) {
    if (option == 0) opt_knn_idx_ = value ? 1 : 0;
    else if (option == 1) {
        GD_CUDA(cudaSetDevice(gpu_));
        if (value && prof_ev_.empty()) {
            prof_ev_.resize(3 * kProfRing);
            for (auto &e : prof_ev_) GD_CUDA(cudaEventCreate(&e));
        }
        if (value && !opt_profile_) { prof_pending_ = 0; prof_agents_ms_ = prof_obs_ms_ = 0; prof_steps_ = 0; }
        opt_profile_ = value ? 1 : 0;
    } else if (option == 2) opt_graph_ = value ? 1 : 0;
    else if (option == 3) opt_async_aux_ = value ? 1 : 0;
    else if (option == 4) {
        GD_CUDA(cudaSetDevice(gpu_));
        if (value && !opt_fused_policy_) {
            // from now on k_obs (live slots, every step / reset) and k_bind (padding slots) keep the buffer current;
            // fill it once from the exports as they stand
            if (!d_policy_obs_) d_policy_obs_ = devAlloc<float>(owned_, (size_t)W_ * S_ * policyWidth());
            k_policy_obs<true><<<W_ * S_, 256, 0, stream_>>>(ex_, d_policy_obs_, S_, 1);
            launches_++;
            GD_CUDA(cudaGetLastError());
        }
        opt_fused_policy_ = value ? 1 : 0;
    }
    else throw std::runtime_error("unknown option");
}

void Engine::foldProfile() {
    if (!prof_pending_) return;
    GD_CUDA(cudaEventSynchronize(prof_ev_[3 * (prof_pending_ - 1) + 2]));
    for (int i = 0; i < prof_pending_; i++) {
        float a = 0, o = 0;
        GD_CUDA(cudaEventElapsedTime(&a, prof_ev_[3 * i], prof_ev_[3 * i + 1]));
        GD_CUDA(cudaEventElapsedTime(&o, prof_ev_[3 * i + 1], prof_ev_[3 * i + 2]));
        prof_agents_ms_ += a; prof_obs_ms_ += o;
    }
    prof_steps_ += prof_pending_;
    prof_pending_ = 0;
}

void Engine::kernelTimes(double *agents_ms, double *obs_ms, int64_t *steps) {
    GD_CUDA(cudaSetDevice(gpu_));
    foldProfile();
    *agents_ms = prof_agents_ms_; *obs_ms = prof_obs_ms_; *steps = prof_steps_;
}

StepStats Engine::stats() {
    GD_CUDA(cudaSetDevice(gpu_));
    StepStats s;
    memset(&s, 0, sizeof(s));
    unsigned long long dev[ST_NUM];
    GD_CUDA(cudaMemcpyAsync(dev, st_.stats, sizeof(dev), cudaMemcpyDeviceToHost, stream_));
    GD_CUDA(cudaStreamSynchronize(stream_));
    s.steps = steps_;
    s.episodes_reset = episodes_reset_ + dev[ST_AUTO_RESETS];
    s.agent_steps = dev[ST_AGENT_STEPS];
    s.controlled_steps = dev[ST_CONTROLLED_STEPS];
    s.collided_road = dev[ST_COLLIDED_ROAD];
    s.collided_vehicle = dev[ST_COLLIDED_VEHICLE];
    s.collided_nonvehicle = dev[ST_COLLIDED_NONVEHICLE];
    s.goals_reached = dev[ST_GOALS_REACHED];
    s.dones = dev[ST_DONES];
    s.sat_tests = dev[ST_SAT_TESTS];
    return s;
}

TensorDesc Engine::tensor(int id) const {
    TensorDesc d;
    memset(&d, 0, sizeof(d));
    d.device = gpu_;
    auto set = [&](void *p, int dtype, std::initializer_list<int64_t> shp) {
        d.ptr = p; d.dtype = dtype; d.ndim = (int)shp.size();
        int i = 0;
        for (int64_t v : shp) d.shape[i++] = v;
    };
    const int64_t W = W_, S = S_;
    switch (id) {
    case T_ACTION: set(ex_.action, 0, {W, S, kActionSize}); break;
    case T_REWARD: set(ex_.reward, 0, {W, S, 1}); break;
    case T_DONE: set(ex_.done, 1, {W, S, 1}); break;
    case T_SELF_OBS: set(ex_.self_obs, 0, {W, S, kSelfObsSize}); break;
    case T_MAP_OBS: set(ex_.map_obs, 0, {W, kMaxRoadEntityCount, kMapObsSize}); break;
    case T_PARTNER_OBS: set(ex_.partner_obs, 0, {W, S, S - 1, kPartnerObsSize}); break;
    case T_AGENT_ROADMAP: set(ex_.agent_roadmap, 0, {W, S, kMaxAgentMapObs, kMapObsSize}); break;
    case T_LIDAR: set(ex_.lidar, 0, {W, S, 3, kNumLidarSamples, kLidarRowSize}); break;
    case T_STEPS_REMAINING: set(ex_.steps_remaining, 1, {W, S, 1}); break;
    case T_SHAPE: set(ex_.shape, 1, {W, 2}); break;
    case T_CONTROLLED: set(ex_.controlled, 1, {W, S, 1}); break;
    case T_ABS_OBS: set(ex_.abs_obs, 0, {W, S, kAbsObsSize}); break;
    case T_VALID_STATE: set(ex_.valid_state, 1, {W, S, 1}); break;
    case T_INFO: set(ex_.info, 1, {W, S, kInfoSize}); break;
    case T_RESPONSE: set(ex_.response, 1, {W, S, 1}); break;
    case T_TRAJ: set(ex_.traj, 0, {W, S, kTrajSize}); break;
    case T_WORLD_MEANS: set(ex_.world_means, 0, {W, 3}); break;
    case T_METADATA: set(ex_.metadata, 1, {W, S, kMetaSize}); break;
    case T_DELETED: set(ex_.deleted, 1, {W, S}); break;
    case T_MAP_NAME: set(ex_.map_name, 1, {W, 32}); break;
    case T_SCENARIO_ID: set(ex_.scenario_id, 1, {W, 32}); break;
    case T_KNN_IDX: set(ex_.knn_idx, 1, {W, S, kMaxAgentMapObs}); break;
    case T_BEV:
        if (!bev_enabled_)
            throw std::runtime_error("bev_observation_tensor: the bird's-eye-view grid is opt-in (160 KB per agent "
                                     "slot): create the manager with GdConfig.flags |= GD_FLAG_BEV / enable_bev=True");
        set(ex_.bev, 0, {W, S, BEV_RES, BEV_RES, 1});
        break;
    default: throw std::runtime_error("unknown tensor id");
    }
    return d;
}

} // namespace gd

// ------------------------------------------------------------------------------------------------------------
// gd_scatter_actions_noalloc (include/gpudrive_b200.h): the action staging of gd_step_host - contiguous
// [n_rows, cols] actions -> a landing buffer -> k_scatter_actions -> the first `cols` columns of [n_rows, 10] action
// rows - with the landing buffer in this file-scope device array instead of cudaMalloc'ed memory, for callers that
// must not allocate (CUDA graph capture, tight serving loops). No counterpart in the reference.
// Every GPU that loads the library holds the array (16 MiB) until exit.
// In host code `g_noalloc_stage` is only a stand-in for the symbol, not a device address: the address is resolved
// once per device with cudaGetSymbolAddress (below) and that pointer is what the copy and the kernel get.
// ------------------------------------------------------------------------------------------------------------
namespace gd {
void setLastError(const std::string &msg);          // capi.cpp
constexpr size_t kNoAllocMaxFloats = size_t(4) << 20;         // 4 Mi floats = 16 MiB
static __device__ __align__(256) float g_noalloc_stage[kNoAllocMaxFloats];
constexpr int kNoAllocMaxDevices = 64;
static float *g_noalloc_ptr[kNoAllocMaxDevices];              // per device: device address of g_noalloc_stage
}

extern "C" int gd_scatter_actions_noalloc(const float *actions, int64_t n_rows, int32_t action_cols,
                                          float *action_rows_out, uint64_t stream)
{
    using namespace gd;
    auto fail = [](const std::string &m) { setLastError("gd_scatter_actions_noalloc: " + m); return 1; };
    if (!actions || !action_rows_out) return fail("null pointer");
    if (action_cols <= 0 || action_cols > kActionSize) return fail("action_cols must be in [1, 10]");
    if (n_rows < 0) return fail("negative row count");
    const unsigned long long n = (unsigned long long)n_rows * (unsigned long long)action_cols;
    if (n > kNoAllocMaxFloats)
        return fail("n_rows * action_cols = " + std::to_string(n) + " floats exceeds the static staging array's maximum of " +
                    std::to_string(kNoAllocMaxFloats) + " floats (16 MiB); split the call or use gd_step_host");
    if (n == 0) return 0;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= kNoAllocMaxDevices) return fail("device ordinal out of range");
    if (!g_noalloc_ptr[dev]) {       // set-up, once per device, with that device current
        float *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_stage);
        if (e != cudaSuccess || !p) return fail(std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_noalloc_ptr[dev] = p;
    }
    float *stage = g_noalloc_ptr[dev];
    cudaStream_t st = (cudaStream_t)(uintptr_t)stream;
    // (nothing in the array has to start at zero: the kernel reads exactly the n floats this call wrote)
    e = cudaMemcpyAsync(stage, actions, n * sizeof(float), cudaMemcpyDefault, st);
    if (e != cudaSuccess) return fail(std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e));
    k_scatter_actions<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(stage, action_cols, action_rows_out, (size_t)n);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    return 0;
}

#ifdef GD_COUNTERS
extern "C" int gd_debug_counters(unsigned long long *out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, gd::g_obs_cnt, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(gd::g_obs_cnt, z, sizeof(z)); }
    return 0;
}
#endif
