// k_bev: collectBevObservationsSystem (src/sim.cpp:462-555) + rasterizer::rasterizeRotatedRectangle
// (src/rasterizer.hpp:12-78). Opt-in (GdConfig.flags & GD_FLAG_BEV): the export is 160 KB per agent slot.
// One CTA per live agent. The 200 x 200 grid lives in shared memory as (order << 4 | type) words; every
// rectangle is stamped over its bounding square with atomicMax, where `order` is the rectangle's position
// in the reference's sequential loop (first <= 200 in-radius road entities in road-index order, then the
// other agents in OtherAgents order), so the surviving value of a cell is the one the reference's
// last-writer-wins loop leaves. All arithmetic follows the reference's operation order (IEEE division and
// square root, glibc-exact cosf / sinf, double atan2), because a cell on a rectangle's edge flips on 1 ulp.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace gd {

constexpr int BEV_RES = 200;                 // consts::bev_rasterization_resolution, src/consts.hpp:49
constexpr int BEV_BLOCK = 512;
constexpr int BEV_CHUNK = 512;               // rectangles staged per round

struct BevRect {
    float cx, cy, cosy, siny, half_l, half_w;
    int32_t gxy;        // gx | gy << 8 | box_radius << 16
    uint32_t value;     // (order + 1) << 4 | type
};

struct BevSmem {
    uint32_t cell[BEV_RES * BEV_RES];
    BevRect rect[BEV_CHUNK];
    int32_t warp_cnt[BEV_BLOCK / 32];
    int32_t order;
};

// toGridCoords + the set-up half of rasterizeRotatedRectangle
__device__ __forceinline__ BevRect bev_make_rect(float cx, float cy, float yaw, float length, float width, int type,
                                                 int order, float radius)
{
    BevRect q;
    const float res = (float)BEV_RES;
    q.half_w = __fdiv_rn(width, 2.0f);
    q.half_l = __fdiv_rn(length, 2.0f);
    const float scale_px = __fdiv_rn(2.f * radius, res);
    const float scale_m = __fdiv_rn(res, 2.f * radius);
    int gx = __float2int_rz(__fmul_rn(__fadd_rn(cx, radius), scale_m));
    int gy = __float2int_rz(__fmul_rn(__fadd_rn(cy, radius), scale_m));
    gx = min(max(0, gx), BEV_RES - 1); gy = min(max(0, gy), BEV_RES - 1);
    const float max_side = fmaxf(q.half_w, q.half_l);
    int br = __float2int_rz(ceilf(__fdiv_rn(__fsqrt_rn(2.f * __fmul_rn(max_side, max_side)), scale_px)));
    br = min(br, 255);                       // beyond the grid anyway (the loop clips to [0, 200))
    q.cx = cx; q.cy = cy;
    q.cosy = cos_hp(-yaw); q.siny = sin_hp(-yaw);
    q.gxy = gx | (gy << 8) | (br << 16);
    q.value = ((uint32_t)(order + 1) << 4) | (uint32_t)type;
    return q;
}

// one warp stamps one rectangle
__device__ __forceinline__ void bev_stamp(uint32_t *cell, const BevRect &q, float radius, int lane)
{
    const int gx = q.gxy & 255, gy = (q.gxy >> 8) & 255, br = q.gxy >> 16;
    const float scale_px = __fdiv_rn(2.f * radius, (float)BEV_RES);
    const int x0 = max(gx - br, 0), x1 = min(gx + br, BEV_RES - 1);
    const int y0 = max(gy - br, 0), y1 = min(gy + br, BEV_RES - 1);
    const int wdt = x1 - x0 + 1, n = wdt * (y1 - y0 + 1);
    const float hl = __fadd_rn(q.half_l, 1e-3f), hw = __fadd_rn(q.half_w, 1e-3f);
    for (int i = lane; i < n; i += 32) {
        const int y = y0 + i / wdt, x = x0 + i % wdt;
        const float px = __fsub_rn(__fmul_rn((float)x, scale_px), radius);
        const float py = __fsub_rn(__fmul_rn((float)y, scale_px), radius);
        const float ldx = __fsub_rn(px, q.cx), ldy = __fsub_rn(py, q.cy);
        const float lx = __fsub_rn(__fmul_rn(ldx, q.cosy), __fmul_rn(ldy, q.siny));
        const float ly = __fadd_rn(__fmul_rn(ldx, q.siny), __fmul_rn(ldy, q.cosy));
        if (fabsf(lx) <= hl && fabsf(ly) <= hw) atomicMax(&cell[y * BEV_RES + x], q.value);
    }
}

__global__ void __launch_bounds__(BEV_BLOCK, 1)
k_bev(const SceneDev *__restrict__ scenes, StateDev st, Exports ex, Params prm, int S, float key_thr)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BevSmem &sm = *reinterpret_cast<BevSmem *>(smem_raw);
    const int w = blockIdx.x / S, a = blockIdx.x % S;
    const SceneDev &sc = scenes[st.world_scene[w]];
    const int A = sc.numAgents, R = sc.numRoads;
    if (a >= A || prm.disableClassicalObs) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = BEV_BLOCK / 32;
    const size_t wa = (size_t)w * S + a;
    const float4 p = st.pos[wa], rv = st.rv[wa];
    const float px = p.x, py = p.y, qw = rv.x, qz = rv.y;
    const float radius = prm.observationRadius;

    {
        uint4 *c4 = reinterpret_cast<uint4 *>(sm.cell);
        for (int i = tid; i < BEV_RES * BEV_RES / 4; i += BEV_BLOCK) c4[i] = make_uint4(0, 0, 0, 0);
        if (tid == 0) sm.order = 0;
    }
    __syncthreads();

    // ---- roads: the first kMaxAgentMapObs in-radius entities in road-index order (sim.cpp:484-525)
    for (int base = 0; base < R; base += BEV_BLOCK) {
        const int order0 = sm.order;
        if (order0 >= kMaxAgentMapObs) break;
        const int r = base + tid;
        bool in = false; float ox = 0.f, oy = 0.f; float4 ra = make_float4(0, 0, 1, 0);
        if (r < R) {
            ra = sc.road_a[r];
            inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
            in = len2_xy(ox, oy) <= key_thr;                  // == !(length() > observationRadius)
        }
        const unsigned bm = __ballot_sync(0xffffffffu, in);
        if (lane == 0) sm.warp_cnt[warp] = __popc(bm);
        __syncthreads();
        int before = 0, total = 0;
        for (int k = 0; k < NW; k++) { const int c = sm.warp_cnt[k]; if (k < warp) before += c; total += c; }
        const int slot = before + __popc(bm & ((1u << lane) - 1u));
        const int take = min(total, kMaxAgentMapObs - order0);
        if (in && slot < take) {
            const float2 rb = sc.road_b[r];
            float rw, rz;
            inv_mul_quat(qw, qz, ra.z, ra.w, rw, rz);
            const float d1 = fmaxf(rb.y, __fdiv_rn(2.f * radius, (float)BEV_RES));     // minimum segment width
            sm.rect[slot] = bev_make_rect(ox, oy, quat_to_yaw_hp(rw, rz), rb.x, d1, (int)(sc.road_c[r] >> 8) & 0xff,
                                          order0 + slot, radius);
        }
        __syncthreads();
        for (int i = warp; i < take; i += NW) bev_stamp(sm.cell, sm.rect[i], radius, lane);
        if (tid == 0) sm.order = order0 + take;
        __syncthreads();
    }

    // ---- the other agents, OtherAgents order (sim.cpp:527-554); always after every road rectangle
    for (int base = 0; base < A; base += BEV_BLOCK) {
        const int j = base + tid;
        bool in = false; float ox = 0.f, oy = 0.f; float4 orv = make_float4(1, 0, 0, 0);
        if (j < A && j != a) {
            const float4 op = st.pos[(size_t)w * S + j];
            orv = st.rv[(size_t)w * S + j];
            inv_rotate_xy(qw, qz, __fsub_rn(op.x, px), __fsub_rn(op.y, py), ox, oy);
            in = !(__fsqrt_rn(len2_xy(ox, oy)) > radius);
        }
        if (j < A) {
            BevRect q; q.gxy = -1;
            if (in) {
                const float4 ga = sc.ag_a[j];
                float rw, rz;
                inv_mul_quat(qw, qz, orv.x, orv.y, rw, rz);
                q = bev_make_rect(ox, oy, quat_to_yaw_hp(rw, rz), ga.x, ga.y, (int)sc.ag_b[j].w,
                                  kMaxAgentMapObs + j, radius);
            }
            sm.rect[tid] = q;
        }
        __syncthreads();
        const int n = min(BEV_BLOCK, A - base);
        for (int i = warp; i < n; i += NW)
            if (sm.rect[i].gxy >= 0) bev_stamp(sm.cell, sm.rect[i], radius, lane);
        __syncthreads();
    }

    // ---- write out: type as float, [200][200] row-major (grid.obs[y][x])
    float4 *out = reinterpret_cast<float4 *>(ex.bev + wa * (size_t)(BEV_RES * BEV_RES));
    const uint4 *c4 = reinterpret_cast<const uint4 *>(sm.cell);
    for (int i = tid; i < BEV_RES * BEV_RES / 4; i += BEV_BLOCK) {
        const uint4 c = c4[i];
        out[i] = make_float4((float)(c.x & 15u), (float)(c.y & 15u), (float)(c.z & 15u), (float)(c.w & 15u));
    }
}

} // namespace gd
