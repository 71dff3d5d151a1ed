// k_lidar_fix: exact re-trace of the few lidar rays k_obs does not settle.
// k_obs finds candidate hits with fast arithmetic widened by a rounding bound; a hit inside that bound ("marginal":
// grazed corner, origin on a face, range limit) that wins its ray - about one ray in 28 000 - is listed in
// StateDev::retrace instead of being exported. Here one warp takes one listed ray and clips it, with the reference's
// exact arithmetic (broadphase.cpp:754-855), against every hull visible in its plane (the other agents, then the
// road entities of every group whose bounding circle the ray can reach), one box per lane - first with the fast clip
// widened by its rounding bound, then the survivors exactly - closest hit kept; equal depths go to the lower id (the reference:
// to the leaf its BVH visits last, broadphase.cpp:700-712 - the depth is the same either way). A separate launch,
// because inside k_obs the exact clip and the re-trace cost every agent up to 25 % more time.
#pragma once
#include "kernels_obs.cuh"

namespace gd {

// latency-bound (a chain of dependent global loads per ray, little arithmetic): one warp per ray, many warps in flight
constexpr int FIX_BLOCK = 128, FIX_WARPS = FIX_BLOCK / 32, FIX_GRID = 148 * 16, FIX_LIST = 96;

__global__ void __launch_bounds__(FIX_BLOCK)
k_lidar_fix(const SceneDev *__restrict__ scenes, StateDev st, Exports ex, int S)
{
    __shared__ uint16_t s_list[FIX_WARPS][FIX_LIST];     // boxes the fast clip cannot rule out (0x4000 | agent slot, or road slot)
    __shared__ uint16_t s_glist[FIX_WARPS][kMaxRoadEntityCount / 32 + 24];   // road groups the ray can reach
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int warp = blockIdx.x * FIX_WARPS + wib, nwarps = FIX_GRID * FIX_WARPS;
    uint16_t *list = s_list[wib], *glist = s_glist[wib];
    const unsigned count = min(st.retrace[0], kRetraceCap);
    for (unsigned it = warp; it < count; it += nwarps) {
        const unsigned code = st.retrace[1 + it];
        const size_t wa = code / (3 * NRAY);
        const int ray = (int)(code % (3 * NRAY)), plane = ray / NRAY, ri = ray % NRAY;
        const int w = (int)(wa / S), a = (int)(wa % S);
        const SceneDev &sc = scenes[st.world_scene[w]];
        const int A = sc.numAgents;
        const float4 p = st.pos[wa], rv = st.rv[wa];
        const float px = p.x, py = p.y, pz = p.z, qw = rv.x, qz = rv.y;
        int zmask = 0;                           // 3 plane bits per z-class
        if (lane < sc.numZClasses) zmask = plane_mask(pz, sc.zc_z[lane], sc.zc_d2[lane], -1.f, 1.f) << (3 * lane);
        zmask = (int)__reduce_or_sync(FULL, (unsigned)zmask);
        // the ray, exactly as k_obs builds it (sim.cpp:409-412)
        const float head = ex.controlled[wa] ? ex.action[wa * kActionSize + 2] : 0.f;
        float rgx, rgy, fwx, fwy;
        rotate_xy(qw, qz, 1.f, 0.f, rgx, rgy);
        rotate_xy(qw, qz, 0.f, 1.f, fwx, fwy);
        const float frac = ri == 0 ? 0.f : __fdiv_rn(2.f * (float)ri, (float)NRAY);
        const float theta = __fadd_rn(__fmul_rn(kLidarAngle, __fsub_rn(frac, 1.f)), head);
        float x, y;
        sincos_hp(theta, y, x);
        float dx = __fadd_rn(__fmul_rn(x, rgx), __fmul_rn(y, fwx));
        float dy = __fadd_rn(__fmul_rn(x, rgy), __fmul_rn(y, fwy));
        const float inv = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(len2_xy(dx, dy), 0.f)));
        dx = __fmul_rn(dx, inv); dy = __fmul_rn(dy, inv);

        // pass 1: the fast clip, widened by its rounding bound, over every box the ray can reach; survivors are listed
        int nl = 0;
        unsigned long long best = ~0ull;         // (depth bits << 32) | box id: depth first, then the lower id
        auto exact = [&](unsigned id) {
            float bx, by, bqw, bqz, d0, d1;
            if (id & 0x4000u) {
                const size_t wj = (size_t)w * S + (id & 0x3fffu);
                const float4 op = st.pos[wj], orv = st.rv[wj]; const float4 gb = sc.ag_b[id & 0x3fffu];
                bx = op.x; by = op.y; bqw = orv.x; bqz = orv.y; d0 = gb.x; d1 = gb.y;
            } else {
                const float4 ra = sc.s_a[id]; const float2 rb = sc.s_b[id];
                bx = ra.x; by = ra.y; bqw = ra.z; bqz = ra.w; d0 = rb.x; d1 = rb.y;
            }
            const float t = ray_box_exact(px, py, bx, by, bqw, bqz, d0, d1, dx, dy);
            if (t >= 0.f) best = min(best, ((unsigned long long)__float_as_uint(t) << 32) | id);
        };
        auto drain = [&]() {                     // the exact clip (eight IEEE divisions), one listed box per lane
            __syncwarp();
            for (int i = lane; i < nl; i += 32) exact(list[i]);
            nl = 0;
            __syncwarp();
        };
        auto consider = [&](bool valid, unsigned id, float bx, float by, float bqw, float bqz, float d0, float d1) {
            bool pass = false;
            if (valid) {
                const float relx = bx - px, rely = by - py;
                const float cphi = 1.f - 2.f * bqz * bqz, sphi = 2.f * bqw * bqz;
                bool marginal;
                pass = ray_box(-fmaf(cphi, relx, sphi * rely), -fmaf(cphi, rely, -(sphi * relx)), cphi, sphi, d0, d1,
                               dx, dy, marginal) >= 0.f;
            }
            const unsigned pm = __ballot_sync(FULL, pass);
            if (nl + __popc(pm) > FIX_LIST) drain();
            if (pass) list[nl + __popc(pm & lt_mask)] = (uint16_t)id;
            nl += __popc(pm);
            GD_CNT(12, __popc(pm));
        };
        for (int jb = 0; jb < A; jb += 32) {
            const int j = jb + lane;
            bool valid = j < A && j != a;        // own hull: every face is a back face
            float4 op = make_float4(0.f, 0.f, 0.f, 0.f), orv = op, gb = op;
            if (valid) {
                const size_t wj = (size_t)w * S + j;
                op = st.pos[wj]; orv = st.rv[wj]; gb = sc.ag_b[j];
                valid = ((plane_mask(pz, op.z, gb.z, 0.f, 2.f) >> plane) & 1) != 0;
            }
            consider(valid, 0x4000u | (unsigned)j, op.x, op.y, orv.x, orv.y, gb.x, gb.y);
        }
        // road groups whose bounding circle the ray can reach (distance of the centre from the ray's line and
        // position along it, with a margin far above the rounding of these few operations), then their boxes
        const int G = sc.numGroups;
        int ng = 0;
        for (int gb0 = 0; gb0 < G; gb0 += 32) {
            const int g = gb0 + lane;
            bool reach = false;
            if (g < G) {
                const float4 gc = sc.grp[2 * g];
                const unsigned cm = __float_as_uint(gc.w) & 0xffu;
                const int zc = cm ? __ffs(cm) - 1 : 0;
                const float relx = gc.x - px, rely = gc.y - py, r = gc.z * 1.001f + 0.05f;
                const float along = fmaf(relx, dx, rely * dy), across = fabsf(fmaf(rely, dx, -(relx * dy)));
                reach = cm && ((zmask >> (3 * zc + plane)) & 1) && across <= r && along >= -r && along <= kLidarDistance + r;
            }
            const unsigned gm = __ballot_sync(FULL, reach);
            if (reach) glist[ng + __popc(gm & lt_mask)] = (uint16_t)g;
            ng += __popc(gm);
        }
        __syncwarp();
        GD_CNT(14, 1); GD_CNT(11, ng);
        // four groups per iteration: their twelve loads are in flight together (the kernel is a chain of dependent
        // loads; with one group at a time a single ray took 80 us whatever the ray count)
        for (int gi = 0; gi < ng; gi += 4) {
            float4 ra[4]; float2 rb[4]; bool valid[4]; int e[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const bool have = gi + q < ng;
                e[q] = (have ? (int)glist[gi + q] : (int)glist[gi]) * 32 + lane;
                valid[q] = have && (sc.s_c[e[q]] & 0xffu) < (unsigned)sc.numZClasses;       // not a dummy slot
                ra[q] = sc.s_a[e[q]]; rb[q] = sc.s_b[e[q]];
            }
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (gi + q < ng) consider(valid[q], (unsigned)e[q], ra[q].x, ra[q].y, ra[q].z, ra[q].w, rb[q].x, rb[q].y);
        }
        drain();
        // pass 2 result: closest exact hit over the warp
        unsigned hi = (unsigned)(best >> 32), lo = (unsigned)best;
        const unsigned whi = __reduce_min_sync(FULL, hi);
        const unsigned wid = __reduce_min_sync(FULL, hi == whi ? lo : 0xffffffffu);
        if (lane == 0) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (whi != 0xffffffffu) {
                const int type = (wid & 0x4000u) ? (int)sc.ag_b[wid & 0x3fffu].w : (int)((sc.s_c[wid] >> 8) & 0xffu);
                const float t = __uint_as_float(whi);
                v = make_float4(t, (float)type, __fmul_rn(t, x), __fmul_rn(t, y));
            }
            reinterpret_cast<float4 *>(ex.lidar + wa * 3 * NRAY * kLidarRowSize)[ray] = v;
        }
        __syncwarp();
    }
}

} // namespace gd
