// k_lidar_fix: exact re-trace of the few lidar rays k_obs does not settle.
// k_obs finds candidate hits with fast arithmetic widened by a rounding bound; a hit inside that bound ("marginal":
// grazed corner, origin on a face, range limit) that wins its ray - about one ray in 200 000 - is listed in
// StateDev::retrace instead of being exported. Here one CTA takes one listed ray and clips it, with the reference's
// exact arithmetic (broadphase.cpp:754-855), against every hull visible in its plane (the other agents, then the
// road entities of every group whose bounding circle the ray can reach), one box per thread, closest hit kept; equal depths go to the lower id (the reference:
// to the leaf its BVH visits last, broadphase.cpp:700-712 - the depth is the same either way). A separate launch,
// because inside k_obs the exact clip and the re-trace cost every agent up to 25 % more time.
#pragma once
#include "kernels_obs.cuh"

namespace gd {

// latency-bound (a chain of dependent global loads per ray, little arithmetic): many small CTAs, one ray each
constexpr int FIX_BLOCK = 128, FIX_GRID = 148 * 12;

__global__ void __launch_bounds__(FIX_BLOCK)
k_lidar_fix(const SceneDev *__restrict__ scenes, StateDev st, Exports ex, int S)
{
    __shared__ unsigned long long s_best;      // (depth bits << 32) | box id: one atomicMin keeps depth, then id, minimal
    __shared__ int s_zmask;                    // 3 plane bits per z-class
    __shared__ int s_ngroups;
    __shared__ uint16_t s_groups[kMaxRoadEntityCount / 32 + 24];   // road groups the ray can reach
    const int tid = threadIdx.x;
    const unsigned count = min(st.retrace[0], kRetraceCap);
    for (unsigned it = blockIdx.x; it < count; it += FIX_GRID) {
        const unsigned code = st.retrace[1 + it];
        const size_t wa = code / (3 * NRAY);
        const int ray = (int)(code % (3 * NRAY)), plane = ray / NRAY, ri = ray % NRAY;
        const int w = (int)(wa / S), a = (int)(wa % S);
        const SceneDev &sc = scenes[st.world_scene[w]];
        const int A = sc.numAgents;
        const float4 p = st.pos[wa], rv = st.rv[wa];
        const float px = p.x, py = p.y, pz = p.z, qw = rv.x, qz = rv.y;
        __syncthreads();                       // the previous ray's result has been read
        if (tid == 0) { s_best = ~0ull; s_zmask = 0; s_ngroups = 0; }
        __syncthreads();
        if (tid < sc.numZClasses) atomicOr(&s_zmask, plane_mask(pz, sc.zc_z[tid], sc.zc_d2[tid], -1.f, 1.f) << (3 * tid));
        // the ray, exactly as k_obs builds it (sim.cpp:409-412)
        const float head = ex.controlled[wa] ? ex.action[wa * kActionSize + 2] : 0.f;
        float rgx, rgy, fwx, fwy;
        rotate_xy(qw, qz, 1.f, 0.f, rgx, rgy);
        rotate_xy(qw, qz, 0.f, 1.f, fwx, fwy);
        const float frac = ri == 0 ? 0.f : __fdiv_rn(2.f * (float)ri, (float)NRAY);
        const float theta = __fadd_rn(__fmul_rn(kLidarAngle, __fsub_rn(frac, 1.f)), head);
        const float x = cos_hp(theta), y = sin_hp(theta);
        float dx = __fadd_rn(__fmul_rn(x, rgx), __fmul_rn(y, fwx));
        float dy = __fadd_rn(__fmul_rn(x, rgy), __fmul_rn(y, fwy));
        const float inv = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(len2_xy(dx, dy), 0.f)));
        dx = __fmul_rn(dx, inv); dy = __fmul_rn(dy, inv);
        __syncthreads();
        const int zmask = s_zmask;

        // Exact clip (eight IEEE divisions) only for boxes the fast clip, widened by its rounding bound, cannot rule out.
        auto clip = [&](float bx, float by, float bqw, float bqz, float d0, float d1) -> float {
            const float relx = bx - px, rely = by - py;
            const float cphi = 1.f - 2.f * bqz * bqz, sphi = 2.f * bqw * bqz;
            bool marginal;
            const float tf = ray_box(-fmaf(cphi, relx, sphi * rely), -fmaf(cphi, rely, -(sphi * relx)), cphi, sphi, d0, d1,
                                     dx, dy, marginal);
            return tf < 0.f ? -1.f : ray_box_exact(px, py, bx, by, bqw, bqz, d0, d1, dx, dy);
        };
        unsigned long long best = ~0ull;
        for (int j = tid; j < A; j += FIX_BLOCK) {
            if (j == a) continue;              // own hull: every face is a back face
            const size_t wj = (size_t)w * S + j;
            const float4 op = st.pos[wj], orv = st.rv[wj]; const float4 gb = sc.ag_b[j];
            if (!((plane_mask(pz, op.z, gb.z, 0.f, 2.f) >> plane) & 1)) continue;
            const float t = clip(op.x, op.y, orv.x, orv.y, gb.x, gb.y);
            if (t >= 0.f) best = min(best, ((unsigned long long)__float_as_uint(t) << 32) | (0x4000u | (unsigned)j));
        }
        // road groups whose bounding circle the ray can reach (distance of the centre from the ray's line and
        // position along it, with a margin far above the rounding of these few operations), then their boxes
        const int G = sc.numGroups;
        for (int g = tid; g < G; g += FIX_BLOCK) {
            const float4 gc = sc.grp[2 * g];
            const unsigned cm = __float_as_uint(gc.w) & 0xffu;
            const int zc = cm ? __ffs(cm) - 1 : 0;
            const float relx = gc.x - px, rely = gc.y - py, r = gc.z * 1.001f + 0.05f;
            const float along = fmaf(relx, dx, rely * dy), across = fabsf(fmaf(rely, dx, -(relx * dy)));
            if (cm && ((zmask >> (3 * zc + plane)) & 1) && across <= r && along >= -r && along <= kLidarDistance + r)
                s_groups[atomicAdd(&s_ngroups, 1)] = (uint16_t)g;
        }
        __syncthreads();
        const int ng = s_ngroups;
        for (int i = tid; i < ng * 32; i += FIX_BLOCK) {
            const int e = (int)s_groups[i >> 5] * 32 + (i & 31);
            const unsigned zc = sc.s_c[e] & 0xffu;
            if (zc >= (unsigned)sc.numZClasses) continue;                     // dummy slot
            const float4 ra = sc.s_a[e]; const float2 rb = sc.s_b[e];
            const float t = clip(ra.x, ra.y, ra.z, ra.w, rb.x, rb.y);
            if (t >= 0.f) best = min(best, ((unsigned long long)__float_as_uint(t) << 32) | (unsigned)e);
        }
        if (best != ~0ull) atomicMin(&s_best, best);
        __syncthreads();
        if (tid == 0) {
            const unsigned long long wb = s_best;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (wb != ~0ull) {
                const unsigned wid = (unsigned)wb;
                const int type = (wid & 0x4000u) ? (int)sc.ag_b[wid & 0x3fffu].w : (int)((sc.s_c[wid] >> 8) & 0xffu);
                const float t = __uint_as_float((unsigned)(wb >> 32));
                v = make_float4(t, (float)type, __fmul_rn(t, x), __fmul_rn(t, y));
            }
            reinterpret_cast<float4 *>(ex.lidar + wa * 3 * NRAY * kLidarRowSize)[ray] = v;
        }
    }
}

} // namespace gd
