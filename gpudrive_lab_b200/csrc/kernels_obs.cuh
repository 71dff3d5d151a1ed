// k_obs: observation fan-out of the step graph, one warp per agent.
//   collectSelfObsSystem / collectAbsoluteObservationsSystem   src/sim.cpp:168-186,769-783
//   collectPartnerObsSystem                                     src/sim.cpp:188-240
//   collectMapObservationsSystem + selectKNearestRoadEntities   src/sim.cpp:242-280, src/knn.hpp:103-158
//   lidarSystem + BVH::traceRay + traceRayIntoConvexPolyhedron  src/sim.cpp:394-460,
//                                   external/madrona/src/physics/broadphase.cpp:656-888
//
// Road entities arrive sorted by (z-class, Morton code) in groups of 32 with two bounds per group (SceneDev::grp):
// an oriented rectangle along the principal axis of the member boxes, used for every distance decision, and the
// circle around the same centre, used for the angular range. A one-level stand-in for the reference's BVH that
// fits a warp: one lane tests one group, and a group that survives is processed by all 32 lanes at once.
//   k-nearest: groups whose rectangle reaches the observation radius -> exact squared-distance keys in the
//     reference's operation order -> in-radius candidates -> histogram selection of the K smallest (key, road
//     index); dense maps refine over the near groups first, an exact radix select is the last resort. The selected
//     SET is bit-exact. The linear algorithm uses the road index as the key and emits rows in index order.
//   lidar: ray directions with glibc's own sincosf (gdtrig::sincosf_glibc). Groups are visited in bands of
//     increasing distance. Before each band the rays that are still "open" (no hit closer than the band) are
//     gathered into one bit mask per plane; a group, and then a box, is skipped with one AND if the rays its
//     angular range covers are all closed. The warp expands the surviving (box, ray range) pairs of a pass into
//     (box, ray) pairs, 32 per round, one per lane (prefix sum + a per-warp owner table: no queue), so 32 lanes
//     clip rays in lock step. The clip (ray_box) is a box-frame slab test on intervals: every slab parameter
//     carries its rounding bound, a hit is "certain" or "marginal". The closest hit per ray is kept with a native
//     32-bit shared atomicMin on the bits of t, the box id (+ the marginal flag) written by the lane that holds
//     the minimum. At write-out the deciding face of a certain winner is evaluated with the reference's exact
//     object-space arithmetic, so exported depths carry its rounding; a ray whose winner is marginal is handed
//     to k_lidar_fix (kernels_lidar_fix.cuh), which re-traces it exactly.
//   rows leave through a per-warp staging tile as 16-byte vector stores (flush_aligned); k_obs<true> also turns
//     each chunk of rows into the policy observation rows (kernels_policy.cuh) on the way out.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace gd {

#ifdef GD_COUNTERS     // kernel experiments only: work counters read back through gd_debug_counters
__device__ unsigned long long g_obs_cnt[16];
#define GD_CNT(i, v) do { const unsigned long long cv_ = (unsigned long long)(v); if (lane == 0) atomicAdd(&g_obs_cnt[i], cv_); } while (0)
#else
#define GD_CNT(i, v) do { } while (0)
#endif

#ifndef GD_OBS_WARPS
#define GD_OBS_WARPS 8
#endif
constexpr int OBS_WARPS = GD_OBS_WARPS;   // agents per CTA
constexpr int OBS_BLOCK = OBS_WARPS * 32;
#ifndef GD_OBS_STAGE_W
#define GD_OBS_STAGE_W 13                 // floats per staged row: the widest row is the fused policy road row
#endif
constexpr int OBS_CAND = 512;             // in-radius candidates kept per agent before the slow path
constexpr int OBS_GCAP = 128;             // group descriptors per pass (4 per lane; larger maps take several passes)
constexpr int NRAY = kNumLidarSamples;    // 50
#ifndef GD_OBS_BANDS
#define GD_OBS_BANDS 7
#endif
constexpr int NBAND = GD_OBS_BANDS;       // distance bands: 7: [0,4) [4,8) [8,16) [16,32) [32,64) [64,128) [128,inf)
                                          //                 4: [0,4) [4,16) [16,64) [64,inf)
// band of a group whose nearest point is dmin away (e = floor(log2(dmin))), and the lower edge of a band
__device__ __forceinline__ int obs_band_of(float dmin, int e) {
    if (NBAND == 7) return dmin < 4.f ? 0 : min(NBAND - 1, e - 1);
    return dmin < 4.f ? 0 : min(NBAND - 1, e >> 1);
}
__device__ __forceinline__ float obs_band_lo(int band) {
    if (NBAND == 7) return band == 0 ? 0.f : (float)(2 << band);
    return band == 0 ? 0.f : (float)(1 << (2 * band));
}
constexpr unsigned long long RAY_ALL = (1ull << NRAY) - 1ull;

struct ObsWarpSmem {
    uint32_t ckey[OBS_CAND];                  // candidate key: squared-distance bits (k-nearest) or road index (linear)
    uint16_t cpos[OBS_CAND];                  // candidate slot in the sorted road arrays
    union {
        struct {
            uint32_t bt[3 * NRAY];            // lidar: bits of the closest hit distance per plane and ray (none: ~0)
            uint16_t bid[3 * NRAY];           // and its box: sorted road slot, or 0x4000 | agent slot; 0x8000: marginal hit
        } l;
        float stage[32 * GD_OBS_STAGE_W + 4];             // row transpose buffer for coalesced stores (+ up to 3 floats of misalignment);
                                              // 13 floats per row: the widest row is the fused policy road row
    } u;
    union {                                   // the histogram is only needed after the lidar is written out
        struct {
            uint32_t glist[OBS_GCAP];         // surviving groups: id | band << 12 | lo << 15 | hi << 21 | near << 27 | planes << 28
            float2 raydir[NRAY];              // normalised world-space direction
            float2 rayxy[NRAY];               // (cos theta, sin theta) in the agent frame
        } g;
        uint32_t hist[256];                   // radix-select histogram
    } v;
    uint8_t own[32];                          // clip stage: the r-th lane that brought a non-empty ray range
};

struct ObsSmem {
    float4 a0[128];     // per agent slot: x, y, qw, qz
    float4 a1[128];     // z, speed, type, id
    ObsWarpSmem wp[OBS_WARPS];
};

// ------------------------------------------------------------------------------------------
// lidar helpers
// ------------------------------------------------------------------------------------------

struct LidarCtx {
    float ox, oy;        // ray origin (agent position)
    float fc, fs;        // cos / sin of the fan-frame x axis (middle of the fan) in world
};

__device__ __forceinline__ float best_t(const uint32_t *bt, int plane, int ray) {
    return __uint_as_float(bt[plane * NRAY + ray]);      // "none" (~0) reads as NaN
}

__device__ __forceinline__ unsigned long long ray_bits(int lo, int hi) {   // bits lo..hi, 0 <= lo <= hi < 64
    return ((2ull << hi) - 1ull) & ~((1ull << lo) - 1ull);
}

__device__ __forceinline__ unsigned long long open_for(const unsigned long long (&open)[3], int pm) {
    return ((pm & 1) ? open[0] : 0ull) | ((pm & 2) ? open[1] : 0ull) | ((pm & 4) ? open[2] : 0ull);
}

// Approximate reciprocal (MUFU.RCP). Only on culling / candidate paths: everything exported is recomputed
// with IEEE operations in the reference's order.
__device__ __forceinline__ float rcp_fast(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Clip one ray against one box in the box frame (pure rotation, unscaled): equivalent to
// traceRayIntoConvexPolyhedron over the four vertical faces (broadphase.cpp:754-855) with
// tfirst = 0, tlast = lidarDistance, "entering t >= tfirst" and "only back faces -> miss" - up to rounding.
// The reference decides hit / miss in its own float arithmetic (object space, IEEE divisions); this fast form
// (approximate reciprocals, fused multiply-adds) can disagree with it when the ray grazes a corner, starts on a
// face or ends at the range limit. So each slab's two parameters carry an error bound - a direction component has
// ~3 eps of absolute error, i.e. 3 eps / |component| relative, which multiplies the parameter (d + |p|) / |component|;
// the box-frame origin adds ~2 eps of its length - and the test is done on intervals: a pair that is a possible hit
// but not a certain one is flagged `marginal`; if it wins its ray, k_lidar_fix re-traces the ray with the
// reference's exact clip. The bounds are per slab: a ray nearly parallel to one axis has huge, imprecise parameters
// for the other slab, but those do not bind and must not blur the decision (a lumped bound re-traced 1 ray in 370).
// Nothing exported ever rests on this function's rounding.
// (px, py): ray origin in the box frame. Explicit fmaf: this translation unit is compiled with --fmad=false.
__device__ __forceinline__ float ray_box(float px, float py, float cphi, float sphi, float d0, float d1,
                                         float dx, float dy, bool &marginal)
{
    // Branch-free slabs. A direction component of exactly 0 gives an infinite reciprocal (and an infinite bound:
    // the pair is settled exactly); fminf / fmaxf drop the NaN of 0 * inf (origin exactly on a plane).
    const float lx = fmaf(cphi, dx, sphi * dy), ly = fmaf(cphi, dy, -(sphi * dx));   // direction in box frame
    const float ivx = rcp_fast(lx), ivy = rcp_fast(ly);
    const float x1 = (-d0 - px) * ivx, x2 = (d0 - px) * ivx;
    const float y1 = (-d1 - py) * ivy, y2 = (d1 - py) * ivy;
    const float xmn = fminf(x1, x2), xmx = fmaxf(x1, x2), ymn = fminf(y1, y2), ymx = fmaxf(y1, y2);
    const float tnear = fmaxf(xmn, ymn);
    const float apx = fabsf(px), apy = fabsf(py), aivx = fabsf(ivx), aivy = fabsf(ivy), rr = apx + apy;
    const float ex = 6e-7f * aivx * fmaf(d0 + apx, aivx, rr), ey = 6e-7f * aivy * fmaf(d1 + apy, aivy, rr);
    const float tn_lo = fmaxf(xmn - ex, ymn - ey), tn_hi = fmaxf(xmn + ex, ymn + ey);
    const float tf_lo = fminf(kLidarDistance, fminf(xmx - ex, ymx - ey));
    const float tf_hi = fminf(kLidarDistance, fminf(xmx + ex, ymx + ey));
    const bool possible = tn_hi >= 0.f && tn_lo <= tf_hi;
    marginal = !(tn_lo >= 0.f && tn_hi <= tf_lo);
    return possible ? fmaxf(tnear, 0.f) : -1.f;
}

// The same clip with the reference's own arithmetic: ray moved into object space (rotate, divide by the
// scale), planes +x, -x, +y, -y with d = 1, IEEE divisions (broadphase.cpp:754-855, 817-845). Run once
// per ray on the winning box, so the exported depth carries the reference's rounding.
__device__ __forceinline__ float ray_box_exact(float ox, float oy, float bx, float by, float qw, float qz,
                                               float d0, float d1, float dirx, float diry)
{
    float lox, loy, ldx, ldy;
    inv_rotate_xy(qw, qz, __fsub_rn(ox, bx), __fsub_rn(oy, by), lox, loy);
    lox = __fdiv_rn(lox, d0); loy = __fdiv_rn(loy, d1);
    inv_rotate_xy(qw, qz, dirx, diry, ldx, ldy);
    ldx = __fdiv_rn(ldx, d0); ldy = __fdiv_rn(ldy, d1);
    // The four vertical faces in the reference's order (+x, -x, +y, -y), without its early exits: a miss is
    // a sticky flag, so the planes after it cannot change the outcome.
    float tfirst = 0.f, tlast = kLidarDistance;
    bool entered = false, miss = false;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float sgn = (k & 1) ? -1.f : 1.f;
        const float denom = sgn * ((k < 2) ? ldx : ldy);
        const float neg_dist = __fsub_rn(1.f, sgn * ((k < 2) ? lox : loy));
        const bool par = denom == 0.0f;
        const float t = __fdiv_rn(neg_dist, par ? 1.f : denom);      // (a zero operand takes the division's slow path)
        miss = miss || (par && neg_dist < 0.0f);
        const bool enter = !par && denom < 0.0f && t >= tfirst;
        const bool leave = !par && denom > 0.0f && t <= tlast;
        tfirst = enter ? t : tfirst;
        tlast = leave ? t : tlast;
        entered = entered || enter;
        miss = miss || (!par && tfirst > tlast);
    }
    return (entered && !miss) ? tfirst : -1.f;
}

// The exported depth of a ray whose winning box is already known: the reference's arithmetic for the one face
// that decides it. A hit's depth is the largest entry parameter over the entering faces (broadphase.cpp:817-845
// keeps the later, larger t); the slab with the larger entry is chosen with cheap arithmetic, then that face's
// t is computed exactly as the reference does (rotate, divide origin and direction by the scale, IEEE division),
// so the result has the reference's rounding. If the two entries tie within rounding, either face's t is the
// depth to ~1e-7. Three IEEE divisions instead of the eight of the full restatement above.
__device__ __forceinline__ float ray_box_exact_depth(float ox, float oy, float bx, float by, float qw, float qz,
                                                     float d0, float d1, float dirx, float diry)
{
    float lox, loy, ldx, ldy;
    inv_rotate_xy(qw, qz, __fsub_rn(ox, bx), __fsub_rn(oy, by), lox, loy);
    inv_rotate_xy(qw, qz, dirx, diry, ldx, ldy);
    const float ex = (-copysignf(d0, ldx) - lox) * rcp_fast(ldx);      // entry parameters, unscaled box frame
    const float ey = (-copysignf(d1, ldy) - loy) * rcp_fast(ldy);
    const bool use_x = ex > ey;                                        // (ties and NaNs go to y, like the reference's order)
    const float sc = use_x ? d0 : d1;
    const float slo = __fdiv_rn(use_x ? lox : loy, sc), sld = __fdiv_rn(use_x ? ldx : ldy, sc);
    const float sgn = sld < 0.f ? 1.f : -1.f;                          // the entering face of the axis: denom < 0
    const float denom = sgn * sld, neg_dist = __fsub_rn(1.f, sgn * slo);
    return denom == 0.f ? -1.f : __fdiv_rn(neg_dist, denom);
}

// atan2 with |error| < 2e-4 rad (odd minimax polynomial on [0, 1] + octant folding)
__device__ __forceinline__ float fast_atan2(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = mn * rcp_fast(fmaxf(mx, 1e-30f));
    const float s = a * a;
    float r = fmaf(fmaf(fmaf(-0.0464964749f, s, 0.15931422f), s, -0.327622764f) * s, a, a);
    if (ay > ax) r = 1.57079637f - r;
    if (x < 0.f) r = 3.14159274f - r;
    return y < 0.f ? -r : r;
}

// Ray index range [lo, hi] an angular interval (centre a, half width w, fan frame) can touch; false if none.
__device__ __forceinline__ bool ray_range(float a, float w, int &lo, int &hi) {
    const float inv_delta = (float)NRAY / (2.f * kLidarAngle);
    const float half = 0.5f * (float)(NRAY - 1);
    const float flo = fmaf(a - w, inv_delta, half), fhi = fmaf(a + w, inv_delta, half);
    lo = max(0, __float2int_ru(flo));
    hi = min(NRAY - 1, __float2int_rd(fhi));
    return !(fhi < 0.f) && !(flo > (float)(NRAY - 1)) && lo <= hi;
}

// Which rays can touch this box? Conservative: never drops a ray the exact clip would accept.
// Returns false if none, else the ray index range [lo, hi]. (relx, rely): box centre relative to the ray
// origin; (cphi, sphi): box axis. Written without early exits: in a warp some lane needs every step anyway,
// and straight-line code with selects is cheaper than the divergence bookkeeping of the branches.
struct BoxCull {      // first half of lidar_consider: the cheap tests, and what the second half reuses
    float fx, fy, maxcross, mindot;
    bool close_box, alive;
};

__device__ __forceinline__ BoxCull lidar_cull(const LidarCtx &c, float relx, float rely, float cphi, float sphi,
                                              float d0, float d1)
{
    BoxCull r;
    const float rbound = d0 + d1;                         // >= half diagonal
    const float D2 = fmaf(relx, relx, rely * rely);
    const float range = kLidarDistance + rbound;
    const bool in_range = D2 <= range * range;
    r.fx = fmaf(c.fc, relx, c.fs * rely); r.fy = fmaf(c.fc, rely, -(c.fs * relx));   // centre in the fan frame
    // box half axes e0 = d0 (cphi, sphi), e1 = d1 (-sphi, cphi)
    const float along = fmaf(relx, cphi, rely * sphi), across = fmaf(rely, cphi, -(relx * sphi));
    r.maxcross = fmaf(d0, fabsf(across), d1 * fabsf(along));
    r.mindot = D2 - fmaf(d0, fabsf(along), d1 * fabsf(across));
    r.close_box = !(r.mindot > 1e-6f * D2) || D2 <= (rbound + 0.01f) * (rbound + 0.01f);
    const float cH = 0.5146014f, sH = 0.8574295f;     // cos / sin of 24.5 * delta + 4e-3 (fan half angle)
    const bool outside = r.fx < -rbound                                  // behind the fan
                         || fmaf(cH, r.fy, -(sH * r.fx)) > rbound         // beyond the last ray
                         || fmaf(cH, r.fy, sH * r.fx) < -rbound;          // before the first ray
    r.alive = in_range && (r.close_box || !outside);
    return r;
}

// second half: the ray index range of a box that passed lidar_cull
__device__ __forceinline__ bool lidar_range(const BoxCull &b, int &lo, int &hi)
{
    const float w = b.maxcross * rcp_fast(b.mindot);      // tan of the half width; atan(w) <= w
    int lo2, hi2;
    const bool some = ray_range(fast_atan2(b.fy, b.fx), w * 1.0001f + 2e-3f, lo2, hi2);
    const bool narrow = !b.close_box && w < 1.0f;
    lo = narrow ? lo2 : 0;
    hi = narrow ? hi2 : NRAY - 1;
    return b.alive && (!narrow || some);
}

__device__ __forceinline__ bool lidar_consider(const LidarCtx &c, float relx, float rely, float cphi, float sphi,
                                               float d0, float d1, int &lo, int &hi)
{
    const BoxCull b = lidar_cull(c, relx, rely, cphi, sphi, d0, d1);
    return lidar_range(b, lo, hi);
}

// z-slab membership of the three lidar planes for a hull with object-space z in [zmin_obj, zmax_obj]
// (cube: [-1, 1]; agent hull: [0, 2]); horizontal rays => denom == 0 => "neg_dist < 0 -> miss".
__device__ __forceinline__ int plane_mask(float pz, float bz, float d2, float zmin_obj, float zmax_obj)
{
    int m = 0;
    const float inv = rcp_fast(d2);
    // (not unrolled, like the lane-stride loops of this kernel: k_obs is sensitive to its instruction footprint -
    // unrolling these few loops cost 3.7 % of the launch in instruction-fetch stalls)
#pragma unroll 1
    for (int p = 0; p < 3; p++) {
        const float oz = __fadd_rn(pz, p == 0 ? kLidarCarOffset : (p == 1 ? kLidarRoadEdgeOffset : kLidarRoadLineOffset));
        // The reference tests lz = (oz - bz) / d2 against the slab [zmin, zmax] (touching counts as inside). Float
        // subtraction never flips a sign, so that is zmin <= lz <= zmax, and an approximate quotient decides it
        // unless lz is within 1e-5 of a face: only then is the IEEE division evaluated (and never with a zero
        // operand, which sends it down its slow path - a plane at exactly the box's height is the common case).
        const float num = __fsub_rn(oz, bz);
        const float q = num * inv;
        bool in = q >= zmin_obj && q <= zmax_obj;
        const float tol = 1e-5f * fmaxf(1.f, fabsf(q));
        // (|q| >= 1e30: an agent parked at the padding position, z = FLT_MAX - far outside, and the IEEE division would
        // overflow, which sends every warp that sees one such partner down the division's slow path)
        if (fabsf(q) < 1e30f && (fabsf(q - zmin_obj) <= tol || fabsf(q - zmax_obj) <= tol)) {
            const float lz = num == 0.f ? 0.f : __fdiv_rn(num, d2);
            in = !(__fsub_rn(zmax_obj, lz) < 0.f) && !(__fsub_rn(lz, zmin_obj) < 0.f);
        }
        m |= in ? (1 << p) : 0;
    }
    return m;
}

// ------------------------------------------------------------------------------------------
// lidar clip stage: every lane brings one box and a ray range [lo, lo + n); the warp expands the 32
// ranges into (box, ray) pairs, 32 pairs per round, one pair per lane. A lane finds its pair's owner in a
// small per-warp table (see lidar_expand) and fetches the box from that lane with shuffles, so
// no queue and no reload of the box are needed. The closest hit per (plane, ray) is kept as a native
// 32-bit shared atomicMin on the bits of t; the box id is written by whichever lane holds the minimum
// after the round's atomics (warp-synchronous, so the last writer is always the current minimum).
// ------------------------------------------------------------------------------------------
// One round of hits: `hit` lanes hold a (ray, t bits, box id, plane mask). (Measured and dropped: a warp-uniform
// single-plane form with one atomic per lane, no gain; a direct path where lanes with at most two rays clip
// their own rays without the expansion, 7 % slower.)
__device__ __forceinline__ void lidar_commit(ObsWarpSmem &wp, bool hit, int ray, unsigned tb, unsigned id, int spm)
{
    if (hit) {
        if (spm & 1) atomicMin(&wp.u.l.bt[ray], tb);
        if (spm & 2) atomicMin(&wp.u.l.bt[NRAY + ray], tb);
        if (spm & 4) atomicMin(&wp.u.l.bt[2 * NRAY + ray], tb);
    }
    __syncwarp();
    if (hit) {
        if ((spm & 1) && wp.u.l.bt[ray] == tb) wp.u.l.bid[ray] = (uint16_t)id;
        if ((spm & 2) && wp.u.l.bt[NRAY + ray] == tb) wp.u.l.bid[NRAY + ray] = (uint16_t)id;
        if ((spm & 4) && wp.u.l.bt[2 * NRAY + ray] == tb) wp.u.l.bid[2 * NRAY + ray] = (uint16_t)id;
    }
}

// The owner of pair j - the lane whose ray range contains it - without a search: the lanes that brought a range are
// listed once (wp.own); per round one warp OR collects the positions at which a new owner's range starts, and a
// lane's owner is the last one that started at or before its position: a popc and a byte load, where a binary
// search over the prefix sums took five dependent shuffles.
__device__ __forceinline__ void lidar_expand(ObsWarpSmem &wp, int lane, int lo, int n, int pm, unsigned id,
                                             float relx, float rely, float cphi, float sphi, float d0, float d1)
{
    const unsigned FULL = 0xffffffffu;
    const unsigned live = __ballot_sync(FULL, n > 0);
    if (!live) return;
    // ray origin in the box frame (the box centre is (relx, rely) away from the origin)
    const float opx = -fmaf(cphi, relx, sphi * rely), opy = -fmaf(cphi, rely, -(sphi * relx));
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
    }
    const int total = __shfl_sync(FULL, incl, 31);
    GD_CNT(5, total); GD_CNT(6, 1);
    const int excl = incl - n;
    if (n > 0) wp.own[__popc(live & ((1u << lane) - 1u))] = (uint8_t)lane;
    __syncwarp();
    const unsigned meta = (unsigned)(lo - excl + 2048) | ((unsigned)pm << 12) | (id << 15);   // lo - excl in [-1600, 49]
    int r0 = 0;                                       // owners whose range starts before this round
#pragma unroll 1
    for (int base = 0; base < total; base += 32) {
        const int j = base + lane;
        const unsigned rel = (unsigned)(excl - base);
        const unsigned starts = __reduce_or_sync(FULL, (n > 0 && rel < 32u) ? (1u << rel) : 0u);
        const int s = wp.own[r0 + __popc(starts & (FULL >> (31 - lane))) - 1];      // (round 0 starts at position 0)
        r0 += __popc(starts);
        const float spx = __shfl_sync(FULL, opx, s), spy = __shfl_sync(FULL, opy, s);
        const float sc_ = __shfl_sync(FULL, cphi, s), ss_ = __shfl_sync(FULL, sphi, s);
        const float sd0 = __shfl_sync(FULL, d0, s), sd1 = __shfl_sync(FULL, d1, s);
        const unsigned smeta = __shfl_sync(FULL, meta, s);
        const int ray = min(max((int)(smeta & 4095u) - 2048 + j, 0), NRAY - 1);   // (clamped for the lanes past `total`)
        const float2 d = wp.v.g.raydir[ray];
        bool marginal;
        const float t = ray_box(spx, spy, sc_, ss_, sd0, sd1, d.x, d.y, marginal);
        GD_CNT(9, 1);
        lidar_commit(wp, j < total && t >= 0.f, ray, __float_as_uint(t), ((smeta >> 15) & 0x7fffu) | (marginal ? 0x8000u : 0u),
                     (int)((smeta >> 12) & 7u));
    }
    __syncwarp();
}

// Rays of each plane with no hit closer than `dist` yet (bit i = ray i).
__device__ __forceinline__ void open_rays(const uint32_t *best, float dist, int lane,
                                          unsigned long long (&open)[3])
{
    const unsigned FULL = 0xffffffffu;
#pragma unroll
    for (int p = 0; p < 3; p++) {
        const bool a = !(best_t(best, p, lane) <= dist);                     // NaN (no hit) -> open
        const bool b = lane + 32 < NRAY && !(best_t(best, p, min(lane + 32, NRAY - 1)) <= dist);
        open[p] = (unsigned long long)__ballot_sync(FULL, a) | ((unsigned long long)__ballot_sync(FULL, b) << 32);
    }
}


// Emit `m` rows held one per lane (lanes with has_row) to consecutive global rows starting at dst, via the
// per-warp transpose buffer. rank = exclusive prefix of has_row. dst is only 4-byte aligned in general (a row is
// 36 bytes): the rows are staged `mis` = (float index of dst) & 3 floats into the buffer, so that shared and global
// addresses are congruent modulo 16 bytes and everything but the <= 3 floats in front of the first and behind the
// last 16-byte boundary moves as LDS.128 / STG.128 (72 vector stores per 32 rows instead of 288 scalar ones).
__device__ __forceinline__ void flush_aligned(const float *stage, float *dst, int mis, int n, int lane)
{
    const int head = min((4 - mis) & 3, n);
    if (lane < head) dst[lane] = stage[mis + lane];
    const int nq = (n - head) >> 2;
    const float4 *s4 = reinterpret_cast<const float4 *>(stage + mis + head);
    float4 *d4 = reinterpret_cast<float4 *>(dst + head);
#pragma unroll 1
    for (int i = lane; i < nq; i += 32) d4[i] = s4[i];
    const int t0 = head + 4 * nq;
    if (lane < n - t0) dst[t0 + lane] = stage[mis + t0 + lane];
}

template <int NF>
__device__ __forceinline__ void flush_rows_n(float *stage, float *dst, const float (&row)[NF], bool has_row,
                                             int rank, int m, int lane)
{
    const int mis = (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u);
    if (has_row) {
#pragma unroll
        for (int k = 0; k < NF; k++) stage[mis + rank * NF + k] = row[k];
    }
    __syncwarp();
    flush_aligned(stage, dst, mis, m * NF, lane);
    __syncwarp();
}

__device__ __forceinline__ void flush_rows(float *stage, float *dst, const float (&row)[9], bool has_row,
                                           int rank, int m, int lane)
{
    flush_rows_n<9>(stage, dst, row, has_row, rank, m, lane);
}

} // namespace gd
