// k_obs: observation fan-out of the step graph for one group of 8 agent slots of one world
// (one warp per agent; the CTA shares shared-memory tiles of the world's road table).
//   collectSelfObsSystem / collectAbsoluteObservationsSystem   src/sim.cpp:168-186,769-783
//   collectPartnerObsSystem                                     src/sim.cpp:188-240
//   collectMapObservationsSystem + selectKNearestRoadEntities   src/sim.cpp:242-280, src/knn.hpp:103-158
//   lidarSystem + BVH::traceRay + traceRayIntoConvexPolyhedron  src/sim.cpp:394-460,
//                                   external/madrona/src/physics/broadphase.cpp:656-888
// The heap of knn.hpp is replaced by: in-radius compaction (index order) -> exact radix select of
// the K smallest (key, index) -> rows emitted in road-index order. The BVH of the lidar is replaced
// by a conservative angular cull per box followed by the reference's exact slab clipping.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace gd {

constexpr int OBS_WARPS = 8;              // agents per CTA
constexpr int OBS_BLOCK = OBS_WARPS * 32;
constexpr int OBS_TILE = 1024;            // road entities per shared-memory tile
constexpr int OBS_CAND = 1024;            // in-radius candidates kept per agent before fallback
constexpr int NRAY = kNumLidarSamples;    // 50
constexpr unsigned long long LIDAR_NONE = 0xFFFFFFFFFFFFFFFFull;

struct ObsWarpSmem {
    uint2 cand[OBS_CAND];                     // (key bits, road index), index order
    union {
        unsigned long long best[3 * NRAY];    // lidar: (t bits << 32) | type, per plane and ray
        float stage[32 * 9];                  // row transpose buffer for coalesced stores
        uint32_t hist[256];                   // radix-select histogram
    } u;
    float2 raydir[NRAY];                      // normalised world-space direction
    float2 rayxy[NRAY];                       // (cos theta, sin theta) in the agent frame
};

struct ObsSmem {
    float4 a0[128];     // per agent slot: x, y, qw, qz
    float4 a1[128];     // z, speed, type, id
    float4 a2[128];     // d0, d1, d2, -
    float4 a3[128];     // length, width, height, -
    float4 ta[OBS_TILE];
    float2 tb[OBS_TILE];
    uint32_t tc[OBS_TILE];
    ObsWarpSmem wp[OBS_WARPS];
};

// One lidar ray against one box, exactly as traceRayIntoLeaf + traceRayIntoConvexPolyhedron clip the
// x/y slabs of a (+-1 scaled) hull in object space (broadphase.cpp:754-855). (lox, loy): ray origin
// in object space (already divided by the scale); returns hit distance or -1.
__device__ __forceinline__ float ray_box_xy(float lox, float loy, float qw, float qz, float d0, float d1,
                                            float dirx, float diry)
{
    float ldx, ldy;
    inv_rotate_xy(qw, qz, dirx, diry, ldx, ldy);
    ldx = __fdiv_rn(ldx, d0);
    ldy = __fdiv_rn(ldy, d1);
    float tfirst = 0.f, tlast = kLidarDistance;
    bool entered = false;
    // planes +x, -x, +y, -y with normals pointing outwards and d = 1
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float sgn = (k & 1) ? -1.f : 1.f;
        const float dcomp = (k < 2) ? ldx : ldy;
        const float ocomp = (k < 2) ? lox : loy;
        const float denom = sgn * dcomp;
        const float neg_dist = __fsub_rn(1.f, sgn * ocomp);
        if (denom == 0.0f) {
            if (neg_dist < 0.0f) return -1.f;
        } else {
            const float t = __fdiv_rn(neg_dist, denom);
            if (denom < 0.0f) {
                if (t >= tfirst) { tfirst = t; entered = true; }
            } else {
                if (t <= tlast) tlast = t;
            }
            if (tfirst > tlast) return -1.f;
        }
    }
    return entered ? tfirst : -1.f;
}

// Conservative range of ray indices whose direction can touch a box seen from the ray origin.
// (fx, fy): box centre in the fan frame (x axis = middle of the fan); ex/ey: half-extent vectors of
// the box in the same frame. Returns false if no ray can hit.
__device__ __forceinline__ bool fan_range(float fx, float fy, float e0x, float e0y, float e1x, float e1y,
                                          float rbound, int &ilo, int &ihi)
{
    const float delta = (2.f * kLidarAngle) / (float)NRAY;
    const float half = 0.5f * (float)(NRAY - 1);
    const float D2 = fx * fx + fy * fy;
    const float maxcross = fabsf(fx * e0y - fy * e0x) + fabsf(fx * e1y - fy * e1x);
    const float mindot = D2 - fabsf(fx * e0x + fy * e0y) - fabsf(fx * e1x + fy * e1y);
    ilo = 0; ihi = NRAY - 1;
    if (!(mindot > 1e-6f * D2) || D2 <= (rbound + 0.01f) * (rbound + 0.01f)) return true;   // too close: test all
    const float wdt = atanf(maxcross / mindot) + 2e-3f;
    if (wdt > 1.0f) return true;
    const float a = atan2f(fy, fx);
    const float lo = (a - wdt) / delta + half, hi = (a + wdt) / delta + half;
    if (hi < 0.f || lo > (float)(NRAY - 1)) return false;
    ilo = max(0, (int)ceilf(lo));
    ihi = min(NRAY - 1, (int)floorf(hi));
    return ilo <= ihi;
}

__device__ __forceinline__ void lidar_commit(unsigned long long *best, int planeMask, int ray, float t, int type)
{
    const unsigned long long v = ((unsigned long long)__float_as_uint(t) << 32) | (unsigned)type;
#pragma unroll
    for (int p = 0; p < 3; p++)
        if (planeMask & (1 << p)) atomicMin(&best[p * NRAY + ray], v);
}

// Process one candidate box for the lidar of one agent (lane-private loop over its ray range).
__device__ __forceinline__ void lidar_box(const ObsWarpSmem &wp, unsigned long long *best, int planeMask,
                                          float ox, float oy, float fc, float fs,
                                          float bx, float by, float qw, float qz, float d0, float d1, int type)
{
    const float relx = bx - ox, rely = by - oy;
    const float rbound = sqrtf(d0 * d0 + d1 * d1);
    const float range = kLidarDistance + rbound;
    if (relx * relx + rely * rely > range * range) return;
    // box frame vectors in world, then everything rotated into the fan frame
    const float cphi = 1.f - 2.f * qz * qz, sphi = 2.f * qw * qz;
    const float fx = fc * relx + fs * rely, fy = -fs * relx + fc * rely;
    const float e0wx = d0 * cphi, e0wy = d0 * sphi, e1wx = -d1 * sphi, e1wy = d1 * cphi;
    int ilo, ihi;
    if (!fan_range(fx, fy, fc * e0wx + fs * e0wy, -fs * e0wx + fc * e0wy,
                   fc * e1wx + fs * e1wy, -fs * e1wx + fc * e1wy, rbound, ilo, ihi)) return;
    // ray origin in object space: rot^-1 (o - p) / scale   (broadphase.cpp:827-833)
    float lox, loy;
    inv_rotate_xy(qw, qz, __fsub_rn(ox, bx), __fsub_rn(oy, by), lox, loy);
    lox = __fdiv_rn(lox, d0);
    loy = __fdiv_rn(loy, d1);
    for (int i = ilo; i <= ihi; i++) {
        const float2 d = wp.raydir[i];
        const float t = ray_box_xy(lox, loy, qw, qz, d0, d1, d.x, d.y);
        if (t >= 0.f) lidar_commit(best, planeMask, i, t, type);
    }
}

// z-slab membership of the three lidar planes for a hull with object-space z in [zmin_obj, zmax_obj]
// (cube: [-1, 1]; agent hull: [0, 2]); horizontal rays => denom == 0 => "neg_dist < 0 -> miss".
__device__ __forceinline__ int plane_mask(float pz, float bz, float d2, float zmin_obj, float zmax_obj)
{
    int m = 0;
    const float offs[3] = {kLidarCarOffset, kLidarRoadEdgeOffset, kLidarRoadLineOffset};
#pragma unroll
    for (int p = 0; p < 3; p++) {
        const float oz = __fadd_rn(pz, offs[p]);
        const float lz = __fdiv_rn(__fsub_rn(oz, bz), d2);
        const bool in = !(__fsub_rn(zmax_obj, lz) < 0.f) && !(__fsub_rn(lz, zmin_obj) < 0.f);
        m |= in ? (1 << p) : 0;
    }
    return m;
}

// Emit `m` rows held one per lane (lanes with has_row) to consecutive global rows starting at dst,
// via the per-warp transpose buffer. rank = exclusive prefix of has_row.
__device__ __forceinline__ void flush_rows(float *stage, float *dst, const float (&row)[9], bool has_row,
                                           int rank, int m, int lane)
{
    __syncwarp();
    if (has_row) {
#pragma unroll
        for (int k = 0; k < 9; k++) stage[rank * 9 + k] = row[k];
    }
    __syncwarp();
    for (int i = lane; i < m * 9; i += 32) dst[i] = stage[i];
    __syncwarp();
}

__global__ void __launch_bounds__(OBS_BLOCK)
k_obs(const SceneDev *__restrict__ scenes, StateDev st, Exports ex, Params prm, int S, int groups,
      int write_knn_idx)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ObsSmem &sm = *reinterpret_cast<ObsSmem *>(smem_raw);

    const int w = blockIdx.x / groups;
    const int g = blockIdx.x % groups;
    const SceneDev &sc = scenes[st.world_scene[w]];
    const int A = sc.numAgents, R = sc.numRoads;
    if (g * OBS_WARPS >= A) return;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned FULL = 0xffffffffu;
    const unsigned lt_mask = (1u << lane) - 1u;

    // ---- stage the world's agents
    for (int j = tid; j < A; j += OBS_BLOCK) {
        const size_t wj = (size_t)w * S + j;
        const float4 p = st.pos[wj], rv = st.rv[wj];
        const float4 ga = sc.ag_a[j], gb = sc.ag_b[j];
        sm.a0[j] = make_float4(p.x, p.y, rv.x, rv.y);
        sm.a1[j] = make_float4(p.z, len3(rv.z, rv.w, p.w), gb.w, ga.w);
        sm.a2[j] = make_float4(gb.x, gb.y, gb.z, 0.f);
        sm.a3[j] = make_float4(ga.x, ga.y, ga.z, 0.f);
    }
    __syncthreads();

    const int a = g * OBS_WARPS + warp;
    const bool active = a < A;            // inactive warps still take part in the tile barriers
    ObsWarpSmem &wp = sm.wp[warp];
    const size_t wa = (size_t)w * S + (active ? a : 0);

    float px = 0, py = 0, pz = 0, qw = 1, qz = 0;
    if (active) {
        const float4 s0 = sm.a0[a];
        px = s0.x; py = s0.y; qw = s0.z; qz = s0.w; pz = sm.a1[a].x;
    }
    const float radius = prm.observationRadius;
    const bool classical = !prm.disableClassicalObs;
    const bool lidar_on = prm.enableLidar != 0;

    // ------------------------------------------------------------ self + absolute observations
    if (active && lane == 0) {
        const float4 s1 = sm.a1[a], s3 = sm.a3[a];
        const float4 gc = sc.ag_c[a];
        float gx, gy;
        inv_rotate_xy(qw, qz, __fsub_rn(gc.x, px), __fsub_rn(gc.y, py), gx, gy);
        float4 *so = reinterpret_cast<float4 *>(ex.self_obs + wa * kSelfObsSize);
        so[0] = make_float4(s1.y, s3.x, s3.y, s3.z);
        so[1] = make_float4(gx, gy, st.collided[wa] ? 1.f : 0.f, s1.w);
        float2 *ab = reinterpret_cast<float2 *>(ex.abs_obs + wa * kAbsObsSize);
        ab[0] = make_float2(px, py);
        ab[1] = make_float2(pz, qw);
        ab[2] = make_float2(0.f * qz, 0.f * qz);            // Quat x, y = normal.{x,y} * sinhalf
        ab[3] = make_float2(qz, quat_to_yaw_hp(qw, qz));
        ab[4] = make_float2(gc.x, gc.y);
        ab[5] = make_float2(s3.x, s3.y);
        ab[6] = make_float2(s3.z, s1.w);
    }

    // ------------------------------------------------------------ partner observations
    if (active && classical) {
        const int P = S - 1;
        float *dstbase = ex.partner_obs + wa * P * kPartnerObsSize;
        for (int base = 0; base < P; base += 32) {
            const int jp = base + lane;
            const bool has = jp < P;
            float row[9];
            if (has && jp < A - 1) {
                const int j = jp < a ? jp : jp + 1;       // OtherAgents order, level_gen.cpp:450-464
                const float4 o0 = sm.a0[j], o1 = sm.a1[j], o3 = sm.a3[j];
                float rx, ry;
                inv_rotate_xy(qw, qz, __fsub_rn(o0.x, px), __fsub_rn(o0.y, py), rx, ry);
                if (__fsqrt_rn(len2_xy(rx, ry)) > radius) {
                    row[0] = row[1] = row[2] = row[3] = row[4] = row[5] = row[6] = row[7] = 0.f;
                    row[8] = -1.f;                          // PartnerObservation::zero()
                } else {
                    float rw, rz;
                    inv_mul_quat(qw, qz, o0.z, o0.w, rw, rz);
                    row[0] = o1.y; row[1] = rx; row[2] = ry; row[3] = quat_to_yaw(rw, rz);
                    row[4] = o3.x; row[5] = o3.y; row[6] = o3.z; row[7] = o1.z; row[8] = o1.w;
                }
            } else {
                row[0] = row[1] = row[2] = row[3] = row[4] = row[5] = row[6] = row[7] = 0.f;
                row[8] = -2.f;                              // zero_nonexist()
            }
            const int m = min(32, P - base);
            flush_rows(wp.u.stage, dstbase + (size_t)base * 9, row, has, lane, m, lane);
        }
    }

    // ------------------------------------------------------------ lidar set-up
    float fc = 1.f, fs = 0.f;      // cos / sin of the fan-frame x axis (world)
    int zmask_classes = 0;         // 3 bits per z-class
    if (active && lidar_on) {
        __syncwarp();
        for (int i = lane; i < 3 * NRAY; i += 32) wp.u.best[i] = LIDAR_NONE;
        const int controlled = ex.controlled[wa];
        const float head = controlled ? ex.action[wa * kActionSize + 2] : 0.f;   // sim.cpp:409
        float rgx, rgy, fwx, fwy;
        rotate_xy(qw, qz, 1.f, 0.f, rgx, rgy);     // rot.rotateVec(math::right)
        rotate_xy(qw, qz, 0.f, 1.f, fwx, fwy);     // rot.rotateVec(math::fwd)
        for (int i = lane; i < NRAY; i += 32) {
            const float theta = __fadd_rn(__fmul_rn(kLidarAngle,
                __fsub_rn(__fdiv_rn(2.f * (float)i, (float)NRAY), 1.f)), head);
            const float x = cos_hp(theta), y = sin_hp(theta);      // sim.cpp:411-412
            const float dx = __fadd_rn(__fmul_rn(x, rgx), __fmul_rn(y, fwx));
            const float dy = __fadd_rn(__fmul_rn(x, rgy), __fmul_rn(y, fwy));
            const float inv = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(len2_xy(dx, dy), 0.f)));
            wp.raydir[i] = make_float2(__fmul_rn(dx, inv), __fmul_rn(dy, inv));
            wp.rayxy[i] = make_float2(x, y);
        }
        const float yaw = quat_to_yaw(qw, qz);
        const float delta = (2.f * kLidarAngle) / (float)NRAY;
        const float mid = yaw + head - kLidarAngle + 0.5f * (float)(NRAY - 1) * delta;
        sincosf(mid, &fs, &fc);
        for (int c = 0; c < sc.numZClasses; c++)
            zmask_classes |= plane_mask(pz, sc.zc_z[c], sc.zc_d2[c], -1.f, 1.f) << (3 * c);
        __syncwarp();
    }

    // ------------------------------------------------------------ road tiles: kNN candidates + lidar
    int cnt = 0;                   // in-radius entities seen so far (warp-uniform)
    const bool want_cands = active && classical;
    const bool first_k_only = prm.roadObservationAlgorithm == RO_AllEntities;
    for (int t0 = 0; t0 < R; t0 += OBS_TILE) {
        const int tl = min(OBS_TILE, R - t0);
        __syncthreads();
        for (int i = tid; i < tl; i += OBS_BLOCK) {
            sm.ta[i] = sc.road_a[t0 + i];
            sm.tb[i] = sc.road_b[t0 + i];
            sm.tc[i] = sc.road_c[t0 + i];
        }
        __syncthreads();
        if (!active) continue;
        for (int base = 0; base < tl; base += 32) {
            const int r = base + lane;
            const bool valid = r < tl;
            float4 ra = make_float4(0, 0, 1, 0);
            if (valid) ra = sm.ta[r];
            if (want_cands) {
                float ox, oy;
                inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
                const float key = len2_xy(ox, oy);
                const bool in = valid && !(__fsqrt_rn(key) > radius);     // knn.hpp:88 / sim.cpp:268-271
                const unsigned m = __ballot_sync(FULL, in);
                if (in) {
                    const int pos = cnt + __popc(m & lt_mask);
                    if (pos < OBS_CAND) wp.cand[pos] = make_uint2(__float_as_uint(key), (unsigned)(t0 + r));
                }
                cnt += __popc(m);
            }
            if (lidar_on && valid) {
                const uint32_t c = sm.tc[r];
                const int pm = (zmask_classes >> (3 * (c & 0xff))) & 7;
                if (pm) {
                    const float2 rb = sm.tb[r];
                    lidar_box(wp, wp.u.best, pm, px, py, fc, fs, ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, (int)(c >> 8));
                }
            }
        }
    }
    if (!active) return;

    // ------------------------------------------------------------ lidar vs agents, write-out
    if (lidar_on) {
        for (int j = lane; j < A; j += 32) {
            const float4 o0 = sm.a0[j], o1 = sm.a1[j], o2 = sm.a2[j];
            const int pm = plane_mask(pz, o1.x, o2.z, 0.f, 2.f);
            if (pm) lidar_box(wp, wp.u.best, pm, px, py, fc, fs, o0.x, o0.y, o0.z, o0.w, o2.x, o2.y, (int)o1.z);
        }
        __syncwarp();
        float4 *out = reinterpret_cast<float4 *>(ex.lidar + wa * 3 * NRAY * kLidarRowSize);
        for (int i = lane; i < 3 * NRAY; i += 32) {
            const unsigned long long b = wp.u.best[i];
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (b != LIDAR_NONE) {
                const float t = __uint_as_float((unsigned)(b >> 32));
                const float2 xy = wp.rayxy[i % NRAY];
                v = make_float4(t, (float)(unsigned)(b & 0xffffffffu), __fmul_rn(t, xy.x), __fmul_rn(t, xy.y));
            }
            out[i] = v;
        }
        __syncwarp();
    }

    // ------------------------------------------------------------ road-map rows
    if (!classical) return;
    const int K = kMaxAgentMapObs;
    float *rows = ex.agent_roadmap + wa * K * kMapObsSize;
    int32_t *kidx = ex.knn_idx + wa * K;
    int emitted = 0;

    // selection threshold: accept key < thr, plus the first need_eq entries with key == thr
    unsigned thr = 0xffffffffu; int need_eq = 0;
    const bool overflow = cnt > OBS_CAND;
    const int C = min(cnt, OBS_CAND);
    bool take_all = cnt <= K;
    if (first_k_only) take_all = true;     // first K in index order (sim.cpp:261-279)

    if (!take_all) {
        // exact radix select (4 x 8 bits, MSB first) of the K-th smallest key among in-radius entities
        unsigned prefix = 0; int remaining = K;
        for (int pass = 0; pass < 4; pass++) {
            const int shift = 24 - 8 * pass;
            __syncwarp();
            for (int i = lane; i < 256; i += 32) wp.u.hist[i] = 0;
            __syncwarp();
            const unsigned hi_mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
            if (!overflow) {
                for (int i = lane; i < C; i += 32) {
                    const unsigned kb = wp.cand[i].x;
                    if ((kb & hi_mask) == prefix) atomicAdd(&wp.u.hist[(kb >> shift) & 255], 1u);
                }
            } else {   // candidate list overflowed: recompute keys from the pooled road table
                for (int r = lane; r < R; r += 32) {
                    const float4 ra = sc.road_a[r];
                    float ox, oy;
                    inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
                    const float key = len2_xy(ox, oy);
                    if (__fsqrt_rn(key) > radius) continue;
                    const unsigned kb = __float_as_uint(key);
                    if ((kb & hi_mask) == prefix) atomicAdd(&wp.u.hist[(kb >> shift) & 255], 1u);
                }
            }
            __syncwarp();
            // locate the bin holding the remaining-th element: lane owns bins [8*lane, 8*lane+8)
            unsigned local[8]; unsigned sum = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) { local[k] = wp.u.hist[lane * 8 + k]; sum += local[k]; }
            unsigned incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned v = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += v;
            }
            const unsigned excl = incl - sum;
            const bool mine = (unsigned)remaining > excl && (unsigned)remaining <= incl;
            int digit = 0; int before = 0;
            if (mine) {
                unsigned acc = excl;
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    if ((unsigned)remaining > acc && (unsigned)remaining <= acc + local[k]) { digit = lane * 8 + k; before = (int)acc; }
                    acc += local[k];
                }
            }
            const unsigned who = __ballot_sync(FULL, mine);
            const int src = __ffs(who) - 1;
            digit = __shfl_sync(FULL, digit, src);
            before = __shfl_sync(FULL, before, src);
            prefix |= (unsigned)digit << shift;
            remaining -= before;
        }
        thr = prefix;
        need_eq = remaining;       // entries equal to thr still to take, lowest road index first
        __syncwarp();
    }

    // emission in road-index order
    {
        int eq_seen = 0;
        const int N = overflow && !take_all ? R : C;
        for (int base = 0; base < N && emitted < K; base += 32) {
            const int i = base + lane;
            bool acc = false; unsigned ridx = 0; unsigned kb = 0;
            if (i < N) {
                if (overflow && !take_all) {
                    const float4 ra = sc.road_a[i];
                    float ox, oy;
                    inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
                    const float key = len2_xy(ox, oy);
                    acc = !(__fsqrt_rn(key) > radius);
                    kb = __float_as_uint(key); ridx = (unsigned)i;
                } else {
                    const uint2 cd = wp.cand[i];
                    acc = true; kb = cd.x; ridx = cd.y;
                }
            }
            if (!take_all) {
                const bool is_eq = acc && kb == thr;
                const unsigned em = __ballot_sync(FULL, is_eq);
                const int eq_rank = eq_seen + __popc(em & lt_mask);
                acc = acc && (kb < thr || (is_eq && eq_rank < need_eq));
                eq_seen += __popc(em);
            }
            const unsigned am = __ballot_sync(FULL, acc);
            int rank = __popc(am & lt_mask);
            int m = __popc(am);
            if (emitted + m > K) { m = K - emitted; if (rank >= m) acc = false; }
            float row[9];
            if (acc) {
                const float4 ra = sc.road_a[ridx]; const float2 rb = sc.road_b[ridx];
                const float4 rd = sc.road_d[ridx]; const uint32_t rc = sc.road_c[ridx];
                float ox, oy, rw, rz;
                inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
                inv_mul_quat(qw, qz, ra.z, ra.w, rw, rz);
                row[0] = ox; row[1] = oy; row[2] = rb.x; row[3] = rb.y; row[4] = rd.x;
                row[5] = quat_to_yaw(rw, rz); row[6] = (float)(rc >> 8); row[7] = rd.z; row[8] = rd.w;
                if (write_knn_idx) kidx[emitted + rank] = (int32_t)ridx;
            }
            flush_rows(wp.u.stage, rows + (size_t)emitted * 9, row, acc, rank, m, lane);
            emitted += m;
        }
    }
    // padding rows: fillZeros (knn.hpp:19-28 -> id = 0, mapType = 0) for the k-nearest algorithm,
    // MapObservation::zero() (id = -1, mapType = -1) for the linear one (sim.cpp:277-279)
    {
        const float padv = first_k_only ? -1.f : 0.f;
        const int n = (K - emitted) * 9;
        float *dst = rows + (size_t)emitted * 9;
        for (int i = lane; i < n; i += 32) dst[i] = (i % 9) >= 7 ? padv : 0.f;
        if (write_knn_idx)
            for (int i = emitted + lane; i < K; i += 32) kidx[i] = -1;
    }
}

} // namespace gd
