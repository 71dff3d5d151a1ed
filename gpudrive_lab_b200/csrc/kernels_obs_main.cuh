// k_obs kernel body (helpers in kernels_obs.cuh).
#pragma once
#include "kernels_obs.cuh"
#include "kernels_policy.cuh"

namespace gd {

#ifndef GD_OBS_MINB
#define GD_OBS_MINB 4
#endif

// key_thr: largest float k with sqrtf(k) <= observationRadius, so `key <= key_thr` is exactly the
// reference's `length() <= radius` (knn.hpp:88, sim.cpp:268-271) without a square root per entity.
// POL: also write the fused policy observation (second output, `pol`); k_obs<false> carries none of that code.
template <bool POL>
__global__ void __launch_bounds__(OBS_BLOCK, GD_OBS_MINB)
k_obs(const SceneDev *__restrict__ scenes, StateDev st, Exports ex, Params prm, int S, int groups,
      int write_knn_idx, float key_thr, int only_flagged, float *__restrict__ pol)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ObsSmem &sm = *reinterpret_cast<ObsSmem *>(smem_raw);

    const int w = blockIdx.x / groups;
    const int g = blockIdx.x % groups;
    if (only_flagged && !st.reset_flag[w]) return;      // auto-reset: only the worlds that were just re-bound
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
    }
    __syncthreads();

    const int turn = g * OBS_WARPS + warp;
    if (turn >= A) return;                // no block-wide barrier below this point
#ifndef GD_OBS_NO_COSTSORT
    const unsigned clk0 = (unsigned)clock();
    const int a = st.order[(size_t)w * S + turn];      // agents in order of decreasing cost (k_agents)
#else
    const int a = turn;
#endif
    const bool active = true;
    ObsWarpSmem &wp = sm.wp[warp];
    const size_t wa = (size_t)w * S + (active ? a : 0);

    float px = 0, py = 0, pz = 0, qw = 1, qz = 0;
    if (active) {
        const float4 s0 = sm.a0[a];
        px = s0.x; py = s0.y; qw = s0.z; qz = s0.w; pz = sm.a1[a].x;
    }
    // optional second output: the flat normalised policy observation GPUDriveTorchEnv.get_obs()
    // assembles from these rows (kernels_policy.cuh), written from the rows while the warp still holds them
    // (the address is recomputed where it is used: a pointer held across the kernel costs two of the 64 registers)
    auto pol_row0 = [&]() { return pol + wa * (size_t)(kPolicyEgo + (S - 1) * kPolicyPartner + kMaxAgentMapObs * kPolicyRoad); };
    const float radius = prm.observationRadius;
    const bool classical = !prm.disableClassicalObs;
    const bool lidar_on = prm.enableLidar != 0;

    // ------------------------------------------------------------ self + absolute observations
    if (active && lane == 0) {
        const float4 s1 = sm.a1[a], s3 = sc.ag_a[a];
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
        ab[3] = make_float2(qz, st.yaw[wa]);                // quatToYaw(rot), computed once per step in k_agents
        ab[4] = make_float2(gc.x, gc.y);
        ab[5] = make_float2(s3.x, s3.y);
        ab[6] = make_float2(s3.z, s1.w);
        if (POL) {
            float o[kPolicyEgo];
            policy_ego_row<true>(s1.y, s3.x, s3.y, gx, gy, st.collided[wa] ? 1.f : 0.f, true, o);
            float *pe = pol_row0();
#pragma unroll
            for (int k = 0; k < kPolicyEgo; k++) pe[k] = o[k];
        }
    }

    // ------------------------------------------------------------ partner observations
    if (active && classical) {
        const int P = S - 1;
        float *dstbase = ex.partner_obs + wa * P * kPartnerObsSize;
        bool const_staged = false;
        for (int base = 0; base < P; base += 32) {
            const int m = min(32, P - base);
            float *dst = dstbase + (size_t)base * 9;
            if (base < A - 1) {       // (warp-uniform) straight-line inside: some lane takes each path anyway
                const int jp = base + lane;
                const bool has = jp < P;
                const bool exists = has && jp < A - 1;
                const int j = exists ? (jp < a ? jp : jp + 1) : 0;   // OtherAgents order, level_gen.cpp:450-464
                const float4 o0 = sm.a0[j], o1 = sm.a1[j];
                const float4 o3 = sc.ag_a[j];
                float rx, ry, rw, rz;
                inv_rotate_xy(qw, qz, __fsub_rn(o0.x, px), __fsub_rn(o0.y, py), rx, ry);
                inv_mul_quat(qw, qz, o0.z, o0.w, rw, rz);
                const bool seen = exists && !(__fsqrt_rn(len2_xy(rx, ry)) > radius);
                float row[9];
                row[0] = seen ? o1.y : 0.f; row[1] = seen ? rx : 0.f; row[2] = seen ? ry : 0.f;
                row[3] = seen ? quat_to_yaw(rw, rz) : 0.f;
                row[4] = seen ? o3.x : 0.f; row[5] = seen ? o3.y : 0.f; row[6] = seen ? o3.z : 0.f;
                row[7] = seen ? o1.z : 0.f;
                row[8] = seen ? o1.w : (exists ? -1.f : -2.f);       // PartnerObservation::zero() / zero_nonexist()
                flush_rows(wp.u.stage, dst, row, has, lane, m, lane);
                if (POL) {
                    float o[kPolicyPartner];
                    policy_partner_row<true>(row, true, o);
                    flush_rows_n<kPolicyPartner>(wp.u.stage, pol_row0() + kPolicyEgo + base * kPolicyPartner, o, has, lane, m, lane);
                }
            } else {
                // every row of the chunk is zero_nonexist(); chunks start 1152 bytes apart, so the staged image
                // (and its alignment) is the same for all of them: stage it once
                const int mis = (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u);
                if (!const_staged) {
#pragma unroll
                    for (int k = 0; k < 9; k++) wp.u.stage[mis + lane * 9 + k] = k == 8 ? -2.f : 0.f;
                    __syncwarp();
                    const_staged = !POL;            // (the policy rows below reuse the tile)
                }
                flush_aligned(wp.u.stage, dst, mis, m * 9, lane);
                if (POL) {
                    __syncwarp();
                    const float zrow[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, -2.f};
                    float o[kPolicyPartner];
                    policy_partner_row<true>(zrow, true, o);
                    flush_rows_n<kPolicyPartner>(wp.u.stage, pol_row0() + kPolicyEgo + base * kPolicyPartner, o, true, lane, m, lane);
                }
            }
        }
        __syncwarp();
    }

    // ------------------------------------------------------------ lidar set-up
    LidarCtx lc;
    lc.ox = px; lc.oy = py; lc.fc = 1.f; lc.fs = 0.f;
    int zmask_classes = 0;         // 3 plane bits per z-class (bits >= 24 stay 0: dummy class 8 has no plane)
    unsigned long long open[3] = {RAY_ALL, RAY_ALL, RAY_ALL};
    if (lidar_on) {
        __syncwarp();
#pragma unroll 1
        for (int i = lane; i < 3 * NRAY; i += 32) wp.u.l.bt[i] = 0xffffffffu;
        const int controlled = ex.controlled[wa];
        const float head = controlled ? ex.action[wa * kActionSize + 2] : 0.f;   // sim.cpp:409
        float rgx, rgy, fwx, fwy;
        rotate_xy(qw, qz, 1.f, 0.f, rgx, rgy);     // rot.rotateVec(math::right)
        rotate_xy(qw, qz, 0.f, 1.f, fwx, fwy);     // rot.rotateVec(math::fwd)
#pragma unroll 1
        for (int i = lane; i < NRAY; i += 32) {
            // (ray 0: 0 / NRAY = 0 without the slow path IEEE division takes for a zero operand)
            const float frac = i == 0 ? 0.f : __fdiv_rn(i == 0 ? 1.f : 2.f * (float)i, (float)NRAY);
            const float theta = __fadd_rn(__fmul_rn(kLidarAngle, __fsub_rn(frac, 1.f)), head);
            // sim.cpp:411-412, with glibc's own cosf / sinf (one-pass restatement, bit-identical). A direction that
            // is 1 ulp off moves the depth of a ray that runs nearly parallel to a long thin box by far more than
            // 1e-5 (CUDA's sincosf, <= 2 ulp away, cost 7 of 2.4 M rays their parity on the benchmarked scenes).
            float x, y;
            sincos_hp(theta, y, x);
            const float dx = __fadd_rn(__fmul_rn(x, rgx), __fmul_rn(y, fwx));
            const float dy = __fadd_rn(__fmul_rn(x, rgy), __fmul_rn(y, fwy));
            const float inv = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(len2_xy(dx, dy), 0.f)));
            wp.v.g.raydir[i] = make_float2(__fmul_rn(dx, inv), __fmul_rn(dy, inv));
            wp.v.g.rayxy[i] = make_float2(x, y);
        }
        const float yaw = st.yaw[wa];
        const float delta = (2.f * kLidarAngle) / (float)NRAY;
        const float mid = yaw + head - kLidarAngle + 0.5f * (float)(NRAY - 1) * delta;
        sincosf(mid, &lc.fs, &lc.fc);
        {   // one lane per z-class, combined with a warp OR
            int mine = 0;
            if (lane < sc.numZClasses) mine = plane_mask(pz, sc.zc_z[lane], sc.zc_d2[lane], -1.f, 1.f) << (3 * lane);
            zmask_classes = (int)__reduce_or_sync(FULL, (unsigned)mine);
        }
        __syncwarp();
        // other agents first: few boxes, and their hits close rays before the road groups are visited
#pragma unroll 1
        for (int base = 0; base < A; base += 32) {
            const int j = base + lane;
            int pm = 0, lo = 0, n = 0;
            float relx = 0.f, rely = 0.f, cphi = 1.f, sphi = 0.f, d0 = 1.f, d1 = 1.f;
            if (j < A) {
                const float4 o0 = sm.a0[j]; const float4 gb = sc.ag_b[j];
                pm = plane_mask(pz, sm.a1[j].x, gb.z, 0.f, 2.f);
                relx = o0.x - px; rely = o0.y - py; d0 = gb.x; d1 = gb.y;
                cphi = 1.f - 2.f * o0.w * o0.w; sphi = 2.f * o0.z * o0.w;
                int hi;
                // the agent's own hull is skipped: the origin is its centre, so every face is a back face -> miss
                if (pm && j != a && lidar_consider(lc, relx, rely, cphi, sphi, d0, d1, lo, hi)) n = hi - lo + 1;
            }
            lidar_expand(wp, lane, lo, n, pm, 0x4000u | (unsigned)j, relx, rely, cphi, sphi, d0, d1);
        }
        __syncwarp();
    }

    // ------------------------------------------------------------ road groups: kNN candidates + lidar
    int cnt = 0;                   // in-radius entities seen so far (warp-uniform)
    const bool want_cands = classical;
    const bool first_k_only = prm.roadObservationAlgorithm == RO_AllEntities;
    const float rpad = radius * 1.00001f + 1e-3f;      // |dx|,|dy| beyond this cannot be in radius
    const int G = sc.numGroups;
    for (int g0 = 0; g0 < G && (want_cands || lidar_on); g0 += OBS_GCAP) {
        // ---- one lane per group: can its bounding circle matter to this agent?
        const int gN = min(OBS_GCAP, G - g0);
        int nlist = 0;
        unsigned band_planes = 0, band_near = 0;     // per band: lidar planes / kNN interest of its groups
        __syncwarp();
        for (int base = 0; base < gN; base += 32) {
            const int gl = base + lane;
            const bool valid = gl < gN;
            // straight-line: selects instead of branches (a warp has lanes on every path)
            const float4 gc = __ldg(&sc.grp[2 * (g0 + (valid ? gl : 0))]);
            const float4 gb = __ldg(&sc.grp[2 * (g0 + (valid ? gl : 0)) + 1]);
            const float relx = gc.x - px, rely = gc.y - py;
            const float d2 = fmaf(relx, relx, rely * rely);
            const float d = sqrtf(d2);
            // nearest point of the group: distance to its oriented rectangle (tighter than the circle for the
            // long thin runs of a polyline); the circle serves the angular tests below
            const float gex = fmaxf(fabsf(fmaf(relx, gb.x, rely * gb.y)) - gb.z, 0.f);
            const float gey = fmaxf(fabsf(fmaf(rely, gb.x, -(relx * gb.y))) - gb.w, 0.f);
            // (inside the rectangle gex = gey = 0, and sqrtf(0) takes its slow path: 1e-10 m decides nothing below)
            const float dmin = fmaxf(sqrtf(fmaxf(fmaf(gex, gex, gey * gey), 1e-20f)), fmaxf(d - gc.z, 0.f));
            const bool nearg = valid && want_cands && dmin <= rpad;
            const unsigned cm = __float_as_uint(gc.w) & 0xffu;       // one z-class per group (classes are padded to 32)
            int pmg = (lidar_on && valid && cm && dmin <= kLidarDistance) ? (zmask_classes >> (3 * (__ffs(cm) - 1))) & 7 : 0;
            const float fx = fmaf(lc.fc, relx, lc.fs * rely), fy = fmaf(lc.fc, rely, -(lc.fs * relx));
            const bool apart = d > gc.z * 1.0001f + 0.01f;          // the origin is outside the circle
            const float cH = 0.5146014f, sH = 0.8574295f;            // fan half angle, as in lidar_consider
            const bool outside = fx < -gc.z || fmaf(cH, fy, -(sH * fx)) > gc.z || fmaf(cH, fy, sH * fx) < -gc.z;
            const float sn = gc.z * rcp_fast(fmaxf(d, 1e-20f));                 // sin of the half width
            const float w = sn * rsqrtf(fmaxf(1.f - sn * sn, 1e-6f));           // tan >= the angle itself
            int lo2, hi2;
            const bool some = ray_range(fast_atan2(fy, fx), w * 1.0001f + 2e-3f, lo2, hi2);
            const bool narrow = apart && w < 3.f;
            if (apart && (outside || (narrow && !some))) pmg = 0;
            const int lo = narrow ? lo2 : 0, hi = narrow ? hi2 : NRAY - 1;
            // (measured and dropped: the angular span of the rectangle's corners instead of the circle's -
            //  four more atan2 per group cost more than the skipped groups save, 7.62 vs 7.40 ms)
            const bool keep = nearg || pmg != 0;
            const int e = (int)(__float_as_uint(dmin) >> 23) - 127;           // floor(log2(dmin))
            const int band = obs_band_of(dmin, e);
            const uint32_t desc = !keep ? 0u : (uint32_t)gl | ((uint32_t)band << 12) | ((uint32_t)lo << 15) | ((uint32_t)hi << 21) |
                                  (nearg ? (1u << 27) : 0u) | ((uint32_t)pmg << 28);
            const unsigned km = __ballot_sync(FULL, keep);
            if (keep) wp.v.g.glist[nlist + __popc(km & lt_mask)] = desc;
            nlist += __popc(km);
            const unsigned bsh = (desc >> 12) & 7u;
            band_planes |= __reduce_or_sync(FULL, keep ? ((desc >> 28) << (3u * bsh)) : 0u);
            band_near |= __reduce_or_sync(FULL, (keep && ((desc >> 27) & 1u)) ? (1u << bsh) : 0u);
        }
        __syncwarp();
        GD_CNT(0, 1); GD_CNT(1, nlist);
        // ---- bands of increasing distance
        for (int band = 0; band < NBAND; band++) {
            const float band_lo = obs_band_lo(band);
            // rays with no hit nearer than the band: only these can still be improved from here on
            if (lidar_on) open_rays(wp.u.l.bt, band_lo * 0.9999f - 1e-3f, lane, open);
            const bool any_open = lidar_on && (open[0] | open[1] | open[2]) != 0ull;
            if (!any_open && !(want_cands && band_lo <= rpad)) break;     // nothing farther can matter
            {   // no group of this band is near, and none lies on a plane that still has open rays: skip it
                const unsigned planes_open = (open[0] ? 1u : 0u) | (open[1] ? 2u : 0u) | (open[2] ? 4u : 0u);
                if (!(((band_planes >> (3 * band)) & planes_open) || ((band_near >> band) & 1u))) continue;
            }
            for (int base = 0; base < nlist; base += 32) {
                const int li = base + lane;
                const uint32_t desc = li < nlist ? wp.v.g.glist[li] : 0u;
                const bool mine = li < nlist && (int)((desc >> 12) & 7u) == band;
                if (!__any_sync(FULL, mine)) continue;       // no group of this band in this chunk of the list
                const int pmg = (int)(desc >> 28);
                const bool lid = mine && pmg != 0 &&
                    (ray_bits((desc >> 15) & 63u, (desc >> 21) & 63u) & open_for(open, pmg)) != 0ull;
                const bool nr = mine && ((desc >> 27) & 1u);
                unsigned todo = __ballot_sync(FULL, lid || nr);
                const unsigned lidm = __ballot_sync(FULL, lid), nrm = __ballot_sync(FULL, nr);
                while (todo) {
                    const int src = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const uint32_t gdesc = __shfl_sync(FULL, desc, src);
                    const int gid = g0 + (int)(gdesc & 0xfffu);
                    const bool do_l = (lidm >> src) & 1u;
                    const bool do_n = (nrm >> src) & 1u;
                    const int e = gid * 32 + lane;
                    GD_CNT(2, 1); GD_CNT(3, do_l ? 1 : 0); GD_CNT(4, do_n ? 1 : 0);
                    // all three columns up front, so their latencies overlap (the kNN-only passes pay two
                    // loads they do not need; they hit L1 / L2)
                    const float4 ra = __ldg(&sc.s_a[e]);
                    const uint32_t c = __ldg(&sc.s_c[e]);
                    const float2 rb = __ldg(&sc.s_b[e]);
                    if (do_n) {
                        const float dx = __fsub_rn(ra.x, px), dy = __fsub_rn(ra.y, py);
                        float ox, oy;
                        inv_rotate_xy(qw, qz, dx, dy, ox, oy);
                        const float key = len2_xy(ox, oy);
                        const bool in = key <= key_thr;         // (a far-away dummy slot gives inf: never in)
                        const unsigned inm = __ballot_sync(FULL, in);
                        const int pos = cnt + __popc(inm & lt_mask);
                        if (in && pos < OBS_CAND) {
                            wp.ckey[pos] = first_k_only ? (c >> 16) : __float_as_uint(key);
                            wp.cpos[pos] = (uint16_t)e;
                        }
                        cnt += __popc(inm);
                    }
                    if (do_l) {
                        // a group holds one z-class, so its plane mask and the open rays of those planes are the same
                        // for all 32 boxes (warp-uniform, from the descriptor); a dummy slot is far away: never alive
                        const int pm = (int)(gdesc >> 28);
                        const unsigned long long om = open_for(open, pm);
                        const float relx = ra.x - px, rely = ra.y - py;
                        const float cphi = 1.f - 2.f * ra.w * ra.w, sphi = 2.f * ra.z * ra.w;
                        // cheap tests first: about one pass in five has no box left after them
                        const BoxCull bc = lidar_cull(lc, relx, rely, cphi, sphi, rb.x, rb.y);
                        if (__any_sync(FULL, bc.alive)) {
                            int lo, hi;
                            const bool some = lidar_range(bc, lo, hi);
                            // trim the range to its first / last open ray (closed rays inside are clipped too)
                            const unsigned long long m = some ? (ray_bits(lo, hi) & om) : 0ull;
                            lo = m ? __ffsll((long long)m) - 1 : 0;
                            const int n = m ? 64 - __clzll((long long)m) - lo : 0;
                            lidar_expand(wp, lane, lo, n, pm, (unsigned)e, relx, rely, cphi, sphi, rb.x, rb.y);
                        }
                    }
                }
            }
            __syncwarp();
        }
    }

    GD_CNT(8, cnt); GD_CNT(10, cnt > OBS_CAND ? 1 : 0);
    // ------------------------------------------------------------ lidar write-out
    if (lidar_on) {
        __syncwarp();
        float4 *out = reinterpret_cast<float4 *>(ex.lidar + wa * 3 * NRAY * kLidarRowSize);
#pragma unroll 1
        for (int i = lane; i < 3 * NRAY; i += 32) {
            const uint32_t tbits = wp.u.l.bt[i];
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            bool retrace = false;
            if (tbits != 0xffffffffu) {
                float t;
                const unsigned id = wp.u.l.bid[i];
                float bx, by, bqw, bqz, d0, d1; int type;
                if (id & 0x4000u) {
                    const int j = id & 0x3fff;
                    const float4 o0 = sm.a0[j]; const float4 gb = sc.ag_b[j];
                    bx = o0.x; by = o0.y; bqw = o0.z; bqz = o0.w; d0 = gb.x; d1 = gb.y; type = (int)gb.w;
                } else {
                    const int e = id & 0x3fff;
                    const float4 ra = __ldg(&sc.s_a[e]); const float2 rb = __ldg(&sc.s_b[e]);
                    bx = ra.x; by = ra.y; bqw = ra.z; bqz = ra.w; d0 = rb.x; d1 = rb.y;
                    type = (int)((__ldg(&sc.s_c[e]) >> 8) & 0xffu);
                }
                // A marginal winner (grazed corner, origin on a face, range limit) is not decided here: the ray goes to
                // k_lidar_fix, which re-traces it with the reference's exact clip over every box. (Doing the exact clip
                // here, and the re-trace as a call, cost every agent up to 25 % more time - 9.2 vs 7.7 ms per launch -
                // for a path that about one ray in 28 000 takes.)
                retrace = (id & 0x8000u) != 0;
                const float2 d = wp.v.g.raydir[i % NRAY];
                t = ray_box_exact_depth(px, py, bx, by, bqw, bqz, d0, d1, d.x, d.y);
                const float2 xy = wp.v.g.rayxy[i % NRAY];
                v = make_float4(t, (float)type, __fmul_rn(t, xy.x), __fmul_rn(t, xy.y));
            }
            if (!retrace) out[i] = v;
            else {
                const unsigned slot = atomicAdd(&st.retrace[0], 1u);
                if (slot < kRetraceCap) st.retrace[1 + slot] = (unsigned)(wa * (3 * NRAY) + i);
            }
        }
        __syncwarp();
    }

    // ------------------------------------------------------------ road-map rows
    if (!classical) {
#ifndef GD_OBS_NO_COSTSORT
        if (lane == 0) st.cost[wa] = (unsigned)clock() - clk0;
#endif
        return;
    }
    const int K = kMaxAgentMapObs;
    float *rows = ex.agent_roadmap + wa * K * kMapObsSize;
    int32_t *kidx = ex.knn_idx + wa * K;
    int emitted = 0;
    const int RS = G * 32;

    // candidate key of sorted slot e (slow path only): in radius? and its selection key
    auto slot_key = [&](int e, unsigned &kb) -> bool {
        const float4 ra = __ldg(&sc.s_a[e]);
        float ox, oy;
        inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
        const float key = len2_xy(ox, oy);
        kb = first_k_only ? (__ldg(&sc.s_c[e]) >> 16) : __float_as_uint(key);
        return key <= key_thr;
    };

    // selection: accept key < thr, plus the entries with key == thr whose road index is <= idx_thr.
    // k-nearest: key = squared-distance bits, ties by road index (the heap of knn.hpp keeps the earlier
    // index on equal distance); linear: key = road index, the first K in index order (sim.cpp:261-279).
    unsigned thr = 0xffffffffu, idx_thr = 0xffffffffu;
    // monotone 8-bit quantisation of a selection key
    const float qscale = key_thr > 0.f ? 255.9f / key_thr : 0.f;
    auto quant = [&](unsigned kb) -> int {
        return first_k_only ? min(255, (int)(kb >> 6)) : min(255, __float2int_rz(__uint_as_float(kb) * qscale));
    };
    // the bin of wp.v.hist that holds the k-th smallest entry (1-based), and how many entries lie in the bins below it
    auto find_bin = [&](int k, int &digit, int &before, int &inbin) {
        unsigned local[8]; unsigned sum = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) { local[q] = wp.v.hist[lane * 8 + q]; sum += local[q]; }
        unsigned incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += v;
        }
        const unsigned excl = incl - sum;
        const bool mine = (unsigned)k > excl && (unsigned)k <= incl;
        digit = 0; before = 0; inbin = 0;
        if (mine) {
            unsigned acc = excl;
#pragma unroll
            for (int q = 0; q < 8; q++) {
                if ((unsigned)k > acc && (unsigned)k <= acc + local[q]) { digit = lane * 8 + q; before = (int)acc; inbin = (int)local[q]; }
                acc += local[q];
            }
        }
        const int src = __ffs(__ballot_sync(FULL, mine)) - 1;
        digit = __shfl_sync(FULL, digit, src);
        before = __shfl_sync(FULL, before, src);
        inbin = __shfl_sync(FULL, inbin, src);
    };

    if (cnt > OBS_CAND) {
        // Dense map: more in-radius entities than the candidate buffer holds (a 10 000-entity scene has ~900 within
        // 50 m). Two more sweeps over the near groups only: a histogram of the quantised keys finds the bin of the
        // K-th smallest, then the entities up to that bin - K plus a few - are collected as the candidates, and
        // the normal path below takes over. Only if even those overflow does the general path scan the whole table
        // (it did for every agent of such a scene before: 71 ms per step on the collision-stress configuration).
        auto sweep_near = [&](auto &&visit) {
            for (int gb = 0; gb < G; gb += 32) {
                const int gi = gb + lane;
                bool nearg = false;
                if (gi < G) {
                    const float4 gc = __ldg(&sc.grp[2 * gi]), gr = __ldg(&sc.grp[2 * gi + 1]);
                    const float relx = gc.x - px, rely = gc.y - py;
                    const float gex = fmaxf(fabsf(fmaf(relx, gr.x, rely * gr.y)) - gr.z, 0.f);
                    const float gey = fmaxf(fabsf(fmaf(rely, gr.x, -(relx * gr.y))) - gr.w, 0.f);
                    const float dmin = fmaxf(sqrtf(fmaf(gex, gex, gey * gey)),
                                             fmaxf(sqrtf(fmaf(relx, relx, rely * rely)) - gc.z, 0.f));
                    nearg = dmin <= rpad;                     // the same test that listed the group above
                }
                for (unsigned gm = __ballot_sync(FULL, nearg); gm; gm &= gm - 1) {
                    const int e = (gb + __ffs(gm) - 1) * 32 + lane;
                    unsigned kb;
                    const bool in = slot_key(e, kb);
                    visit(e, kb, in);
                }
            }
        };
        __syncwarp();
#pragma unroll 1
        for (int i = lane; i < 256; i += 32) wp.v.hist[i] = 0;
        __syncwarp();
        sweep_near([&](int, unsigned kb, bool in) { if (in) atomicAdd(&wp.v.hist[quant(kb)], 1u); });
        __syncwarp();
        int digit, before, inbin;
        find_bin(K, digit, before, inbin);
        __syncwarp();
        if (before + inbin <= OBS_CAND) {
            int c2 = 0;
            sweep_near([&](int e, unsigned kb, bool in) {
                const bool take = in && quant(kb) <= digit;
                const unsigned tm = __ballot_sync(FULL, take);
                if (take) { const int pos = c2 + __popc(tm & lt_mask); wp.ckey[pos] = kb; wp.cpos[pos] = (uint16_t)e; }
                c2 += __popc(tm);
            });
            cnt = c2;                                         // >= K entries, every one of the K smallest among them
            __syncwarp();
        }
    }
    const bool overflow = cnt > OBS_CAND;
    const int C = min(cnt, OBS_CAND);
    const bool take_all = cnt <= K;
    const int N = overflow ? RS : C;

    bool selected = take_all;
    if (!take_all && !overflow) {
        // Fast path: one histogram pass over a monotone 8-bit quantisation of the key finds the bin that
        // holds the K-th smallest; its (few) members are then ranked exactly by (key, road index).
        __syncwarp();
#pragma unroll 1
        for (int i = lane; i < 256; i += 32) wp.v.hist[i] = 0;
        __syncwarp();
#pragma unroll 1
        for (int i = lane; i < C; i += 32) atomicAdd(&wp.v.hist[quant(wp.ckey[i])], 1u);
        __syncwarp();
        int digit, before, inbin;
        find_bin(K, digit, before, inbin);
        if (inbin <= 32) {
            uint32_t *mk = reinterpret_cast<uint32_t *>(wp.u.stage), *mi = mk + 32;
            int nm = 0;
            for (int base = 0; base < C; base += 32) {
                const int i = base + lane;
                const bool is = i < C && quant(wp.ckey[i]) == digit;
                const unsigned bm = __ballot_sync(FULL, is);
                if (is) {
                    const int slot = nm + __popc(bm & lt_mask);
                    mk[slot] = wp.ckey[i];
                    mi[slot] = __ldg(&sc.s_c[wp.cpos[i]]) >> 16;
                }
                nm += __popc(bm);
            }
            __syncwarp();
            const unsigned myk = lane < nm ? mk[lane] : 0xffffffffu, myi = lane < nm ? mi[lane] : 0xffffffffu;
            int less = 0;
            for (int m = 0; m < nm; m++) {
                const unsigned k2 = mk[m], i2 = mi[m];
                less += (k2 < myk || (k2 == myk && i2 < myi)) ? 1 : 0;
            }
            const int hit = __ffs(__ballot_sync(FULL, lane < nm && less == K - before - 1)) - 1;
            thr = __shfl_sync(FULL, myk, hit);
            idx_thr = __shfl_sync(FULL, myi, hit);
            selected = true;
            __syncwarp();
        }
    }
    if (!selected) {
        // general path (candidate overflow, or many equal keys): exact radix select (4 x 8 bits, MSB first)
        // of the K-th smallest key among in-radius entities
        unsigned prefix = 0; int remaining = K; int count_eq = 0;
        for (int pass = 0; pass < 4; pass++) {
            const int shift = 24 - 8 * pass;
            __syncwarp();
#pragma unroll 1
            for (int i = lane; i < 256; i += 32) wp.v.hist[i] = 0;
            __syncwarp();
            const unsigned hi_mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
#pragma unroll 1
            for (int i = lane; i < N; i += 32) {
                unsigned kb; bool in = true;
                if (overflow) in = slot_key(i, kb); else kb = wp.ckey[i];
                if (in && (kb & hi_mask) == prefix) atomicAdd(&wp.v.hist[(kb >> shift) & 255], 1u);
            }
            __syncwarp();
            // locate the bin holding the remaining-th element: lane owns bins [8*lane, 8*lane+8)
            unsigned local[8]; unsigned sum = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) { local[k] = wp.v.hist[lane * 8 + k]; sum += local[k]; }
            unsigned incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned v = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += v;
            }
            const unsigned excl = incl - sum;
            const bool mine = (unsigned)remaining > excl && (unsigned)remaining <= incl;
            int digit = 0, before = 0, inbin = 0;
            if (mine) {
                unsigned acc = excl;
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    if ((unsigned)remaining > acc && (unsigned)remaining <= acc + local[k]) {
                        digit = lane * 8 + k; before = (int)acc; inbin = (int)local[k];
                    }
                    acc += local[k];
                }
            }
            const unsigned who = __ballot_sync(FULL, mine);
            const int src = __ffs(who) - 1;
            digit = __shfl_sync(FULL, digit, src);
            before = __shfl_sync(FULL, before, src);
            count_eq = __shfl_sync(FULL, inbin, src);
            prefix |= (unsigned)digit << shift;
            remaining -= before;
        }
        thr = prefix;
        __syncwarp();
        if (remaining < count_eq) {
            // a true tie at the K-th key: take the `remaining` lowest road indices among the equal ones
            unsigned last = 0; bool first = true;
            for (int r = 0; r < remaining; r++) {
                unsigned best_idx = 0xffffffffu;
#pragma unroll 1
                for (int i = lane; i < N; i += 32) {
                    unsigned kb; bool in = true; int e = i;
                    if (overflow) in = slot_key(i, kb); else { kb = wp.ckey[i]; e = wp.cpos[i]; }
                    if (in && kb == thr) {
                        const unsigned oi = __ldg(&sc.s_c[e]) >> 16;
                        if ((first || oi > last) && oi < best_idx) best_idx = oi;
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) best_idx = min(best_idx, __shfl_xor_sync(FULL, best_idx, o));
                last = best_idx; first = false;
            }
            idx_thr = last;
        }
    }

    // One chunk of 32 candidates: is entry i selected, and which sorted slot / road word does it carry?
    auto candidate = [&](int i, int &e, uint32_t &rc) -> bool {
        bool acc = false; unsigned kb = 0;
        e = 0;
        if (i < N) {
            if (overflow) { e = i; acc = slot_key(i, kb); }
            else { acc = true; kb = wp.ckey[i]; e = wp.cpos[i]; }
        }
        rc = acc ? __ldg(&sc.s_c[e]) : 0u;
        if (!take_all) acc = acc && (kb < thr || (kb == thr && (rc >> 16) <= idx_thr));
        return acc;
    };
    auto make_row = [&](int e, uint32_t rc, float (&row)[9]) {
        // straight-line (most lanes of a chunk emit a row; e is a valid slot on the others)
        const float4 ra = __ldg(&sc.s_a[e]); const float2 rb = __ldg(&sc.s_b[e]);
        const float4 rd = __ldg(&sc.s_d[e]);
        float ox, oy, rw, rz;
        inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
        inv_mul_quat(qw, qz, ra.z, ra.w, rw, rz);
        row[0] = ox; row[1] = oy; row[2] = rb.x; row[3] = rb.y; row[4] = rd.x;
        row[5] = quat_to_yaw(rw, rz); row[6] = (float)((rc >> 8) & 0xffu); row[7] = rd.z; row[8] = rd.w;
    };

    // one chunk of <= 32 finished rows: the export rows, then (optionally) their policy rows
    auto pol_road0 = [&]() { return pol_row0() + kPolicyEgo + (S - 1) * kPolicyPartner; };
    auto emit_chunk = [&](const float (&row)[9], bool acc, int rank, int m, int at) {
        flush_rows(wp.u.stage, rows + (size_t)at * 9, row, acc, rank, m, lane);
        if (POL) {
            float o[kPolicyRoad];
            policy_road_row<true>(row, true, o);
            flush_rows_n<kPolicyRoad>(wp.u.stage, pol_road0() + (size_t)at * kPolicyRoad, o, acc, rank, m, lane);
        }
    };
    if (first_k_only) {
        // Linear algorithm: the reference emits the first K in-radius entities in road-index order
        // (sim.cpp:261-279), so the order is part of the contract. The selected road indices go into a bit map
        // (<= 10 000 bits, in the dead group-list / ray buffers) with a per-word prefix count; the row at output
        // position p is then the p-th set bit: rows are produced in output order, 32 positions at a time, from the
        // index-ordered road table, and leave through the same coalesced flush as the k-nearest rows.
        constexpr int BMW = (kMaxRoadEntityCount + 31) / 32;        // 313 words
        static_assert(sizeof(wp.v) >= BMW * 4, "bit map does not fit");
        static_assert(sizeof(wp.cpos) >= 320 * 2, "prefix table does not fit");
        uint32_t *bm = reinterpret_cast<uint32_t *>(&wp.v);
        uint16_t *pre = wp.cpos;                                    // (the candidates are dead once the bit map is built)
        __syncwarp();
#pragma unroll 1
        for (int i = lane; i < BMW; i += 32) bm[i] = 0u;
        __syncwarp();
        for (int base = 0; base < N; base += 32) {
            int e; uint32_t rc;
            if (candidate(base + lane, e, rc)) atomicOr(&bm[rc >> 21], 1u << ((rc >> 16) & 31u));
        }
        __syncwarp();
        {   // exclusive prefix of the per-word counts: lane owns words [10 * lane, 10 * lane + 10)
            unsigned cntw[10]; unsigned sum = 0;
#pragma unroll
            for (int k = 0; k < 10; k++) { const int wi = lane * 10 + k; cntw[k] = wi < BMW ? __popc(bm[wi]) : 0; sum += cntw[k]; }
            unsigned incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned v = __shfl_up_sync(FULL, incl, o);
                if (lane >= o) incl += v;
            }
            emitted = min(K, (int)__shfl_sync(FULL, incl, 31));
            __syncwarp();                                           // every lane has read the candidates it needed
            unsigned acc = incl - sum;
#pragma unroll
            for (int k = 0; k < 10; k++) { pre[lane * 10 + k] = (uint16_t)min(acc, 65535u); acc += cntw[k]; }
        }
        __syncwarp();
        for (int base = 0; base < emitted; base += 32) {
            const int pos = base + lane;
            const bool acc = pos < emitted;
            // the word that holds the pos-th set bit: the last one whose prefix count is <= pos
            int lo = 0, hi = BMW - 1;
#pragma unroll 1
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if ((int)pre[mid] <= pos) lo = mid; else hi = mid - 1;
            }
            const unsigned bit = __fns(bm[lo], 0u, (acc ? pos - (int)pre[lo] : 0) + 1);
            const int r = acc ? min(lo * 32 + (int)(bit & 31u), R - 1) : 0;
            float row[9];
            {   // the row of road entity r, from the index-ordered table (same values as make_row over the sorted one)
                const float4 ra = __ldg(&sc.road_a[r]); const float2 rb = __ldg(&sc.road_b[r]);
                const float4 rd = __ldg(&sc.road_d[r]);
                float ox, oy, rw, rz;
                inv_rotate_xy(qw, qz, __fsub_rn(ra.x, px), __fsub_rn(ra.y, py), ox, oy);
                inv_mul_quat(qw, qz, ra.z, ra.w, rw, rz);
                row[0] = ox; row[1] = oy; row[2] = rb.x; row[3] = rb.y; row[4] = rd.x;
                row[5] = quat_to_yaw(rw, rz); row[6] = (float)(__ldg(&sc.road_c[r]) >> 8); row[7] = rd.z; row[8] = rd.w;
            }
            if (acc && write_knn_idx) kidx[pos] = (int32_t)r;
            emit_chunk(row, acc, lane, min(32, emitted - base), base);
        }
    } else
    {
        // k-nearest: the reference leaves the rows in heap order; here by distance band, then by sorted slot
        for (int base = 0; base < N && emitted < K; base += 32) {
            int e; uint32_t rc;
            bool acc = candidate(base + lane, e, rc);
            const unsigned am = __ballot_sync(FULL, acc);
            const int rank = __popc(am & lt_mask);
            int m = __popc(am);
            if (emitted + m > K) { m = K - emitted; if (rank >= m) acc = false; }
            float row[9];
            make_row(e, rc, row);
            if (acc && write_knn_idx) kidx[emitted + rank] = (int32_t)(rc >> 16);
            emit_chunk(row, acc, rank, m, emitted);
            emitted += m;
        }
    }
    {
        // padding rows: fillZeros (knn.hpp:19-28 -> all zero, id = 0, mapType = 0) for the k-nearest algorithm,
        // MapObservation::zero() (id = -1, mapType = -1) for the linear one (sim.cpp:277-279). An agent's block of
        // rows is 16-byte aligned (7 200 bytes per agent): <= 3 scalars up to the boundary, then STG.128
        const float padv = first_k_only ? -1.f : 0.f;
        const int n = (K - emitted) * 9, off = emitted * 9;
        float *dst = rows + (size_t)off;
        auto padval = [&](int i) { return ((off + i) % 9) >= 7 ? padv : 0.f; };
        const int head = min((4 - (off & 3)) & 3, n);
        if (lane < head) dst[lane] = padval(lane);
        float4 *d4 = reinterpret_cast<float4 *>(dst + head);
        const int nq = (n - head) >> 2;              // (K * 9 is a multiple of 4: no tail)
        if (first_k_only) {
#pragma unroll 1
            for (int i = lane; i < nq; i += 32) {
                const int j = head + 4 * i;
                d4[i] = make_float4(padval(j), padval(j + 1), padval(j + 2), padval(j + 3));
            }
        } else {
#pragma unroll 1
            for (int i = lane; i < nq; i += 32) d4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (POL) {
            // the policy row of a padding row: type 0 either way, everything else the image of zero
            const float zrow[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, padv, padv};
            float o[kPolicyRoad];
            policy_road_row<true>(zrow, true, o);
            float *pd = pol_road0() + (size_t)emitted * kPolicyRoad;
            const int np = (K - emitted) * kPolicyRoad;
#pragma unroll 1
            for (int i = lane; i < np; i += 32) {
                const int k = i % kPolicyRoad;
                pd[i] = k < 2 ? o[0] : (k == 6 ? o[6] : o[2]);      // (o[1] == o[0]; o[2..5] and o[7..12] are equal too)
            }
        }
    }
    if (write_knn_idx)
#pragma unroll 1
        for (int i = emitted + lane; i < K; i += 32) kidx[i] = -1;
#ifndef GD_OBS_NO_COSTSORT
    if (lane == 0) st.cost[wa] = (unsigned)clock() - clk0;
#endif
}

} // namespace gd
