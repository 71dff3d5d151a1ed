// Device restatement of the reference's scalar math for agents that only ever rotate about +z
// (Quat::angleAxis(yaw, up) => q.x = q.y = 0). Operation order follows the reference so that, with
// FMA contraction off on both sides, results are bit-identical given identical inputs:
//   src/utils.hpp:11-25, external/madrona/include/madrona/math.inl:659-691,837-845.
// The translation unit is compiled with --fmad=false; the *_rn intrinsics are used where an
// expression must never be contracted even if that flag is relaxed later.
#pragma once
#include <cuda_runtime.h>
#include "gd_types.h"
#include "glibc_trig.h"

namespace gd {

// Per-agent trigonometry. sinf / cosf reproduce glibc's results bit for bit (glibc_trig.h); the
// remaining functions go through fp64 and are rounded once: (float)atan2((double)y, (double)x) is
// what the reference's g++ build computes (its unqualified atan2 / fmod bind to the double
// overloads), and fp64 tan / atan rounded to float equal glibc's tanf / atanf for all but a
// vanishing fraction of inputs (Classic model only). These run once per agent per step, never
// per road entity.
__device__ __forceinline__ float cos_hp(float x) {
    return gdtrig::in_fast_range(x) ? gdtrig::cosf_glibc(x) : (float)cos((double)x);
}
__device__ __forceinline__ float sin_hp(float x) {
    return gdtrig::in_fast_range(x) ? gdtrig::sinf_glibc(x) : (float)sin((double)x);
}
// sin and cos of one angle with glibc's results: the one-pass restatement where it is valid, fp64 beyond
// (|x| >= 120: only a caller-supplied head angle can get there; kept out of line, it is never on a hot path)
__device__ __noinline__ void sincos_far(float x, float *s, float *c) {
    *s = (float)sin((double)x); *c = (float)cos((double)x);
}
__device__ __forceinline__ void sincos_hp(float x, float &s, float &c) {
    if (gdtrig::in_fast_range(x)) gdtrig::sincosf_glibc(x, s, c);
    else sincos_far(x, &s, &c);
}
__device__ __forceinline__ float tan_hp(float x) { return (float)tan((double)x); }
__device__ __forceinline__ float atan_hp(float x) { return (float)atan((double)x); }
__device__ __forceinline__ float atan2_hp(float y, float x) { return (float)atan2((double)y, (double)x); }

// utils::quatToYaw with x = y = 0 (src/utils.hpp:20-25). hp: per-agent state; fast: observation rows.
__device__ __forceinline__ float quat_to_yaw_hp(float w, float z) {
    float a = 2.0f * __fmul_rn(w, z);
    float b = __fsub_rn(1.0f, 2.0f * __fmul_rn(z, z));
    return atan2_hp(a, b);
}
// atan2 for observation rows: degree-13 odd minimax polynomial on [0, 1] + octant folding, |error| < 5e-7 rad
// (CUDA's atan2f is ~2 ulp but three times the instructions; rows tolerate 1e-5). Never used for agent state.
__device__ __forceinline__ float atan2_row(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = mx > 0.f ? __fdividef(mn, mx) : 0.f;
    const float s = a * a;
    float p = 0.006811790633946657f;
    p = fmaf(p, s, -0.0336042158305645f);
    p = fmaf(p, s, 0.07962366938591003f);
    p = fmaf(p, s, -0.1323334276676178f);
    p = fmaf(p, s, 0.19807815551757812f);
    p = fmaf(p, s, -0.3331736922264099f);
    p = fmaf(p, s, 0.9999961256980896f);
    float r = p * a;
    if (ay > ax) r = 1.57079637f - r;
    if (x < 0.f) r = 3.14159274f - r;
    return y < 0.f ? -r : r;
}
__device__ __forceinline__ float quat_to_yaw(float w, float z) {
    float a = 2.0f * __fmul_rn(w, z);
    float b = __fsub_rn(1.0f, 2.0f * __fmul_rn(z, z));
    return atan2_row(a, b);
}

// utils::NormalizeAngle<float> / AngleAdd (src/utils.hpp:11-18); fmod is exact in any precision.
__device__ __forceinline__ float normalize_angle(float angle) {
    const float ret = fmodf(angle, kPiM2);
    return ret > kPi ? __fsub_rn(ret, kPiM2) : (ret < -kPi ? __fadd_rn(ret, kPiM2) : ret);
}
__device__ __forceinline__ float angle_add(float a, float b) { return normalize_angle(__fadd_rn(a, b)); }

// Quat::angleAxis(angle, up): (cos(angle/2), 0, 0, sin(angle/2))
__device__ __forceinline__ void angle_axis_z_hp(float angle, float &qw, float &qz) {
    float h = angle / 2.f;
    qw = cos_hp(h);
    qz = sin_hp(h);
}

// rot.inv().rotateVec({vx, vy, 0}).xy() for rot = (w, 0, 0, z)   (math.inl:659-678)
// inv() negates the vector part: pure = (0, 0, -z). With c = cross(pure, v) = (z*vy, -(z*vx), 0) and
// cross(pure, c) = (z*c.y, -(z*c.x), 0):
//   out = v + 2 * ((c * w) + cross(pure, c))
__device__ __forceinline__ void inv_rotate_xy(float w, float z, float vx, float vy, float &ox, float &oy) {
    const float pz = -z;
    const float cx = -__fmul_rn(pz, vy);
    const float cy = __fmul_rn(pz, vx);
    const float ccx = -__fmul_rn(pz, cy);
    const float ccy = __fmul_rn(pz, cx);
    ox = __fadd_rn(vx, 2.f * __fadd_rn(__fmul_rn(cx, w), ccx));
    oy = __fadd_rn(vy, 2.f * __fadd_rn(__fmul_rn(cy, w), ccy));
}

// rot.rotateVec({vx, vy, 0}).xy() for rot = (w, 0, 0, z)
__device__ __forceinline__ void rotate_xy(float w, float z, float vx, float vy, float &ox, float &oy) {
    inv_rotate_xy(w, -z, vx, vy, ox, oy);
}

// (rot.inv() * other) for z-only quaternions -> (w, z) of the product (math.inl:837-845)
__device__ __forceinline__ void inv_mul_quat(float w, float z, float ow, float oz, float &rw, float &rz) {
    const float az = -z;
    rw = __fsub_rn(__fmul_rn(w, ow), __fmul_rn(az, oz));
    rz = __fadd_rn(__fmul_rn(w, oz), __fmul_rn(az, ow));
}

__device__ __forceinline__ float len2_xy(float x, float y) {
    return __fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y));
}
__device__ __forceinline__ float len3(float x, float y, float z) {
    return __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
}

// 2-D oriented box as OrientedBoundingBox2D::from + updateAxes build it (src/obb.hpp:12-50)
struct Obb2 {
    float cx[4], cy[4];   // corners
    float ax[2], ay[2];   // axes scaled by 1/len^2
    float org[2];         // corners[0] . axes[a]
};

__device__ __forceinline__ void obb_from(float px, float py, float c, float s, float d0, float d1, Obb2 &o) {
    const float Xx = __fmul_rn(c, d0), Xy = __fmul_rn(s, d0);
    const float Yx = __fmul_rn(-s, d1), Yy = __fmul_rn(c, d1);
    o.cx[0] = __fsub_rn(__fsub_rn(px, Xx), Yx); o.cy[0] = __fsub_rn(__fsub_rn(py, Xy), Yy);
    o.cx[1] = __fsub_rn(__fadd_rn(px, Xx), Yx); o.cy[1] = __fsub_rn(__fadd_rn(py, Xy), Yy);
    o.cx[2] = __fadd_rn(__fadd_rn(px, Xx), Yx); o.cy[2] = __fadd_rn(__fadd_rn(py, Xy), Yy);
    o.cx[3] = __fadd_rn(__fsub_rn(px, Xx), Yx); o.cy[3] = __fadd_rn(__fsub_rn(py, Xy), Yy);
    float a0x = __fsub_rn(o.cx[1], o.cx[0]), a0y = __fsub_rn(o.cy[1], o.cy[0]);
    float a1x = __fsub_rn(o.cx[3], o.cx[0]), a1y = __fsub_rn(o.cy[3], o.cy[0]);
    float inv0 = __fdiv_rn(1.f, len2_xy(a0x, a0y));
    float inv1 = __fdiv_rn(1.f, len2_xy(a1x, a1y));
    o.ax[0] = __fmul_rn(a0x, inv0); o.ay[0] = __fmul_rn(a0y, inv0);
    o.ax[1] = __fmul_rn(a1x, inv1); o.ay[1] = __fmul_rn(a1y, inv1);
    o.org[0] = __fadd_rn(__fmul_rn(o.cx[0], o.ax[0]), __fmul_rn(o.cy[0], o.ay[0]));
    o.org[1] = __fadd_rn(__fmul_rn(o.cx[0], o.ax[1]), __fmul_rn(o.cy[0], o.ay[1]));
}

// a.overlaps(b): projects b's corners on a's axes (src/obb.hpp:51-82)
__device__ __forceinline__ bool obb_overlaps_1way(const Obb2 &a, const Obb2 &b) {
#pragma unroll
    for (int k = 0; k < 2; k++) {
        float t = __fadd_rn(__fmul_rn(b.cx[0], a.ax[k]), __fmul_rn(b.cy[0], a.ay[k]));
        float tMin = t, tMax = t;
#pragma unroll
        for (int c = 1; c < 4; c++) {
            t = __fadd_rn(__fmul_rn(b.cx[c], a.ax[k]), __fmul_rn(b.cy[c], a.ay[k]));
            if (t < tMin) tMin = t;
            else if (t > tMax) tMax = t;
        }
        if ((tMin > __fadd_rn(1.f, a.org[k])) || (tMax < a.org[k])) return false;
    }
    return true;
}
__device__ __forceinline__ bool obb_collided(const Obb2 &a, const Obb2 &b) {
    return obb_overlaps_1way(a, b) && obb_overlaps_1way(b, a);
}

} // namespace gd
