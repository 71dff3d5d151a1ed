// k_policy_obs: the flat, normalised policy observation GPUDriveTorchEnv.get_obs() assembles in Python from
// three exported tensors (gpudrive/env/env_torch.py:756-891,1172-1216 with the default ego_state, partner_obs,
// road_map_obs, norm_obs; gpudrive/datatypes/observation.py:29-92,170-258; gpudrive/datatypes/roadgraph.py:
// 281-362; gpudrive/env/constants.py). SURVEY.md section 8(f)-1: the step right after the hot path, done there
// by ~25 torch kernels that re-read the exports several times. One pass here: per agent slot
//   ego      6      speed/100, length*0.7/30, width*0.7/15, mm(goal x), mm(goal y), collided
//   partner  127*6  speed/100, mm(x), mm(y), heading/2pi, length*0.7/30, width*0.7/15
//   road     200*13 mm(x), mm(y), d0/100, d1/100, d2/100, heading/2pi, one_hot7(type)
// with mm(v) = 2 * ((v + 1000) / 2000) - 1. Same float32 operations in the same order as torch's, so the
// result is bit-identical to the Python path on the same input tensors.
#pragma once
#include "engine.h"

namespace gd {

constexpr int kPolicyEgo = 6, kPolicyPartner = 6, kPolicyRoad = 13;

// `tensor / python_scalar` in torch: an IEEE division on the CPU; on CUDA the tensor times the float reciprocal of the
// scalar (ATen/native/cuda/BinaryDivTrueKernel.cu: "a * reciprocal(b) ... may lose one bit"). The env's get_obs()
// runs on whichever device it was given, so both arithmetics exist: normalize = 1 reproduces the CPU path (what
// the golden vectors pin), normalize = 2 the CUDA path, each bit for bit.
template <bool RCP>
__device__ __forceinline__ float tdiv(float v, float c) {
    return RCP ? __fmul_rn(v, __fdiv_rn(1.f, c)) : __fdiv_rn(v, c);
}
template <bool RCP>
__device__ __forceinline__ float norm_min_max(float v) {      // normalize_min_max(v, -1000, 1000)
    return __fsub_rn(__fmul_rn(2.f, tdiv<RCP>(__fsub_rn(v, -1000.f), 2000.f)), 1.f);
}

// The three row formulas, shared by k_policy_obs below and by the fused second output of k_obs (which uses the
// CUDA arithmetic, RCP = true: the env that consumes it runs on the device).
template <bool RCP>
__device__ __forceinline__ void policy_ego_row(float speed, float length, float width, float gx, float gy,
                                               float collided, bool normalize, float (&o)[kPolicyEgo])
{
    const float len = __fmul_rn(length, 0.7f), wid = __fmul_rn(width, 0.7f);        // madrona_gpudrive.vehicleScale
    o[0] = normalize ? tdiv<RCP>(speed, 100.f) : speed;
    o[1] = normalize ? tdiv<RCP>(len, 30.f) : len;
    o[2] = normalize ? tdiv<RCP>(wid, 15.f) : wid;
    o[3] = normalize ? norm_min_max<RCP>(gx) : gx;
    o[4] = normalize ? norm_min_max<RCP>(gy) : gy;
    o[5] = collided;
}
// r: a PartnerObservation row (speed, x, y, heading, length, width, height, type, id)
template <bool RCP>
__device__ __forceinline__ void policy_partner_row(const float *r, bool normalize, float (&o)[kPolicyPartner])
{
    const float two_pi = 6.283185307179586f;        // float32(2 * np.pi)
    const float len = __fmul_rn(r[4], 0.7f), wid = __fmul_rn(r[5], 0.7f);
    o[0] = normalize ? tdiv<RCP>(r[0], 100.f) : r[0];
    o[1] = normalize ? norm_min_max<RCP>(r[1]) : r[1];
    o[2] = normalize ? norm_min_max<RCP>(r[2]) : r[2];
    o[3] = normalize ? tdiv<RCP>(r[3], two_pi) : r[3];
    o[4] = normalize ? tdiv<RCP>(len, 30.f) : len;
    o[5] = normalize ? tdiv<RCP>(wid, 15.f) : wid;
}
// r: an AgentMapObservation row (x, y, d0, d1, d2, heading, type, id, mapType)
template <bool RCP>
__device__ __forceinline__ void policy_road_row(const float *r, bool normalize, float (&o)[kPolicyRoad])
{
    const float two_pi = 6.283185307179586f;
    o[0] = normalize ? norm_min_max<RCP>(r[0]) : r[0];
    o[1] = normalize ? norm_min_max<RCP>(r[1]) : r[1];
    o[2] = normalize ? tdiv<RCP>(r[2], 100.f) : r[2];
    o[3] = normalize ? tdiv<RCP>(r[3], 100.f) : r[3];
    o[4] = normalize ? tdiv<RCP>(r[4], 100.f) : r[4];
    o[5] = normalize ? tdiv<RCP>(r[5], two_pi) : r[5];
    const int ty = (int)r[6];
#pragma unroll
    for (int k = 0; k < 7; k++) o[6 + k] = ty == k ? 1.f : 0.f;
}

// One CTA per agent slot. Inputs are read and the output written with fully coalesced accesses; the row
// arithmetic happens between two shared-memory images (needs S <= 128: the engine's limit).
template <bool RCP>
__global__ void __launch_bounds__(256)
k_policy_obs(Exports ex, float *__restrict__ out, int S, int normalize)
{
    constexpr int MAXP = 127;
    __shared__ __align__(16) float s_in[8 + MAXP * kPartnerObsSize + kMaxAgentMapObs * kMapObsSize];
    __shared__ __align__(16) float s_out[kPolicyEgo + MAXP * kPolicyPartner + kMaxAgentMapObs * kPolicyRoad];
    const size_t slot = blockIdx.x;
    const int P = S - 1, tid = threadIdx.x;
    const int n_po = P * kPartnerObsSize, n_ro = kMaxAgentMapObs * kMapObsSize;
    const int width = kPolicyEgo + P * kPolicyPartner + kMaxAgentMapObs * kPolicyRoad;
    const float *so = ex.self_obs + slot * kSelfObsSize;
    const float *po = ex.partner_obs + slot * (size_t)n_po;
    const float *ro = ex.agent_roadmap + slot * (size_t)n_ro;
    float *s_po = s_in + 8, *s_ro = s_in + 8 + n_po;
    if (tid < 8) s_in[tid] = so[tid];
    for (int i = tid; i < n_po; i += 256) s_po[i] = po[i];
    for (int i = tid; i < n_ro; i += 256) s_ro[i] = ro[i];
    __syncthreads();
    if (tid == 0) {                                 // speed, length, width, goal x, goal y, collided
        float o[kPolicyEgo];
        policy_ego_row<RCP>(s_in[0], s_in[1], s_in[2], s_in[4], s_in[5], s_in[6], normalize != 0, o);
#pragma unroll
        for (int k = 0; k < kPolicyEgo; k++) s_out[k] = o[k];
    }
    if (tid < P) {                                  // speed, x, y, heading, length, width
        float o[kPolicyPartner];
        policy_partner_row<RCP>(s_po + tid * kPartnerObsSize, normalize != 0, o);
#pragma unroll
        for (int k = 0; k < kPolicyPartner; k++) s_out[kPolicyEgo + tid * kPolicyPartner + k] = o[k];
    }
    if (tid < kMaxAgentMapObs) {                    // x, y, d0, d1, d2, heading, one_hot7(type)
        float o[kPolicyRoad];
        policy_road_row<RCP>(s_ro + tid * kMapObsSize, normalize != 0, o);
#pragma unroll
        for (int k = 0; k < kPolicyRoad; k++) s_out[kPolicyEgo + P * kPolicyPartner + tid * kPolicyRoad + k] = o[k];
    }
    __syncthreads();
    float *dst = out + slot * (size_t)width;
    if ((width & 3) == 0) {                         // S = 128: 3368 floats, 16-byte aligned rows
        float4 *d4 = reinterpret_cast<float4 *>(dst);
        const float4 *s4 = reinterpret_cast<const float4 *>(s_out);
        for (int i = tid; i < width / 4; i += 256) d4[i] = s4[i];
    } else {
        for (int i = tid; i < width; i += 256) dst[i] = s_out[i];
    }
}

} // namespace gd
