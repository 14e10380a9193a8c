// point_bwd.cu -- per visible point chain rule from screen-space gradients to xyz, q, s, opacity
// logit and SH coefficients, with the SH band mask and the constant gradient factors fused in.
//
// Replaces loop 2 of gaussian_point_rasterisation_backward (reference R.py:707-772, with
// GaussianPoint3D.py:132-159,237-331,351-373), _clear_grad_by_color_max_sh_band (R.py:1167-1182),
// the factor multiplies (R.py:1105-1125) and the hook-payload gathers (R.py:1127-1142).
//
// The reference multiplies dense 4x9, 9x9 and 9x{4,3} Jacobians; here the same derivative is
// evaluated in closed form: with U = J W, G the symmetric 2x2 gradient of Sigma', M = R S,
//   dL/dSigma = U^T G U,  dL/dM = 2 (dL/dSigma) M,
//   dL/ds_l = sum_k dL/dM[k,l] R[k,l] exp(s_l),  dL/dq = sum_kl dL/dM[k,l] exp(s_l) dR[k,l]/dq.
// Fast-fp32 profile (FMA contraction allowed); compared with the oracle at 1e-4 relative.
#include "gs_common.cuh"

namespace {

struct PointBwdParams {
    const float *xyz;
    const float *features;
    const int32_t *object_id;
    const float *intrinsics;
    const float *q_cam_pc;
    const float *t_cam_pc;
    const float *t_pc_cam;
    const int32_t *ids;
    long long m;
    const float *point_in_camera;
    const float *splat_grads;
    int keep;  // SH coefficients kept per channel: 1, 4, 9 or 16
    int accumulate;  // add into grad_xyz / grad_features instead of storing
    float q_factor, s_factor, alpha_factor, color_factor, high_order_color_factor;
    float *grad_xyz;
    float *grad_features;
    float *hook_grad_xyz;
    float *hook_grad_features;
    float *hook_grad_viewspace;
    float *hook_magnitude;
    int32_t *hook_num_affected;
};

// store, or add without reading back: one 16-byte vector reduction resolved in L2 (rows are
// unique within a launch, so there is no contention; the SM never waits for the old value)
__device__ __forceinline__ void put4(float4 *dst, const float4 &v, bool accumulate)
{
    if (accumulate) {
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(v.x), "f"(v.y),
                     "f"(v.z), "f"(v.w)
                     : "memory");
    } else {
        *dst = v;
    }
}

constexpr int kRowStride4 = 15;  // float4 per staged row: 14 used + 1 pad (240 B: conflict-free 16-byte lanes)

// Chain rule of one visible point.  `row` is the point's 56-float feature row staged in shared
// memory; it is overwritten IN PLACE with the row of feature gradients (already masked / scaled).
__device__ __forceinline__ void point_chain(const PointBwdParams &p, long long idx, long long pid,
                                            float4 *row)
{
    const float4 *sg4 = reinterpret_cast<const float4 *>(p.splat_grads + (size_t)kGradSlots * idx);
    const float4 ga = __ldg(sg4), gb = __ldg(sg4 + 1), gc = __ldg(sg4 + 2);
    const float gu = ga.x, gv = ga.y, c00 = ga.z, c01 = ga.w, c11 = gb.x;
    const float gcol[3] = {gb.y, gb.z, gb.w};
    const float glogit = gc.x, gmag = gc.y, gcount = gc.z;

    const int obj = __ldg(p.object_id + pid);
    float K[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) K[k] = __ldg(p.intrinsics + k);
    float W[9];
    {
        const float *q = p.q_cam_pc + 4 * obj;
        gs_rotation_from_quaternion(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3), W);
    }
    const float px = __ldg(p.xyz + 3 * pid), py = __ldg(p.xyz + 3 * pid + 1), pz = __ldg(p.xyz + 3 * pid + 2);
    const float cx = __ldg(p.point_in_camera + 3 * idx), cy = __ldg(p.point_in_camera + 3 * idx + 1),
                cz = __ldg(p.point_in_camera + 3 * idx + 2);
    const float iz = 1.0f / cz, iz2 = iz * iz;

    // d uv / d xyz = (d uv / d p_cam) W   (GaussianPoint3D.py:149-159, full K)
    float d[6] = {K[0] * iz, K[1] * iz, (-K[0] * cx - K[1] * cy) * iz2,
                  K[3] * iz, K[4] * iz, (-K[3] * cx - K[4] * cy) * iz2};
    float gxyz[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        float d0 = d[0] * W[j] + d[1] * W[3 + j] + d[2] * W[6 + j];
        float d1 = d[3] * W[j] + d[4] * W[3 + j] + d[5] * W[6 + j];
        gxyz[j] = gu * d0 + gv * d1;
    }

    // U = J W
    const float J0 = K[0] * iz, J2 = -(K[0] * cx) * iz2, J4 = K[4] * iz, J5 = -(K[4] * cy) * iz2;
    float U[6];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        U[j] = J0 * W[j] + J2 * W[6 + j];
        U[3 + j] = J4 * W[3 + j] + J5 * W[6 + j];
    }
    // A = U^T G U
    float GU[6];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        GU[j] = c00 * U[j] + c01 * U[3 + j];
        GU[3 + j] = c01 * U[j] + c11 * U[3 + j];
    }
    float A[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) A[3 * i + j] = U[i] * GU[j] + U[3 + i] * GU[3 + j];

    const float4 q4 = row[0], s4 = row[1];
    const float qx = q4.x, qy = q4.y, qz = q4.z, qw = q4.w;
    float R[9];
    gs_rotation_from_quaternion(qx, qy, qz, qw, R);
    const float es[3] = {__expf(s4.x), __expf(s4.y), __expf(s4.z)};
    // D[k][l] = dL/dM[k][l] * exp(s_l) = 2 (A M)[k][l] exp(s_l),  M[j][l] = R[j][l] exp(s_l)
    float D[9];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int l = 0; l < 3; ++l) {
            float am = A[3 * k] * R[l] + A[3 * k + 1] * R[3 + l] + A[3 * k + 2] * R[6 + l];
            D[3 * k + l] = 2.0f * am * es[l] * es[l];
        }
    float gs3[3];
#pragma unroll
    for (int l = 0; l < 3; ++l) gs3[l] = D[l] * R[l] + D[3 + l] * R[3 + l] + D[6 + l] * R[6 + l];
    // dR/dq contracted with D (table of GaussianPoint3D.py:319-329)
    float gq[4];
    gq[0] = 2.0f * (D[1] * qy + D[2] * qz + D[3] * qy - D[5] * qw + D[6] * qz + D[7] * qw) -
            4.0f * (D[4] * qx + D[8] * qx);
    gq[1] = 2.0f * (D[1] * qx + D[2] * qw + D[3] * qx + D[5] * qz - D[6] * qw + D[7] * qz) -
            4.0f * (D[0] * qy + D[8] * qy);
    gq[2] = 2.0f * (-D[1] * qw + D[2] * qx + D[3] * qw + D[5] * qy + D[6] * qx + D[7] * qy) -
            4.0f * (D[0] * qz + D[4] * qz);
    gq[3] = 2.0f * (-D[1] * qz + D[2] * qy + D[3] * qz - D[5] * qx - D[6] * qy + D[7] * qx);

    // SH colour Jacobian: sigmoid'(raw) * basis (GaussianPoint3D.py:351-373); the ray starts at
    // the input translation t_pointcloud_camera (R.py:731-732,749)
    const float *org = p.t_pc_cam + 3 * obj;
    float basis[16];
    gs_sh_basis(px - __ldg(org), py - __ldg(org + 1), pz - __ldg(org + 2), basis);

    float4 *grow = row;
    {
        const float4 o0 = make_float4(gq[0] * p.q_factor, gq[1] * p.q_factor, gq[2] * p.q_factor,
                                      gq[3] * p.q_factor);
        const float4 o1 = make_float4(gs3[0] * p.s_factor, gs3[1] * p.s_factor, gs3[2] * p.s_factor,
                                      glogit * p.alpha_factor);
        grow[0] = o0;
        grow[1] = o1;
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        float raw = 0.f;
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
            float4 c = row[2 + 4 * ch + k4];
            raw += c.x * basis[4 * k4] + c.y * basis[4 * k4 + 1] + c.z * basis[4 * k4 + 2] +
                   c.w * basis[4 * k4 + 3];
        }
        const float sg = 1.0f / (1.0f + __expf(-raw));
        const float scale = gcol[ch] * sg * (1.0f - sg);
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
            float o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int k = 4 * k4 + e;
                const float f = (k == 0) ? p.color_factor : p.high_order_color_factor;
                o[e] = (k < p.keep) ? scale * basis[k] * f : 0.0f;
            }
            const float4 o4 = make_float4(o[0], o[1], o[2], o[3]);
            grow[2 + 4 * ch + k4] = o4;  // the channel's coefficients were all read above
        }
    }
    if (p.accumulate) {
        atomicAdd(p.grad_xyz + 3 * pid, gxyz[0]);
        atomicAdd(p.grad_xyz + 3 * pid + 1, gxyz[1]);
        atomicAdd(p.grad_xyz + 3 * pid + 2, gxyz[2]);
    } else {
        p.grad_xyz[3 * pid] = gxyz[0];
        p.grad_xyz[3 * pid + 1] = gxyz[1];
        p.grad_xyz[3 * pid + 2] = gxyz[2];
    }

    if (p.hook_grad_xyz) {
        p.hook_grad_xyz[3 * idx] = gxyz[0];
        p.hook_grad_xyz[3 * idx + 1] = gxyz[1];
        p.hook_grad_xyz[3 * idx + 2] = gxyz[2];
    }
    if (p.hook_grad_viewspace) {
        p.hook_grad_viewspace[2 * idx] = gu;
        p.hook_grad_viewspace[2 * idx + 1] = gv;
    }
    if (p.hook_magnitude) p.hook_magnitude[idx] = gmag;
    if (p.hook_num_affected) p.hook_num_affected[idx] = (int32_t)gcount;
}

// One thread per visible point for the arithmetic; the 224-byte feature rows in and the 224-byte
// gradient rows out move through shared memory so that every global access is a coalesced run:
// a half warp handles one row per instruction (14 lanes x 16 B), two rows per instruction.
__global__ void __launch_bounds__(128) point_bwd_kernel(PointBwdParams p)
{
    __shared__ float4 s_rows[4][32 * kRowStride4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = idx < p.m;
    const long long pid = valid ? (long long)p.ids[idx] : -1;
    float4 *rows = s_rows[warp];
    const int half = lane >> 4, col = lane & 15;

    // asynchronous 16-byte copies (LDGSTS): the warp's 448 loads are all in flight at once and
    // never pass through registers
#pragma unroll 8
    for (int r = 0; r < 32; r += 2) {
        const long long rpid = __shfl_sync(0xffffffffu, pid, r + half);
        if (col < 14 && rpid >= 0) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(rows + (r + half) * kRowStride4 + col);
            const float4 *src = reinterpret_cast<const float4 *>(p.features + (size_t)GS_FEATURES * rpid) + col;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    if (valid) point_chain(p, idx, pid, rows + lane * kRowStride4);
    __syncwarp();
    const long long idx0 = idx - lane;
#pragma unroll 8
    for (int r = 0; r < 32; r += 2) {
        const long long rpid = __shfl_sync(0xffffffffu, pid, r + half);
        if (col < 14 && rpid >= 0) {
            const float4 v = rows[(r + half) * kRowStride4 + col];
            put4(reinterpret_cast<float4 *>(p.grad_features + (size_t)GS_FEATURES * rpid) + col, v,
                 p.accumulate);
            if (p.hook_grad_features)
                reinterpret_cast<float4 *>(p.hook_grad_features + (size_t)GS_FEATURES * (idx0 + r + half))[col] = v;
        }
    }
}

}  // namespace

extern "C" int gs_point_bwd(const float *xyz, const float *features, const int32_t *object_id,
                            const float *intrinsics, const float *q_cam_pc, const float *t_cam_pc,
                            const float *t_pc_cam, int n_obj, const int32_t *ids, int64_t m,
                            const float *point_in_camera, const float *splat_grads,
                            int color_max_sh_band, float q_factor, float s_factor,
                            float alpha_factor, float color_factor, float high_order_color_factor,
                            float *grad_xyz, float *grad_features, float *hook_grad_xyz,
                            float *hook_grad_features, float *hook_grad_viewspace,
                            float *hook_magnitude, int32_t *hook_num_affected, int accumulate,
                            void *stream)
{
    if (m < 0 || n_obj <= 0) return GS_E_INVALID;
    if (m == 0) return GS_OK;
    if (!xyz || !features || !object_id || !intrinsics || !q_cam_pc || !t_cam_pc || !t_pc_cam ||
        !ids || !point_in_camera || !splat_grads || !grad_xyz || !grad_features)
        return GS_E_INVALID;
    // rows move as 16-byte vectors (cp.async in, st / red.global.add.v4 out): a 224-byte row pitch
    // keeps every row aligned exactly when the base pointers are
    if (((uintptr_t)features | (uintptr_t)grad_features | (uintptr_t)splat_grads) & 15u) return GS_E_INVALID;
    if (hook_grad_features && ((uintptr_t)hook_grad_features & 15u)) return GS_E_INVALID;
    PointBwdParams p;
    p.xyz = xyz; p.features = features; p.object_id = object_id; p.intrinsics = intrinsics;
    p.q_cam_pc = q_cam_pc; p.t_cam_pc = t_cam_pc; p.t_pc_cam = t_pc_cam; p.ids = ids; p.m = m;
    p.point_in_camera = point_in_camera; p.splat_grads = splat_grads;
    p.keep = color_max_sh_band <= 0 ? 1 : color_max_sh_band == 1 ? 4 : color_max_sh_band == 2 ? 9 : 16;
    p.q_factor = q_factor; p.s_factor = s_factor; p.alpha_factor = alpha_factor;
    p.color_factor = color_factor; p.high_order_color_factor = high_order_color_factor;
    p.grad_xyz = grad_xyz; p.grad_features = grad_features;
    p.hook_grad_xyz = hook_grad_xyz; p.hook_grad_features = hook_grad_features;
    p.hook_grad_viewspace = hook_grad_viewspace; p.hook_magnitude = hook_magnitude;
    p.hook_num_affected = hook_num_affected;
    p.accumulate = accumulate;
    int threads = 128;
    long long blocks = (m + threads - 1) / threads;
    point_bwd_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
