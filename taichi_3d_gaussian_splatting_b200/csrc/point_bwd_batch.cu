// point_bwd_batch.cu -- the per-point chain rule of a whole BATCH of views in one pass over the points.
//
// The view-sharded step renders V views against one cloud and wants the SUM of
// their point gradients.  gs_point_bwd in accumulate mode does that view by view: every view reads the
// 224-byte feature row of each of its visible points again and adds its 236-byte gradient row into the
// N x 59 buffer with RED (a read-modify-write in L2 / HBM), and the buffer needs a zero fill per step:
// per step 4 x 59 N + V x (280 + 2 x 236) M bytes.  Here ONE thread owns a point for the whole batch:
//
//   * its feature row is fetched once (cp.async into shared memory, as in point_bwd.cu);
//   * for every view in which the point is visible (slot_of_point[v][p] >= 0, written by
//     gs_project_cull_sh) it reads the view's 48-byte reduced splat gradients and its camera-space
//     position, and accumulates what depends on the view IN REGISTERS: d xyz (3), the symmetric
//     dL/dSigma = U^T G U (6 -- the chain from dL/dSigma to the quaternion and the scales is linear in
//     it and does not depend on the view, so it runs once per point, not once per view), the opacity
//     logit (1) and the 48 SH coefficients (the basis depends on the view direction);
//   * the finished 236-byte row is STORED, once, for every point of the range whether it was visible
//     or not: no zero fill of the buffer, no RED.
//
// Traffic per step: (224 + 236 + 4 V) N + 60 V M bytes (C2, V = 8: 0.93 GB against 4.1 GB).  A call
// covers a range of point ids, so that a step can be cut into chunks whose results are exchanged
// (all-reduce over the ranks) while the next chunk is computed.
//
// Same arithmetic as point_bwd.cu (fast-fp32 profile; reference R.py:707-772, 1102-1125, 1167-1182,
// GaussianPoint3D.py:132-159, 237-331, 351-373); the sums over views are formed in view order.
#include "gs_common.cuh"

namespace {

constexpr int kBatchBlock = 128;
constexpr int kRowStride4 = 15;  // float4 per staged row: 14 used + 1 pad (conflict-free 16-byte lanes)

struct BatchParams {
    const float *xyz;
    const float *features;
    const int32_t *object_id;
    long long first, count;       // point ids [first, first + count)
    const gs_view_t *views;
    int num_views;
    int keep;
    float q_factor, s_factor, alpha_factor, color_factor, high_order_color_factor;
    float *grad_xyz;
    float *grad_features;
};

__global__ void __launch_bounds__(kBatchBlock) point_bwd_batch_kernel(BatchParams p)
{
    __shared__ float4 s_rows[kBatchBlock / 32][32 * kRowStride4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long local = (long long)blockIdx.x * kBatchBlock + threadIdx.x;
    const bool valid = local < p.count;
    const long long pid = p.first + local;
    float4 *rows = s_rows[warp];
    const int half = lane >> 4, col = lane & 15;
    const long long pid0 = pid - lane;  // the warp's 32 points are consecutive ids

    // the warp's 32 feature rows: 448 asynchronous 16-byte copies, all in flight at once
#pragma unroll 8
    for (int r = 0; r < 32; r += 2) {
        const long long rp = pid0 + r + half;
        if (col < 14 && rp < p.first + p.count) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(rows + (r + half) * kRowStride4 + col);
            const float4 *src = reinterpret_cast<const float4 *>(p.features + (size_t)GS_FEATURES * rp) + col;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");

    float px = 0.f, py = 0.f, pz = 0.f;
    int obj = 0;
    if (valid) {
        px = __ldg(p.xyz + 3 * pid); py = __ldg(p.xyz + 3 * pid + 1); pz = __ldg(p.xyz + 3 * pid + 2);
        obj = __ldg(p.object_id + pid);
    }
    float gxyz[3] = {0.f, 0.f, 0.f};
    float A[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // dL/dSigma: 00 01 02 11 12 22
    float glogit = 0.f;
    float sh[48];
#pragma unroll
    for (int k = 0; k < 48; ++k) sh[k] = 0.f;

    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    float4 *row = rows + lane * kRowStride4;

    for (int v = 0; v < p.num_views; ++v) {
        const gs_view_t &view = p.views[v];
        const int slot = valid ? __ldg(view.slot_of_point + pid) : -1;
        if (slot < 0) continue;
        const float4 *sg4 = reinterpret_cast<const float4 *>(view.splat_grads + (size_t)kGradSlots * slot);
        const float4 ga = __ldg(sg4), gb = __ldg(sg4 + 1), gc = __ldg(sg4 + 2);
        const float gu = ga.x, gv = ga.y, c00 = ga.z, c01 = ga.w, c11 = gb.x;
        const float gcol[3] = {gb.y, gb.z, gb.w};
        glogit += gc.x;
        float K[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) K[k] = __ldg(view.intrinsics + k);
        float W[9];
        {
            const float *q = view.q_cam_pc + 4 * obj;
            gs_rotation_from_quaternion(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3), W);
        }
        const float *pic = view.point_in_camera + 3 * (size_t)slot;
        const float cx = __ldg(pic), cy = __ldg(pic + 1), cz = __ldg(pic + 2);
        const float iz = 1.0f / cz, iz2 = iz * iz;
        // d uv / d xyz = (d uv / d p_cam) W   (GaussianPoint3D.py:149-159, full K)
        const float d[6] = {K[0] * iz, K[1] * iz, (-K[0] * cx - K[1] * cy) * iz2,
                            K[3] * iz, K[4] * iz, (-K[3] * cx - K[4] * cy) * iz2};
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const float d0 = d[0] * W[j] + d[1] * W[3 + j] + d[2] * W[6 + j];
            const float d1 = d[3] * W[j] + d[4] * W[3 + j] + d[5] * W[6 + j];
            gxyz[j] += gu * d0 + gv * d1;
        }
        // U = J W;  dL/dSigma += U^T G U
        const float J0 = K[0] * iz, J2 = -(K[0] * cx) * iz2, J4 = K[4] * iz, J5 = -(K[4] * cy) * iz2;
        float U[6], GU[6];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            U[j] = J0 * W[j] + J2 * W[6 + j];
            U[3 + j] = J4 * W[3 + j] + J5 * W[6 + j];
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            GU[j] = c00 * U[j] + c01 * U[3 + j];
            GU[3 + j] = c01 * U[j] + c11 * U[3 + j];
        }
        A[0] += U[0] * GU[0] + U[3] * GU[3];
        A[1] += U[0] * GU[1] + U[3] * GU[4];
        A[2] += U[0] * GU[2] + U[3] * GU[5];
        A[3] += U[1] * GU[1] + U[4] * GU[4];
        A[4] += U[1] * GU[2] + U[4] * GU[5];
        A[5] += U[2] * GU[2] + U[5] * GU[5];
        // SH colour Jacobian: sigmoid'(raw) * basis (GaussianPoint3D.py:351-373); the ray starts at the
        // input translation t_pointcloud_camera (R.py:731-732,749)
        const float *org = view.t_pc_cam + 3 * obj;
        float basis[16];
        gs_sh_basis(px - __ldg(org), py - __ldg(org + 1), pz - __ldg(org + 2), basis);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            float raw = 0.f;
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                const float4 c = row[2 + 4 * ch + k4];
                raw += c.x * basis[4 * k4] + c.y * basis[4 * k4 + 1] + c.z * basis[4 * k4 + 2] +
                       c.w * basis[4 * k4 + 3];
            }
            const float sg = 1.0f / (1.0f + __expf(-raw));
            const float scale = gcol[ch] * sg * (1.0f - sg);
#pragma unroll
            for (int k = 0; k < 16; ++k) sh[16 * ch + k] = fmaf(scale, basis[k], sh[16 * ch + k]);
        }
    }

    if (valid) {
        // once per point: dL/dM = 2 (dL/dSigma) M, then to the scales and the (normalised) quaternion
        const float4 q4 = row[0], s4 = row[1];
        const float qx = q4.x, qy = q4.y, qz = q4.z, qw = q4.w;
        float R[9];
        gs_rotation_from_quaternion(qx, qy, qz, qw, R);
        const float es[3] = {__expf(s4.x), __expf(s4.y), __expf(s4.z)};
        const float Af[9] = {A[0], A[1], A[2], A[1], A[3], A[4], A[2], A[4], A[5]};
        float D[9];
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int l = 0; l < 3; ++l) {
                const float am = Af[3 * k] * R[l] + Af[3 * k + 1] * R[3 + l] + Af[3 * k + 2] * R[6 + l];
                D[3 * k + l] = 2.0f * am * es[l] * es[l];
            }
        float gs3[3];
#pragma unroll
        for (int l = 0; l < 3; ++l) gs3[l] = D[l] * R[l] + D[3 + l] * R[3 + l] + D[6 + l] * R[6 + l];
        float gq[4];
        gq[0] = 2.0f * (D[1] * qy + D[2] * qz + D[3] * qy - D[5] * qw + D[6] * qz + D[7] * qw) -
                4.0f * (D[4] * qx + D[8] * qx);
        gq[1] = 2.0f * (D[1] * qx + D[2] * qw + D[3] * qx + D[5] * qz - D[6] * qw + D[7] * qz) -
                4.0f * (D[0] * qy + D[8] * qy);
        gq[2] = 2.0f * (-D[1] * qw + D[2] * qx + D[3] * qw + D[5] * qy + D[6] * qx + D[7] * qy) -
                4.0f * (D[0] * qz + D[4] * qz);
        gq[3] = 2.0f * (-D[1] * qz + D[2] * qy + D[3] * qz - D[5] * qx - D[6] * qy + D[7] * qx);
        // the gradient row replaces the feature row in shared memory (band mask and factors applied)
        row[0] = make_float4(gq[0] * p.q_factor, gq[1] * p.q_factor, gq[2] * p.q_factor, gq[3] * p.q_factor);
        row[1] = make_float4(gs3[0] * p.s_factor, gs3[1] * p.s_factor, gs3[2] * p.s_factor,
                             glogit * p.alpha_factor);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                float o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int k = 4 * k4 + e;
                    const float f = (k == 0) ? p.color_factor : p.high_order_color_factor;
                    o[e] = (k < p.keep) ? sh[16 * ch + k] * f : 0.0f;
                }
                row[2 + 4 * ch + k4] = make_float4(o[0], o[1], o[2], o[3]);
            }
        p.grad_xyz[3 * pid] = gxyz[0];
        p.grad_xyz[3 * pid + 1] = gxyz[1];
        p.grad_xyz[3 * pid + 2] = gxyz[2];
    }
    __syncwarp();
#pragma unroll 8
    for (int r = 0; r < 32; r += 2) {
        const long long rp = pid0 + r + half;
        if (col < 14 && rp < p.first + p.count)
            reinterpret_cast<float4 *>(p.grad_features + (size_t)GS_FEATURES * rp)[col] =
                rows[(r + half) * kRowStride4 + col];
    }
}

}  // namespace

extern "C" int gs_point_bwd_batch(const float *xyz, const float *features, const int32_t *object_id,
                                  int64_t first_point, int64_t num_points, const gs_view_t *views_device,
                                  int num_views, int color_max_sh_band, float q_factor, float s_factor,
                                  float alpha_factor, float color_factor, float high_order_color_factor,
                                  float *grad_xyz, float *grad_features, void *stream)
{
    if (first_point < 0 || num_points < 0 || num_views < 0) return GS_E_INVALID;
    if (num_points == 0) return GS_OK;
    if (!xyz || !features || !object_id || !grad_xyz || !grad_features || (num_views > 0 && !views_device))
        return GS_E_INVALID;
    if (((uintptr_t)features | (uintptr_t)grad_features | (uintptr_t)views_device) & 15u) return GS_E_INVALID;
    BatchParams p;
    p.xyz = xyz; p.features = features; p.object_id = object_id;
    p.first = first_point; p.count = num_points; p.views = views_device; p.num_views = num_views;
    p.keep = color_max_sh_band <= 0 ? 1 : color_max_sh_band == 1 ? 4 : color_max_sh_band == 2 ? 9 : 16;
    p.q_factor = q_factor; p.s_factor = s_factor; p.alpha_factor = alpha_factor;
    p.color_factor = color_factor; p.high_order_color_factor = high_order_color_factor;
    p.grad_xyz = grad_xyz; p.grad_features = grad_features;
    const long long blocks = (num_points + kBatchBlock - 1) / kBatchBlock;
    point_bwd_batch_kernel<<<(unsigned)blocks, kBatchBlock, 0, (cudaStream_t)stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
