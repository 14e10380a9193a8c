// The two streaming steps that follow the rasteriser's backward in a training iteration
// (SURVEY.md section 8(f) rows 2 and 3).  Both are HBM-bound element-wise passes.
//
//   gs_controller_accumulate : the six scatter-adds of GaussianPointAdaptiveController.update
//                              (reference GaussianPointAdaptiveController.py:130-146) in one pass
//                              over the backward hook's payload;
//   gs_adam_step             : torch.optim.Adam's single-tensor update (the reference trainer's two
//                              optimisers, GaussianPointTrainer.py:131-134,183-184) in one pass:
//                              28 B per element (read p, g, m, v; write p, m, v).
#include <cuda_runtime.h>
#include <stdint.h>

#include "gs_b200.h"

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
controller_accumulate_kernel(const int32_t *__restrict__ ids, const long long m,
                             const int32_t *__restrict__ num_affected, const float *__restrict__ magnitude,
                             const float *__restrict__ grad_xyz, int32_t *__restrict__ acc_num_in_camera,
                             int32_t *__restrict__ acc_num_pixels, float *__restrict__ acc_view_grad,
                             float *__restrict__ acc_view_grad_avg, float *__restrict__ acc_pos_grad,
                             float *__restrict__ acc_pos_grad_norm)
{
    const long long i = (long long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= m) return;
    // ids are distinct within ONE launch (R.py:864), but launches of different views may run
    // concurrently on different streams (ViewShardedTrainerStep lanes) and then touch the same
    // points: every update is a reduction resolved in L2 (RED, no return value), so concurrent
    // launches cannot lose updates
    const long long id = ids[i];
    const int32_t px = num_affected[i];
    const float mag = magnitude[i];
    atomicAdd(acc_num_in_camera + id, 1);
    atomicAdd(acc_num_pixels + id, px);
    atomicAdd(acc_view_grad + id, mag);
    float avg = mag / (float)px;  // 0/0 -> NaN -> 0, as `avg[isnan(avg)] = 0` (AC.py:137-138)
    if (avg != avg) avg = 0.0f;
    atomicAdd(acc_view_grad_avg + id, avg);
    const float gx = grad_xyz[3 * i], gy = grad_xyz[3 * i + 1], gz = grad_xyz[3 * i + 2];
    atomicAdd(acc_pos_grad + 3 * id, gx);
    atomicAdd(acc_pos_grad + 3 * id + 1, gy);
    atomicAdd(acc_pos_grad + 3 * id + 2, gz);
    atomicAdd(acc_pos_grad_norm + id, sqrtf(gx * gx + gy * gy + gz * gz));
}

struct AdamScalars {
    float step_size, beta1, beta2, one_minus_beta1, one_minus_beta2, inv_sqrt_bc2, eps;
};

__device__ __forceinline__ void adam_one(float &p, const float g, float &m, float &v, const AdamScalars &a)
{
    m = m + a.one_minus_beta1 * (g - m);                // exp_avg.lerp_(grad, 1 - beta1)
    v = a.beta2 * v + a.one_minus_beta2 * (g * g);      // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
    const float denom = sqrtf(v) * a.inv_sqrt_bc2 + a.eps;
    p = p - a.step_size * (m / denom);                  // param.addcdiv_(exp_avg, denom, -step_size)
}

// 4 elements per thread, 16-byte accesses; `rows_per_mask` > 0 with a mask skips the elements of
// rows whose mask byte is non-zero (row = element / rows_per_mask).
template <bool kMasked>
__global__ void __launch_bounds__(kThreads)
adam_vec_kernel(float4 *__restrict__ p, const float4 *__restrict__ g, float4 *__restrict__ m,
                float4 *__restrict__ v, const long long count4, const AdamScalars a,
                const int8_t *__restrict__ row_mask, const int row_len)
{
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < count4;
         i += (long long)gridDim.x * kThreads) {
        float4 pp = p[i], mm = m[i], vv = v[i];
        const float4 gg = g[i];
        bool k0 = true, k1 = true, k2 = true, k3 = true;
        if (kMasked) {
            const long long e = 4 * i;
            k0 = row_mask[e / row_len] == 0;
            k1 = row_mask[(e + 1) / row_len] == 0;
            k2 = row_mask[(e + 2) / row_len] == 0;
            k3 = row_mask[(e + 3) / row_len] == 0;
            if (!(k0 || k1 || k2 || k3)) continue;
        }
        if (k0) adam_one(pp.x, gg.x, mm.x, vv.x, a);
        if (k1) adam_one(pp.y, gg.y, mm.y, vv.y, a);
        if (k2) adam_one(pp.z, gg.z, mm.z, vv.z, a);
        if (k3) adam_one(pp.w, gg.w, mm.w, vv.w, a);
        p[i] = pp;
        m[i] = mm;
        v[i] = vv;
    }
}

template <bool kMasked>
__global__ void __launch_bounds__(kThreads)
adam_scalar_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                   float *__restrict__ v, const long long begin, const long long count,
                   const AdamScalars a, const int8_t *__restrict__ row_mask, const int row_len)
{
    for (long long i = begin + (long long)blockIdx.x * kThreads + threadIdx.x; i < count;
         i += (long long)gridDim.x * kThreads) {
        if (kMasked && row_mask[i / row_len] != 0) continue;
        float pp = p[i], mm = m[i], vv = v[i];
        adam_one(pp, g[i], mm, vv, a);
        p[i] = pp;
        m[i] = mm;
        v[i] = vv;
    }
}

inline bool aligned16(const void *q) { return ((uintptr_t)q & 15u) == 0; }

}  // namespace

extern "C" int gs_controller_accumulate(const int32_t *ids, int64_t m, const int32_t *num_affected,
                                        const float *magnitude, const float *grad_xyz,
                                        int32_t *acc_num_in_camera, int32_t *acc_num_pixels,
                                        float *acc_view_grad, float *acc_view_grad_avg,
                                        float *acc_pos_grad, float *acc_pos_grad_norm, void *stream)
{
    if (m < 0) return GS_E_INVALID;
    if (m == 0) return GS_OK;
    if (!ids || !num_affected || !magnitude || !grad_xyz || !acc_num_in_camera || !acc_num_pixels ||
        !acc_view_grad || !acc_view_grad_avg || !acc_pos_grad || !acc_pos_grad_norm)
        return GS_E_INVALID;
    const long long blocks = (m + kThreads - 1) / kThreads;
    if (blocks > 0x7fffffffLL) return GS_E_CAPACITY;
    controller_accumulate_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(
        ids, m, num_affected, magnitude, grad_xyz, acc_num_in_camera, acc_num_pixels, acc_view_grad,
        acc_view_grad_avg, acc_pos_grad, acc_pos_grad_norm);
    return (int)cudaGetLastError();
}

extern "C" int gs_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                            int64_t count, float lr, float beta1, float beta2, float eps, int64_t step,
                            const int8_t *row_mask, int row_len, void *stream)
{
    if (count < 0 || step < 1) return GS_E_INVALID;
    if (count == 0) return GS_OK;
    if (!param || !grad || !exp_avg || !exp_avg_sq) return GS_E_INVALID;
    if (row_mask && (row_len <= 0 || count % row_len != 0)) return GS_E_INVALID;
    // torch/optim/adam.py _single_tensor_adam: the scalar factors are formed on the host in double
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    AdamScalars a;
    a.step_size = (float)((double)lr / bc1);
    a.beta1 = beta1;
    a.beta2 = beta2;
    a.one_minus_beta1 = (float)(1.0 - (double)beta1);
    a.one_minus_beta2 = (float)(1.0 - (double)beta2);
    a.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    a.eps = eps;
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = aligned16(param) && aligned16(grad) && aligned16(exp_avg) && aligned16(exp_avg_sq);
    const long long count4 = vec ? count / 4 : 0;
    const int max_blocks = 148 * 16;
    if (count4 > 0) {
        long long blocks = (count4 + kThreads - 1) / kThreads;
        if (blocks > max_blocks) blocks = max_blocks;
        if (row_mask)
            adam_vec_kernel<true><<<(unsigned)blocks, kThreads, 0, st>>>(
                (float4 *)param, (const float4 *)grad, (float4 *)exp_avg, (float4 *)exp_avg_sq, count4, a,
                row_mask, row_len);
        else
            adam_vec_kernel<false><<<(unsigned)blocks, kThreads, 0, st>>>(
                (float4 *)param, (const float4 *)grad, (float4 *)exp_avg, (float4 *)exp_avg_sq, count4, a,
                nullptr, 1);
    }
    const long long begin = 4 * count4;
    if (begin < count) {
        long long blocks = (count - begin + kThreads - 1) / kThreads;
        if (blocks > max_blocks) blocks = max_blocks;
        if (row_mask)
            adam_scalar_kernel<true><<<(unsigned)blocks, kThreads, 0, st>>>(param, grad, exp_avg, exp_avg_sq,
                                                                            begin, count, a, row_mask, row_len);
        else
            adam_scalar_kernel<false><<<(unsigned)blocks, kThreads, 0, st>>>(param, grad, exp_avg, exp_avg_sq,
                                                                             begin, count, a, nullptr, 1);
    }
    return (int)cudaGetLastError();
}
