// blend_bwd.cu -- reverse-order blend backward (reference R.py:531-705, utils.py:331-348).
//
// Fast-fp32 profile (FMA contraction, ex2.approx): gradients are compared with the oracle at
// 1e-4 relative, the atomic summation order differs from the reference anyway.
//
// One 256-thread CTA per 16x16 tile, one pixel per thread.  The tile's list is walked from
// max-over-pixels(last effective index) down to the tile start (entries past that maximum are
// skipped by every pixel in the reference as well, R.py:609-610) in batches of 256 staged in
// shared memory.  For each splat the 10 per-pixel gradient terms are reduced across the warp
// with a recursive-halving butterfly (16 shuffles instead of 50), added to a per-batch
// shared-memory accumulator (one shared atomic per warp per term), and flushed once per
// (tile, splat) with three 16-byte vector reductions into the [M,12] splat-gradient buffer.
// The reference issues 11 global atomics per (pixel, splat) pair (R.py:674-696).
#include "gs_common.cuh"

namespace {

constexpr int kBatch = 256;

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(addr), "f"(a), "f"(b),
                 "f"(c), "f"(d)
                 : "memory");
}

// Sum 16 per-lane values over the 32 lanes of a warp.  On return lane L holds in v[0] the total
// of value index L >> 1.
__device__ __forceinline__ void warp_reduce16(float (&v)[16], int lane)
{
    {
        const bool up = lane & 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float send = up ? v[i] : v[i + 8];
            float keep = up ? v[i + 8] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float send = up ? v[i] : v[i + 4];
            float keep = up ? v[i + 4] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float send = up ? v[i] : v[i + 2];
            float keep = up ? v[i + 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    {
        const bool up = lane & 2;
        float send = up ? v[0] : v[1];
        float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

__global__ void __launch_bounds__(kTilePixels)
blend_bwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, const float *__restrict__ image_grad,
                 const float *__restrict__ acc_alpha, const int32_t *__restrict__ last_effective,
                 float *__restrict__ splat_grads, float *__restrict__ map_on_image)
{
    __shared__ float4 s_rec[kBatch * 3];
    __shared__ __align__(16) float s_acc[kBatch * kGradSlots];
    __shared__ int32_t s_idx[kBatch];
    __shared__ int s_hi;

    const int tid = threadIdx.x, lane = tid & 31;
    const int tiles_per_row = width / kTile;
    const int tile = blockIdx.x;
    const int tile_u = tile % tiles_per_row, tile_v = tile / tiles_per_row;
    const int pu = tile_u * kTile + (tid % kTile), pv = tile_v * kTile + (tid / kTile);
    const float fx = (float)pu + 0.5f, fy = (float)pv + 0.5f;
    const int start = tile_start[tile], end = tile_end[tile];
    const size_t px = (size_t)pv * width + pu;

    const int my_last = (end > start) ? last_effective[px] : start;
    float T = 1.0f - acc_alpha[px];
    const float g0 = image_grad[3 * px], g1 = image_grad[3 * px + 1], g2 = image_grad[3 * px + 2];
    float w0 = 0.f, w1 = 0.f, w2 = 0.f, tot0 = 0.f, tot1 = 0.f;

    if (tid == 0) s_hi = start;
    __syncthreads();
    atomicMax(&s_hi, min(my_last, end));
    __syncthreads();
    const int hi = s_hi;

    for (int top = hi; top > start; top -= kBatch) {
        const int lo = max(top - kBatch, start);
        const int n = top - lo;
        __syncthreads();  // previous batch fully flushed
        if (tid < n) {
            int g = payload[top - 1 - tid];
            s_idx[tid] = g;
            const float4 *src = records + (size_t)3 * g;
            s_rec[3 * tid] = __ldg(src);
            s_rec[3 * tid + 1] = __ldg(src + 1);
            s_rec[3 * tid + 2] = __ldg(src + 2);
        }
#pragma unroll
        for (int q = 0; q < kGradSlots / 4; ++q)
            reinterpret_cast<float4 *>(s_acc)[(kGradSlots / 4) * tid + q] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();

        for (int k = 0; k < n; ++k) {
            const int i = top - 1 - k;
            const float4 r0 = s_rec[3 * k];      // u v a b
            const float4 r1 = s_rec[3 * k + 1];  // c rescale alpha red
            const float dx = fx - r0.x, dy = fy - r0.y;
            const float v0 = r0.z * dx + r0.w * dy;
            const float v1 = r0.w * dx + r1.x * dy;
            const float quad = dx * v0 + dy * v1;
            const float p = __expf(-0.5f * quad) * r1.y;
            const float prod = p * r1.z;
            const bool contributes = (i < my_last) && (prod >= 1.0f / 255.0f);
            const unsigned int mask = __ballot_sync(0xffffffffu, contributes);
            if (mask == 0u) continue;

            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = 0.f;
            if (contributes) {
                const float4 r2 = s_rec[3 * k + 2];  // green blue z pad
                const float a = fminf(prod, 0.99f);
                const float inv = __frcp_rn(1.0f - a);
                T = T * inv;
                const float dcol = a * T;
                const float agr = (r1.w * T - w0 * inv) * g0 + (r2.x * T - w1 * inv) * g1 +
                                  (r2.y * T - w2 * inv) * g2;
                w0 += r1.w * dcol;
                w1 += r2.x * dcol;
                w2 += r2.y * dcol;
                const float sa = agr * prod;  // dL/dp * p
                const float gm0 = sa * v0, gm1 = sa * v1;
                tot0 += fabsf(gm0);
                tot1 += fabsf(gm1);
                const float hs = 0.5f * sa;
                v[0] = gm0;
                v[1] = gm1;
                v[2] = hs * v0 * v0;
                v[3] = hs * v0 * v1;
                v[4] = hs * v1 * v1;
                v[5] = dcol * g0;
                v[6] = dcol * g1;
                v[7] = dcol * g2;
                v[8] = sa * (1.0f - r1.z);
                v[9] = sqrtf(gm0 * gm0 + gm1 * gm1);
            }
            warp_reduce16(v, lane);
            const int q = lane >> 1;
            if (!(lane & 1)) {
                if (q < 10)
                    atomicAdd(&s_acc[kGradSlots * k + q], v[0]);
                else if (q == 10)
                    atomicAdd(&s_acc[kGradSlots * k + 10], (float)__popc(mask));
            }
        }
        __syncthreads();
        if (tid < n) {
            const float4 *acc = reinterpret_cast<const float4 *>(s_acc) + (kGradSlots / 4) * tid;
            const float4 a2 = acc[2];
            if (a2.z > 0.f) {  // at least one pixel of this tile touched the splat
                const float4 a0 = acc[0], a1 = acc[1];
                float *dst = splat_grads + (size_t)kGradSlots * s_idx[tid];
                red_add_v4(dst, a0.x, a0.y, a0.z, a0.w);
                red_add_v4(dst + 4, a1.x, a1.y, a1.z, a1.w);
                red_add_v4(dst + 8, a2.x, a2.y, a2.z, 0.f);
            }
        }
    }
    if (map_on_image) {
        map_on_image[2 * px] = tot0;
        map_on_image[2 * px + 1] = tot1;
    }
}

}  // namespace

extern "C" int gs_blend_bwd(int height, int width, const int32_t *tile_start,
                            const int32_t *tile_end, const int32_t *sorted_payload,
                            const float *records, const float *image_grad, const float *acc_alpha,
                            const int32_t *last_effective, float *splat_grads, float *map_on_image,
                            void *stream)
{
    if (height <= 0 || width <= 0 || height % kTile || width % kTile) return GS_E_INVALID;
    if (!tile_start || !tile_end || !image_grad || !acc_alpha || !last_effective) return GS_E_INVALID;
    int tiles = (width / kTile) * (height / kTile);
    blend_bwd_kernel<<<tiles, kTilePixels, 0, (cudaStream_t)stream>>>(
        height, width, tile_start, tile_end, sorted_payload,
        reinterpret_cast<const float4 *>(records), image_grad, acc_alpha, last_effective,
        splat_grads, map_on_image);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
