// blend_bwd.cu -- reverse-order blend backward (reference R.py:531-705, utils.py:331-348).
//
// Fast-fp32 profile (FMA contraction, ex2.approx): gradients are compared with the oracle at
// 1e-4 relative, the atomic summation order differs from the reference anyway.
//
// One WARP per 16x16 tile; lane l owns the 8 pixels (x = l % 16, y = 2 j + l / 16), j = 0..7
// ("slot" j = the 16x2 pixel strip of rows 2j, 2j+1).  The tile's list is walked from
// max-over-pixels(last effective index) down to the tile start (entries past that maximum are
// skipped by every pixel in the reference too, R.py:609-610) in batches of 32: lane l stages
// entry l of the batch and computes an 8-bit slot mask for it -- bit j is set only if the splat
// can reach alpha >= 1/255 on some pixel centre of strip j (exact minimum of the conic's
// quadratic form over each pixel row, with a safety margin) and some pixel of the strip still
// has that entry before its last effective index.  The warp then visits only (splat, strip)
// combinations whose bit is set, sums the 10 per-pixel gradient terms over the strips in
// registers, and reduces them across the warp once per (tile, splat) with a recursive-halving
// butterfly (16 shuffles), followed by three 16-byte vector reductions into the [M,12]
// splat-gradient buffer.  The reference issues 11 global atomics per (pixel, splat) pair
// (R.py:674-696).
#include <cstdlib>

#include "gs_common.cuh"

namespace {

constexpr int kBatch = 32;

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(addr), "f"(a), "f"(b),
                 "f"(c), "f"(d)
                 : "memory");
}

// single-instruction MUFU approximations (fast-fp32 profile; ~1 ulp, gradients are checked at 1e-4)
__device__ __forceinline__ float fast_ex2(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_rcp(float x)
{
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_sqrt(float x)
{
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
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

// Bit j set <=> the splat may reach alpha >= 1/255 at a pixel centre of strip j of this tile.
// q(d) = a dx^2 + 2 b dx dy + c dy^2 must stay <= tau = 2 ln(255 * rescale * alpha) there; per
// pixel row the minimum over the row's 16 centres is taken at dx clamped to the row, so the test
// is exact up to the margin (0.05 absolute + 4e-6 relative to the positive terms, far above the
// rounding of either arithmetic profile).  Degenerate conics keep every bit.
template <int kRows>
__device__ __forceinline__ unsigned strip_mask(float u, float v, float a, float b, float c,
                                               float opacity, float x_first, float y_first)
{
    if (!(opacity * 255.0f >= 0.97f)) return 0u;         // alpha <= rescale*alpha < 1/255 everywhere
    if (!(a > 0.0f) || !(c > 0.0f) || !isfinite(a + b + c + u + v)) return (1u << (kRows / 2)) - 1u;
    const float tau = 2.0f * __logf(opacity * 255.0f) + 0.05f;
    const float inv_a = 1.0f / a;
    const float lo = x_first - u, hi = lo + 15.0f;
    unsigned mask = 0u;
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
        const float dy = (y_first + (float)r) - v;
        const float dx = fminf(fmaxf(-b * dy * inv_a, lo), hi);
        const float pos = a * dx * dx + c * dy * dy;
        const float q = pos + 2.0f * b * dx * dy;
        if (q <= tau + 4e-6f * pos) mask |= 1u << (r >> 1);
    }
    return mask;
}

// kSlots strips (2 kSlots pixel rows) of one tile per warp: 8 = whole tile, 4 = half, 2 = quarter.
template <int kSlots>
__global__ void __launch_bounds__(32)
blend_bwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, const float *__restrict__ image_grad,
                 const float *__restrict__ acc_alpha, const int32_t *__restrict__ last_effective,
                 float *__restrict__ splat_grads, float *__restrict__ map_on_image)
{
    __shared__ float4 s_rec[kBatch * 3];
    __shared__ __align__(16) float s_out[kBatch * kGradSlots];
    __shared__ unsigned s_mask[kBatch];

    const int lane = threadIdx.x;
    const int tiles_per_row = width / kTile;
    constexpr int kParts = 8 / kSlots;  // warps (CTAs) per tile
    const int tile = blockIdx.x / kParts, part = blockIdx.x % kParts;
    const int tile_u = tile % tiles_per_row, tile_v = tile / tiles_per_row;
    const int row0 = tile_v * kTile + part * (2 * kSlots);  // first pixel row of this warp
    const int pu = tile_u * kTile + (lane & 15);
    const int pv0 = row0 + (lane >> 4);
    const float fx = (float)pu + 0.5f, fy0 = (float)pv0 + 0.5f;
    const float x_first = (float)(tile_u * kTile) + 0.5f, y_first = (float)row0 + 0.5f;
    const int start = tile_start[tile], end = tile_end[tile];

    float T[kSlots], w0[kSlots], w1[kSlots], w2[kSlots], g0[kSlots], g1[kSlots], g2[kSlots];
    float tot0[kSlots], tot1[kSlots];
    int last[kSlots], slot_last[kSlots];
    int hi = start;
#pragma unroll
    for (int j = 0; j < kSlots; ++j) {
        const size_t px = (size_t)(pv0 + 2 * j) * width + pu;
        last[j] = (end > start) ? min(last_effective[px], end) : start;
        T[j] = 1.0f - acc_alpha[px];
        g0[j] = image_grad[3 * px];
        g1[j] = image_grad[3 * px + 1];
        g2[j] = image_grad[3 * px + 2];
        w0[j] = w1[j] = w2[j] = tot0[j] = tot1[j] = 0.f;
        slot_last[j] = __reduce_max_sync(0xffffffffu, last[j]);
        hi = max(hi, slot_last[j]);
    }

    for (int top = hi; top > start; top -= kBatch) {
        const int n = min(kBatch, top - start);
        // ---- stage: lane l takes entry top-1-l (reverse order) ------------------------------
        int my_g = 0;
        unsigned my_mask = 0u;
        if (lane < n) {
            const int i = top - 1 - lane;
            my_g = payload[i];
            const float4 *src = records + (size_t)3 * my_g;
            const float4 r0 = __ldg(src), r1 = __ldg(src + 1), r2 = __ldg(src + 2);
            s_rec[3 * lane] = r0;
            s_rec[3 * lane + 1] = r1;
            s_rec[3 * lane + 2] = r2;
            unsigned live = 0u;
#pragma unroll
            for (int j = 0; j < kSlots; ++j) live |= (i < slot_last[j] ? 1u : 0u) << j;
            my_mask = live ? (live & strip_mask<2 * kSlots>(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y * r1.z, x_first,
                                                y_first))
                           : 0u;
        }
        s_mask[lane] = my_mask;
        s_out[kGradSlots * lane + 10] = 0.f;
        __syncwarp();

        for (int k = 0; k < n; ++k) {
            const unsigned m = s_mask[k];
            if (m == 0u) continue;  // warp-uniform
            const int i = top - 1 - k;
            const float4 r0 = s_rec[3 * k];      // u v a b
            const float4 r1 = s_rec[3 * k + 1];  // c rescale alpha red
            const float4 r2 = s_rec[3 * k + 2];  // green blue z pad
            const float dx = fx - r0.x;
            const float adx = r0.z * dx, bdx = r0.w * dx;
            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = 0.f;
            int count = 0;
#pragma unroll
            for (int j = 0; j < kSlots; ++j) {
                if (!((m >> j) & 1u)) continue;  // warp-uniform
                const float dy = (fy0 + (float)(2 * j)) - r0.y;
                const float v0 = adx + r0.w * dy;
                const float v1 = bdx + r1.x * dy;
                const float quad = dx * v0 + dy * v1;
                const float p = fast_ex2(-0.72134752044448170368f * quad) * r1.y;  // exp(-quad/2)
                const float prod = p * r1.z;
                const bool contributes = (i < last[j]) && (prod >= 1.0f / 255.0f);
                count += __popc(__ballot_sync(0xffffffffu, contributes));
                if (contributes) {
                    const float a = fminf(prod, 0.99f);
                    const float inv = fast_rcp(1.0f - a);
                    const float Tn = T[j] * inv;
                    T[j] = Tn;
                    const float dcol = a * Tn;
                    const float agr = (r1.w * Tn - w0[j] * inv) * g0[j] +
                                      (r2.x * Tn - w1[j] * inv) * g1[j] +
                                      (r2.y * Tn - w2[j] * inv) * g2[j];
                    w0[j] += r1.w * dcol;
                    w1[j] += r2.x * dcol;
                    w2[j] += r2.y * dcol;
                    const float sa = agr * prod;  // dL/dp * p
                    const float gm0 = sa * v0, gm1 = sa * v1;
                    tot0[j] += fabsf(gm0);
                    tot1[j] += fabsf(gm1);
                    const float hs = 0.5f * sa;
                    v[0] += gm0;
                    v[1] += gm1;
                    v[2] += hs * v0 * v0;
                    v[3] += hs * v0 * v1;
                    v[4] += hs * v1 * v1;
                    v[5] += dcol * g0[j];
                    v[6] += dcol * g1[j];
                    v[7] += dcol * g2[j];
                    v[8] += sa;
                    v[9] += fast_sqrt(gm0 * gm0 + gm1 * gm1);
                }
            }
            if (count == 0) continue;  // warp-uniform
            v[8] *= (1.0f - r1.z);     // dlogit = sum(sa) * (1 - opacity)
            warp_reduce16(v, lane);
            const int q = lane >> 1;
            if (!(lane & 1)) {
                if (q < 10)
                    s_out[kGradSlots * k + q] = v[0];
                else if (q == 10)
                    s_out[kGradSlots * k + 10] = (float)count;
            }
        }
        __syncwarp();
        // ---- flush: lane l owns entry l of the batch ------------------------------------------
        if (lane < n && my_mask != 0u) {
            const float4 *acc = reinterpret_cast<const float4 *>(s_out) + (kGradSlots / 4) * lane;
            const float4 a2 = acc[2];
            if (a2.z > 0.f) {  // at least one pixel of this tile touched the splat
                const float4 a0 = acc[0], a1 = acc[1];
                float *dst = splat_grads + (size_t)kGradSlots * my_g;
                red_add_v4(dst, a0.x, a0.y, a0.z, a0.w);
                red_add_v4(dst + 4, a1.x, a1.y, a1.z, a1.w);
                red_add_v4(dst + 8, a2.x, a2.y, a2.z, 0.f);
            }
        }
        __syncwarp();
    }
    if (map_on_image) {
#pragma unroll
        for (int j = 0; j < kSlots; ++j) {
            const size_t px = (size_t)(pv0 + 2 * j) * width + pu;
            reinterpret_cast<float2 *>(map_on_image)[px] = make_float2(tot0[j], tot1[j]);
        }
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
    static int slots = 0;
    if (slots == 0) {
        const char *e = getenv("GS_BWD_SLOTS");  // tuning knob: strips per warp (2, 4 or 8)
        slots = e ? atoi(e) : 4;
        if (slots != 2 && slots != 4 && slots != 8) slots = 4;
    }
    const float4 *rec = reinterpret_cast<const float4 *>(records);
    cudaStream_t st = (cudaStream_t)stream;
    if (slots == 8)
        blend_bwd_kernel<8><<<tiles, 32, 0, st>>>(height, width, tile_start, tile_end, sorted_payload, rec,
                                                  image_grad, acc_alpha, last_effective, splat_grads,
                                                  map_on_image);
    else if (slots == 4)
        blend_bwd_kernel<4><<<tiles * 2, 32, 0, st>>>(height, width, tile_start, tile_end, sorted_payload,
                                                      rec, image_grad, acc_alpha, last_effective,
                                                      splat_grads, map_on_image);
    else
        blend_bwd_kernel<2><<<tiles * 4, 32, 0, st>>>(height, width, tile_start, tile_end, sorted_payload,
                                                      rec, image_grad, acc_alpha, last_effective,
                                                      splat_grads, map_on_image);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
