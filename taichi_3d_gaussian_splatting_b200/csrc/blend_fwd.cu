// blend_fwd.cu -- per-tile front-to-back alpha blend (reference R.py:318-485, utils.py:275-284).
//
// Literal-fp32 profile: build with -fmad=false; the per-pair arithmetic follows
// gso_blend_fwd in oracle/gs_oracle.c operation for operation, so image, depth, accumulated
// alpha, last-effective index and per-pixel counts are bit-exact with the oracle.
//
// One 256-thread CTA per 16x16 tile, one pixel per thread (warp = 2 rows of 16 pixels).  The
// tile's sorted list is streamed through shared memory in batches of 256 packed 48-byte splat
// records, double buffered with cp.async so the gather of batch b+1 overlaps the blend of
// batch b.  A pair whose exponent is below -5.6 cannot reach alpha >= 1/255 (rescale <= 1,
// opacity <= 1), so it is skipped before the exp -- the same decision the reference makes
// after it.  The CTA stops as soon as all 256 pixels are saturated (__syncthreads_and); the
// reference keeps looping without effect (R.py:387-394).
#include "gs_common.cuh"

namespace {

constexpr int kBatch = 256;

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

template <bool kRgbOnly>
__global__ void __launch_bounds__(kTilePixels)
blend_fwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, float *__restrict__ image,
                 float *__restrict__ depth, float *__restrict__ acc_alpha,
                 int32_t *__restrict__ last_effective, int32_t *__restrict__ valid_count)
{
    __shared__ float4 s_rec[2][kBatch * 3];

    const int tid = threadIdx.x;
    const int tiles_per_row = width / kTile;
    const int tile = blockIdx.x;
    const int tile_u = tile % tiles_per_row, tile_v = tile / tiles_per_row;
    const int pu = tile_u * kTile + (tid % kTile), pv = tile_v * kTile + (tid / kTile);
    const float fx = (float)pu + 0.5f, fy = (float)pv + 0.5f;
    const int start = tile_start[tile], end = tile_end[tile];

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f, Wd = 0.f;
    int last = start, cnt = 0;
    bool saturated = false;

    const int num_batches = (end - start + kBatch - 1) / kBatch;
    // prologue: stage batch 0
    if (num_batches > 0) {
        int i = start + tid;
        if (i < end) {
            const float4 *src = records + (size_t)3 * payload[i];
            cp_async16(&s_rec[0][3 * tid], src);
            cp_async16(&s_rec[0][3 * tid + 1], src + 1);
            cp_async16(&s_rec[0][3 * tid + 2], src + 2);
        }
    }
    cp_async_commit();

    for (int b = 0; b < num_batches; ++b) {
        const int buf = b & 1;
        // stage batch b+1 into the other buffer (its last readers finished before the barrier
        // that ended iteration b-1)
        if (b + 1 < num_batches) {
            int i = start + (b + 1) * kBatch + tid;
            if (i < end) {
                const float4 *src = records + (size_t)3 * payload[i];
                cp_async16(&s_rec[buf ^ 1][3 * tid], src);
                cp_async16(&s_rec[buf ^ 1][3 * tid + 1], src + 1);
                cp_async16(&s_rec[buf ^ 1][3 * tid + 2], src + 2);
            }
        }
        cp_async_commit();
        cp_async_wait<1>();  // batch b has landed (for this thread's copies)
        __syncthreads();

        const int base = start + b * kBatch;
        const int n = min(kBatch, end - base);
        if (!saturated) {
            for (int k = 0; k < n; ++k) {
                const float4 r0 = s_rec[buf][3 * k];      // u v a b
                const float4 r1 = s_rec[buf][3 * k + 1];  // c rescale alpha red
                float dx = fx - r0.x, dy = fy - r0.y;
                float e = -0.5f * (dx * dx * r0.z + dy * dy * r1.x) - dx * dy * r0.w;
                // exp(-5.6) * 1.0001 < 1/255 and opacity <= 1: same decision as the a < 1/255 test
                if (e < -5.6f && r1.y <= 1.0001f) continue;
                float ga = gs_expf(e) * r1.y;
                float a = ga * r1.z;
                if (a < 1.0f / 255.0f) continue;
                a = fminf(a, 0.99f);
                float nT = T * (1.0f - a);
                if (nT < 0.0001f) {
                    saturated = true;
                    break;
                }
                const float4 r2 = s_rec[buf][3 * k + 2];  // green blue z pad
                last = base + k + 1;
                C0 += r1.w * a * T;
                C1 += r2.x * a * T;
                C2 += r2.y * a * T;
                if (!kRgbOnly) {
                    D += r2.z * a * T;
                    Wd += a * T;
                    cnt += 1;
                }
                T = nT;
            }
        }
        // everyone is done with buffer `buf`; also the whole-tile early exit
        if (__syncthreads_and(saturated ? 1 : 0)) break;
    }
    cp_async_wait<0>();

    const size_t px = (size_t)pv * width + pu;
    image[3 * px] = C0;
    image[3 * px + 1] = C1;
    image[3 * px + 2] = C2;
    if (!kRgbOnly) {
        depth[px] = D / fmaxf(Wd, 1e-6f);
        acc_alpha[px] = 1.0f - T;
        last_effective[px] = last;
        valid_count[px] = cnt;
    }
}

}  // namespace

extern "C" int gs_blend_fwd(int height, int width, const int32_t *tile_start,
                            const int32_t *tile_end, const int32_t *sorted_payload,
                            const float *records, float *image, float *depth, float *acc_alpha,
                            int32_t *last_effective, int32_t *valid_count, int rgb_only,
                            void *stream)
{
    if (height <= 0 || width <= 0 || height % kTile || width % kTile) return GS_E_INVALID;
    if (!tile_start || !tile_end || !image) return GS_E_INVALID;
    if (!rgb_only && (!depth || !acc_alpha || !last_effective || !valid_count)) return GS_E_INVALID;
    int tiles = (width / kTile) * (height / kTile);
    const float4 *rec = reinterpret_cast<const float4 *>(records);
    cudaStream_t st = (cudaStream_t)stream;
    if (rgb_only)
        blend_fwd_kernel<true><<<tiles, kTilePixels, 0, st>>>(height, width, tile_start, tile_end,
                                                              sorted_payload, rec, image, depth,
                                                              acc_alpha, last_effective, valid_count);
    else
        blend_fwd_kernel<false><<<tiles, kTilePixels, 0, st>>>(height, width, tile_start, tile_end,
                                                               sorted_payload, rec, image, depth,
                                                               acc_alpha, last_effective, valid_count);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
