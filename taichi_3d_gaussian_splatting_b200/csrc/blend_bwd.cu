// blend_bwd.cu -- reverse-order blend backward (reference R.py:531-705, utils.py:331-348).
//
// Fast-fp32 profile (FMA contraction, ex2.approx, rcp.approx): gradients are compared with the
// oracle at 1e-4 relative, the atomic summation order differs from the reference anyway.
// The DECISION whether a (pixel, entry) pair takes part (alpha >= 1/255, R.py:634) is nevertheless
// the reference's: a pair whose fast alpha lies within a per-entry error band of 1/255, and every
// pair of an entry with an irregular conic / non-finite field, is re-decided with the reference's
// backward arithmetic in literal fp32 (literal_alpha below = utils.py:331-345 with gs_expf), so
// the set of contributing pairs -- num_affected_pixels -- equals the literal-fp32 oracle's exactly.
//
// Instruction-issue bound like the forward blend, and laid out the same way: one WARP owns a
// column of kCells cells of 8x4 pixels (kCells = 4: an 8x16 half tile); lane l owns pixel
// (l % 8, l / 8) of every cell.  The tile's list is walked from max-over-pixels(last effective
// index) down to the tile start (entries past that maximum are skipped by every pixel in the
// reference too, R.py:609-610) in batches of 32: lane l stages entry l of the batch (records
// prefetched one batch ahead into registers) and computes a kCells-bit mask -- bit j is set only
// if the splat can reach alpha >= 1/255 on some pixel centre of cell j (gs_cell_mask) and some
// pixel of the cell still has that entry before its last effective index.  The warp then visits
// only (entry, cell) combinations whose bit is set, sums the per-pixel gradient terms over its
// cells in registers, and reduces them across the warp once per (warp, entry) with a
// recursive-halving butterfly over 10 values (12 shuffles), followed by three 16-byte vector
// reductions into the [M,12] splat-gradient buffer.  The reference issues 11 global atomics per
// (pixel, splat) pair (R.py:674-696).
//
// Per pixel the three colour channels are folded: with cg = colour . dL/dpixel, the reference's
// sum_c (colour_c T - w_c / (1 - alpha)) dL/dpixel_c equals T cg - wg / (1 - alpha), where
// wg = sum over the entries behind of (alpha T) cg, so one accumulator replaces three.
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

// alpha of a pair exactly as the reference's backward forms it (utils.py:331-345, R.py:628-631) in
// literal fp32: v = conic d (Taichi's mat @ vec: a[i,0] b[0] + a[i,1] b[1]), quad = d . v,
// p = exp(-0.5 quad) rescale, alpha = p opacity.  _rn intrinsics are never contracted.
__device__ __forceinline__ float literal_alpha(float dx, float dy, float a, float b, float c,
                                               float rescale, float opacity)
{
    const float v0 = __fadd_rn(__fmul_rn(a, dx), __fmul_rn(b, dy));
    const float v1 = __fadd_rn(__fmul_rn(b, dx), __fmul_rn(c, dy));
    const float quad = __fadd_rn(__fmul_rn(dx, v0), __fmul_rn(dy, v1));
    return __fmul_rn(__fmul_rn(gs_expf(__fmul_rn(-0.5f, quad)), rescale), opacity);
}

// Relative half-width of the band around 1/255 inside which the fast alpha (FMA-contracted quadratic
// form, ex2.approx, pre-multiplied peak) and the literal one may fall on different sides.  The two
// exponents differ by at most ~5e-7 of the sum P of the magnitudes of the quadratic form's terms, and
// P <= quad (1 + rho) / (1 - rho) <= 4 quad a c / (a c - b^2) with rho = |b| / sqrt(a c); near the
// threshold quad = 2 ln(255 peak) <= 11.1.  The constants leave a factor > 4 of slack; the exp
// approximations and the products add < 2e-6.
__device__ __forceinline__ float decision_band(float a, float b, float c)
{
    const float ac = a * c;
    return fmaf(2.7e-5f, ac * fast_rcp(ac - b * b), 3e-6f);
}

// Sum 10 per-lane values over the 32 lanes of a warp with 12 shuffles.  On return v[0] of lane L
// holds the total of value index q = 5 b4 + 3 b3 + 2 b2 + b1 (b_i = bit i of L) when
// (b3, b2, b1) is one of 000 001 010 100 101, and padding otherwise; see reduced_slot().
__device__ __forceinline__ void warp_reduce10(float (&v)[10], int lane)
{
    {
        const bool up = lane & 16;  // 10 -> 5: lower lanes keep 0..4, upper lanes keep 5..9
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const float send = up ? v[i] : v[i + 5];
            const float keep = up ? v[i + 5] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool up = lane & 8;  // 5 -> 3: lower keep x0 x1 x2, upper keep x3 x4 (pad)
        const float s0 = up ? v[0] : v[3], k0 = up ? v[3] : v[0];
        const float s1 = up ? v[1] : v[4], k1 = up ? v[4] : v[1];
        const float s2 = up ? v[2] : 0.0f, k2 = up ? 0.0f : v[2];
        v[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 8);
        v[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 8);
        v[2] = k2 + __shfl_xor_sync(0xffffffffu, s2, 8);
    }
    {
        const bool up = lane & 4;  // 3 -> 2: lower keep y0 y1, upper keep y2 (pad)
        const float s0 = up ? v[0] : v[2], k0 = up ? v[2] : v[0];
        const float s1 = up ? v[1] : 0.0f, k1 = up ? 0.0f : v[1];
        v[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 4);
        v[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 4);
    }
    {
        const bool up = lane & 2;  // 2 -> 1
        const float s0 = up ? v[0] : v[1], k0 = up ? v[1] : v[0];
        v[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// value index held by lane L after warp_reduce10, or -1 for padding / duplicate lanes
__device__ __forceinline__ int reduced_slot(int lane)
{
    if (lane & 1) return -1;
    const int b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
    if (b2 && (b1 || b3)) return -1;
    return 5 * (lane >> 4) + 3 * b3 + 2 * b2 + b1;
}

template <int kCells, bool kMagnitude, bool kMasks>
__global__ void __launch_bounds__(32 * (kTilePixels / (kCells * 32)))
blend_bwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, const float *__restrict__ image_grad,
                 const float *__restrict__ acc_alpha, const int32_t *__restrict__ last_effective,
                 float *__restrict__ splat_grads, float *__restrict__ map_on_image,
                 const uint8_t *__restrict__ cell_masks, const long long mask_stride,
                 GsWorkCounter *__restrict__ work, const unsigned num_units)
{
    constexpr int kParts = kTilePixels / (kCells * 32);  // warps per tile = units per tile
    constexpr int kPartsPerColumn = kParts / 2;
    __shared__ float4 s_rec_all[kParts][kBatch * 3];
    __shared__ __align__(16) float s_out_all[kParts][kBatch * kGradSlots];

    const int lane = threadIdx.x & 31;
    float4 *s_rec = s_rec_all[threadIdx.x >> 5];
    float *s_out = s_out_all[threadIdx.x >> 5];
    const int tiles_per_row = width / kTile;
    const int my_slot = reduced_slot(lane);
    // unit = (tile, part).  Persistent launch (work != nullptr): every warp pulls units from the
    // counter until they run out; otherwise one CTA per tile, warp w takes part w.
    for (;;) {
    unsigned unit = blockIdx.x * kParts + (threadIdx.x >> 5);
    if (work) {
        unit = gs_next_unit(work);
        if (unit >= num_units) break;
    }
    const int tile = unit / kParts, part = unit % kParts;
    const int tile_u = tile % tiles_per_row, tile_v = tile / tiles_per_row;
    const int col0 = tile_u * kTile + (part / kPartsPerColumn) * kCellW;
    const int row0 = tile_v * kTile + (part % kPartsPerColumn) * (kCells * kCellH);
    const int pu = col0 + (lane & 7);
    const int pv0 = row0 + (lane >> 3);
    const float fx = (float)pu + 0.5f, fy0 = (float)pv0 + 0.5f;
    const float x_first = (float)col0 + 0.5f, y_first = (float)row0 + 0.5f;
    const int start = tile_start[tile], end = tile_end[tile];

    float T[kCells], wg[kCells], g0[kCells], g1[kCells], g2[kCells];
    float tot0[kCells], tot1[kCells];
    int last[kCells], cell_last[kCells];
    int hi = start;
#pragma unroll
    for (int j = 0; j < kCells; ++j) {
        const size_t px = (size_t)(pv0 + kCellH * j) * width + pu;
        last[j] = (end > start) ? min(last_effective[px], end) : start;
        T[j] = 1.0f - acc_alpha[px];
        g0[j] = image_grad[3 * px];
        g1[j] = image_grad[3 * px + 1];
        g2[j] = image_grad[3 * px + 2];
        wg[j] = tot0[j] = tot1[j] = 0.f;
        cell_last[j] = __reduce_max_sync(0xffffffffu, last[j]);
        hi = max(hi, cell_last[j]);
    }

    // software pipeline: records of the next batch in registers, payload two batches ahead
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0;
    int g_cur = 0, g_next = 0;
    {
        const int i = hi - 1 - lane;
        if (i >= start) {
            g_cur = __ldg(payload + i);
            const float4 *src = records + (size_t)3 * g_cur;
            n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
        }
        if (i - kBatch >= start) g_next = __ldg(payload + i - kBatch);
    }

    for (int top = hi; top > start; top -= kBatch) {
        // ---- stage: lane l holds entry top-1-l (reverse order) ---------------------------------
        const int my_i = top - 1 - lane;
        const int my_g = g_cur;
        unsigned my_mask = 0u;  // cell bits | kCellMaskIrregular
        if (my_i >= start) {
            unsigned live = 0u;
#pragma unroll
            for (int j = 0; j < kCells; ++j) live |= (my_i < cell_last[j] ? 1u : 0u) << j;
            if (live) {
                unsigned m = 0u;
                if (kMasks) {
                    // the forward's culling bits for these cells: plane q = 2 * column + row half
                    // of the tile's 8x8 quarters, two bits (upper / lower 8x4 cell) per entry, and
                    // the entry's irregular flag
                    unsigned irr = 0u;
#pragma unroll
                    for (int h = 0; h < kCells / 2; ++h) {
                        const int q = 2 * (part / kPartsPerColumn) + (part % kPartsPerColumn) * (kCells / 2) + h;
                        const unsigned byte = (unsigned)__ldg(cell_masks + q * mask_stride + my_i);
                        m |= (byte & 3u) << (2 * h);
                        irr |= byte & kMaskByteIrregular;
                    }
                    if (irr) m |= kCellMaskIrregular;
                } else {
                    m = gs_entry_mask<kCells>(n0, n1, n2, x_first, y_first);
                }
                if (live & m) my_mask = (live & m) | (m & kCellMaskIrregular);
            }
        }
        s_rec[3 * lane] = n0;
        s_rec[3 * lane + 1] = n1;
        // the record's pad word carries the entry's decision band to the visits
        s_rec[3 * lane + 2] = make_float4(n2.x, n2.y, n2.z, decision_band(n0.z, n0.w, n1.x));
        s_out[kGradSlots * lane + 10] = 0.f;
        // ---- prefetch the next batch -----------------------------------------------------------
        {
            const int i = my_i - kBatch;
            g_cur = g_next;
            if (i >= start) {
                const float4 *src = records + (size_t)3 * g_cur;
                n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
            }
            if (i - kBatch >= start) g_next = __ldg(payload + i - kBatch);
        }
        __syncwarp();

        unsigned pending = __ballot_sync(0xffffffffu, my_mask != 0u);
        while (pending) {
            const int k = __ffs(pending) - 1;
            pending &= pending - 1u;
            const unsigned m = __shfl_sync(0xffffffffu, my_mask, k);
            const int i = top - 1 - k;
            const float4 r0 = s_rec[3 * k];      // u v a b
            const float4 r1 = s_rec[3 * k + 1];  // c rescale alpha red
            const float4 r2 = s_rec[3 * k + 2];  // green blue z band
            const float dx = fx - r0.x;
            const float adx = r0.z * dx, bdx = r0.w * dx;
            const float peak = r1.y * r1.z;
            // fast alpha decides outside [lo, hi); inside, and for irregular entries (hi = NaN: no
            // alpha is ever "safely above"), the literal arithmetic does
            const bool irregular = (m & kCellMaskIrregular) != 0u;  // warp-uniform
            const float lo = irregular ? -__int_as_float(0x7f800000) : fmaf(-r2.w, 1.0f / 255.0f, 1.0f / 255.0f);
            const float hi = irregular ? __int_as_float(0x7fc00000) : fmaf(r2.w, 1.0f / 255.0f, 1.0f / 255.0f);
            float v[10];
#pragma unroll
            for (int q = 0; q < 10; ++q) v[q] = 0.f;
            int hits = 0;
#pragma unroll
            for (int j = 0; j < kCells; ++j) {
                if (!((m >> j) & 1u)) continue;  // warp-uniform
                const float dy = (fy0 + (float)(kCellH * j)) - r0.y;
                const float v0 = fmaf(r0.w, dy, adx);
                const float v1 = fmaf(r1.x, dy, bdx);
                const float quad = fmaf(dy, v1, dx * v0);
                float prod = fast_ex2(-0.72134752044448170368f * quad) * peak;  // alpha
                bool take = prod >= lo;
                if (!(prod >= hi) && (take || irregular)) {  // rare: on the edge, or an irregular entry
                    const float lit = literal_alpha(dx, dy, r0.z, r0.w, r1.x, r1.y, r1.z);
                    take = lit >= 1.0f / 255.0f;
                    if (irregular) prod = lit;  // +inf / NaN handling of the literal exp
                }
                if ((i < last[j]) && take) {
                    hits += 1;
                    const float a = fminf(prod, 0.99f);
                    const float inv = fast_rcp(1.0f - a);
                    const float Tn = T[j] * inv;
                    T[j] = Tn;
                    const float dcol = a * Tn;
                    const float cg = fmaf(r2.y, g2[j], fmaf(r2.x, g1[j], r1.w * g0[j]));
                    const float sa = (Tn * cg - inv * wg[j]) * prod;  // dL/dalpha * alpha
                    wg[j] = fmaf(dcol, cg, wg[j]);
                    const float gm0 = sa * v0, gm1 = sa * v1;
                    v[0] += gm0;
                    v[1] += gm1;
                    v[2] = fmaf(gm0, v0, v[2]);
                    v[3] = fmaf(gm0, v1, v[3]);
                    v[4] = fmaf(gm1, v1, v[4]);
                    v[5] = fmaf(dcol, g0[j], v[5]);
                    v[6] = fmaf(dcol, g1[j], v[6]);
                    v[7] = fmaf(dcol, g2[j], v[7]);
                    v[8] += sa;
                    if (kMagnitude) {
                        tot0[j] += fabsf(gm0);
                        tot1[j] += fabsf(gm1);
                        v[9] += fast_sqrt(fmaf(gm0, gm0, gm1 * gm1));
                    }
                }
            }
            const int count = __reduce_add_sync(0xffffffffu, hits);
            if (count == 0) continue;  // warp-uniform
            v[2] *= 0.5f;
            v[3] *= 0.5f;
            v[4] *= 0.5f;
            v[8] *= (1.0f - r1.z);  // dlogit = sum(sa) * (1 - opacity)
            warp_reduce10(v, lane);
            if (my_slot >= 0) s_out[kGradSlots * k + my_slot] = v[0];
            if (lane == 1) s_out[kGradSlots * k + 10] = (float)count;
        }
        __syncwarp();
        // ---- flush: lane l owns entry l of the batch ------------------------------------------
        if (my_mask != 0u) {
            const float4 *acc = reinterpret_cast<const float4 *>(s_out) + (kGradSlots / 4) * lane;
            const float4 a2 = acc[2];
            if (a2.z > 0.f) {  // at least one pixel of this warp touched the splat
                const float4 a0 = acc[0], a1 = acc[1];
                float *dst = splat_grads + (size_t)kGradSlots * my_g;
                red_add_v4(dst, a0.x, a0.y, a0.z, a0.w);
                red_add_v4(dst + 4, a1.x, a1.y, a1.z, a1.w);
                red_add_v4(dst + 8, a2.x, kMagnitude ? a2.y : 0.f, a2.z, 0.f);
            }
        }
        __syncwarp();
    }
    if (kMagnitude) {
#pragma unroll
        for (int j = 0; j < kCells; ++j) {
            const size_t px = (size_t)(pv0 + kCellH * j) * width + pu;
            reinterpret_cast<float2 *>(map_on_image)[px] = make_float2(tot0[j], tot1[j]);
        }
    }
    if (!work) return;
    __syncwarp();  // the next unit restages s_rec / s_out
    }
    gs_retire_warp(work, gridDim.x * kParts);
}

template <int kCells>
int launch_bwd(int tiles, cudaStream_t st, int height, int width, const int32_t *tile_start,
               const int32_t *tile_end, const int32_t *payload, const float4 *rec,
               const float *image_grad, const float *acc_alpha, const int32_t *last_effective,
               float *splat_grads, float *map_on_image, const uint8_t *cell_masks, long long mask_stride)
{
    constexpr int kParts = kTilePixels / (kCells * 32);
    constexpr int kThreads = 32 * kParts;
    // persistent launch: as many CTAs as are resident at once, warps pull (tile, part) units
    static int resident[4] = {0, 0, 0, 0};
    int &res = resident[(map_on_image ? 2 : 0) + (cell_masks ? 1 : 0)];
    if (res == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        GS_CUDA_TRY(cudaGetDevice(&dev));
        GS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
#define GS_BWD_OCC(MAG, MSK) \
    GS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, blend_bwd_kernel<kCells, MAG, MSK>, kThreads, 0))
        if (map_on_image) {
            if (cell_masks) GS_BWD_OCC(true, true); else GS_BWD_OCC(true, false);
        } else {
            if (cell_masks) GS_BWD_OCC(false, true); else GS_BWD_OCC(false, false);
        }
#undef GS_BWD_OCC
        res = sms * (per_sm > 0 ? per_sm : 1);
    }
    const char *pe = getenv("GS_BLEND_PERSISTENT");  // 0: one CTA per tile (A/B and tests; read per call)
    GsWorkCounter *work = (pe && atoi(pe) == 0) ? nullptr : gs_work_counter(st);
    const unsigned units = (unsigned)tiles * kParts;
    const int grid = work ? (tiles < res ? tiles : res) : tiles;
#define GS_BWD_LAUNCH(MAG, MSK)                                                                        \
    blend_bwd_kernel<kCells, MAG, MSK><<<grid, kThreads, 0, st>>>(                                      \
        height, width, tile_start, tile_end, payload, rec, image_grad, acc_alpha, last_effective,      \
        splat_grads, map_on_image, cell_masks, mask_stride, work, units)
    if (map_on_image) {
        if (cell_masks) GS_BWD_LAUNCH(true, true); else GS_BWD_LAUNCH(true, false);
    } else {
        if (cell_masks) GS_BWD_LAUNCH(false, true); else GS_BWD_LAUNCH(false, false);
    }
#undef GS_BWD_LAUNCH
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace

extern "C" int gs_blend_bwd(int height, int width, const int32_t *tile_start,
                            const int32_t *tile_end, const int32_t *sorted_payload,
                            const float *records, const float *image_grad, const float *acc_alpha,
                            const int32_t *last_effective, float *splat_grads, float *map_on_image,
                            const uint8_t *cell_masks, int64_t mask_stride, void *stream)
{
    if (cell_masks && mask_stride < 0) return GS_E_INVALID;
    if (height <= 0 || width <= 0 || height % kTile || width % kTile) return GS_E_INVALID;
    if (!tile_start || !tile_end || !image_grad || !acc_alpha || !last_effective || !splat_grads)
        return GS_E_INVALID;
    int tiles = (width / kTile) * (height / kTile);
    static int cells = 0;
    if (cells == 0) {
        const char *e = getenv("GS_BWD_CELLS");  // tuning knob: 8x4 cells per warp (2 or 4)
        cells = e ? atoi(e) : 4;
        if (cells != 2 && cells != 4) cells = 4;
    }
    const float4 *rec = reinterpret_cast<const float4 *>(records);
    cudaStream_t st = (cudaStream_t)stream;
    if (cells == 2)
        return launch_bwd<2>(tiles, st, height, width, tile_start, tile_end, sorted_payload, rec,
                             image_grad, acc_alpha, last_effective, splat_grads, map_on_image, cell_masks,
                             mask_stride);
    return launch_bwd<4>(tiles, st, height, width, tile_start, tile_end, sorted_payload, rec,
                         image_grad, acc_alpha, last_effective, splat_grads, map_on_image, cell_masks,
                         mask_stride);
}
