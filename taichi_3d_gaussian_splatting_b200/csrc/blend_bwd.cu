// blend_bwd.cu -- reverse-order blend backward (reference R.py:531-705, utils.py:331-348).
//
// Two kernels:
//
// blend_bwd_kernel (hot).  Fast-fp32 profile (FMA contraction, ex2.approx, rcp.approx): gradients are
// compared with the oracle at 1e-4 relative, the atomic summation order differs from the reference
// anyway.  Instruction-issue bound like the forward blend, and laid out the same way: one WARP owns a
// column of kCells cells of 8x4 pixels (kCells = 4: an 8x16 half tile); lane l owns pixel
// (l % 8, l / 8) of every cell.  The tile's list is walked from max-over-pixels(last effective index)
// down to the tile start (entries past that maximum are skipped by every pixel in the reference too,
// R.py:609-610) in batches of 32: lane l stages entry l of the batch (records prefetched one batch
// ahead into registers) together with the culling bits the FORWARD blend stored for it (which 8x4 cells
// the splat can reach at all).  The warp then visits only (entry, cell) combinations whose bit is set,
// sums the per-pixel gradient terms over its cells in registers, and reduces them across the warp once
// per (warp, entry) with a recursive-halving butterfly over 10 values (12 shuffles), followed by three
// 16-byte vector reductions into the [M,12] splat-gradient buffer.  The reference issues 11 global
// atomics per (pixel, splat) pair (R.py:674-696).  Per pixel the three colour channels are folded: with
// cg = colour . dL/dpixel, the reference's sum_c (colour_c T - w_c / (1 - alpha)) dL/dpixel_c equals
// T cg - wg / (1 - alpha), where wg = sum over the entries behind of (alpha T) cg.
//
// The hot loop is kept SMALL on purpose (it is sensitive to its instruction footprint: every variant
// that put the rare cases below inside it, inline or as a call, cost 8-25 %).  What it does not do:
//
//   * Which pairs take part is the reference's decision `alpha >= 1/255` (R.py:634) on ITS fp32
//     arithmetic.  The hot loop takes every pair whose fast alpha is >= lo = (1 - band)/255 and only
//     REMEMBERS, per pixel, whether it met a pair below hi = (1 + band)/255, i.e. one whose literal
//     alpha may fall on the other side (band: gs_decision_band_log2() in gs_common.cuh, a per-splat
//     bound of the difference between the two arithmetics, computed by the projection kernel and
//     carried in the record's 12th float; ~1e-5 of the contributing pairs).  That costs the hot
//     block ONE instruction: alpha is formed as 2^x' hi with x' = log2(alpha / hi), and the sign bits
//     of the x' of the pairs taken are OR-ed together.
//   * Quarters whose list holds an irregular entry (conic not positive definite, non-finite field:
//     the forward blend flags them) are not processed here at all.
//
// blend_bwd_fixup_kernel (cold, one warp per flagged pixel).  Re-does the flagged pixels exactly as the
// reference does -- every pair decided and formed with literal_alpha(), the reference's backward
// arithmetic in literal fp32 with the guarded exp -- and, for a pixel the hot kernel did process, takes
// the hot kernel's own contribution for that pixel out again by replaying its arithmetic with the
// opposite sign.  The set of contributing pairs (num_affected_pixels) therefore EQUALS the literal
// oracle's, irregular conics included.
#include <cstdlib>

#include "gs_common.cuh"

namespace {

constexpr int kBatch = 32;
constexpr float kAlphaMin = 1.0f / 255.0f;
// fix list: scratch = {unsigned count; unsigned entries[H * W]}, entry = pixel index | kFixWholeBit
constexpr unsigned kFixWholeBit = 1u << 31;

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

// Per-entry constants of the decision band.  The record's 12th float, written by gs_project_cull_sh, is
// s = log2(peak / hi) with peak = rescale * opacity and hi = (1 + band) / 255: the upper edge of the band
// inside which the fast alpha and the literal one may fall on different sides of 1/255.  From it:
//   hi  = peak 2^-s                      (so that alpha = 2^x' hi with x' = log2(alpha / hi))
//   xlo = log2(lo / hi), lo = (1 - band)/255, i.e. ~ -2 band log2(e) (rounded away from zero)
// A pair is TAKEN iff x' >= xlo and AMBIGUOUS iff it is taken and x' < 0 (its sign bit).  The lane that
// stages an entry computes (hi, xlo) once and passes them to the visits in the two record slots the
// backward does not need (rescale and depth).
struct BandConsts {
    float hi, xlo;
};
__device__ __forceinline__ BandConsts band_consts(float peak, float s)
{
    BandConsts k;
    k.hi = peak * fast_ex2(-s);
    k.xlo = fmaf(k.hi, -2.8856f * 255.0f, 2.8856f - 4e-6f);  // -2.8856 (255 hi - 1) = -2.8856 band, minus slack
    return k;
}

// alpha of a pair in the fast profile, as (x', alpha); v = conic d comes back for the gradient terms.
// One function for the hot kernel and for the fix-up's replay of it: both must take the same decisions.
__device__ __forceinline__ float fast_alpha(float dx, float dy, float adx, float bdx, float b, float c,
                                            float s, float hi, float &v0, float &v1, float &xp)
{
    v0 = fmaf(b, dy, adx);
    v1 = fmaf(c, dy, bdx);
    const float quad = fmaf(dy, v1, dx * v0);
    xp = fmaf(quad, -0.72134752044448170368f, s);
    return fast_ex2(xp) * hi;
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

// Two entries at once: 20 values in 21 shuffles (against 2 x 12).  The first exchange (xor 16) leaves the
// lower half-warp with entry A's ten half-sums and the upper one with entry B's; the four remaining
// levels are those of warp_reduce10 one bit further down (no duplicate level is left).  On return a[0] of
// lane L holds, of entry (L & 16 ? B : A), the total of value index q = 5 b3 + 3 b2 + 2 b1 + b0 when
// (b2, b1, b0) is one of 000 001 010 100 101; see reduced_slot_pair().
__device__ __forceinline__ void warp_reduce10_pair(float (&a)[10], const float (&b)[10], int lane)
{
    {
        const bool up = lane & 16;
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const float send = up ? a[i] : b[i];
            const float keep = up ? b[i] : a[i];
            a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const float send = up ? a[i] : a[i + 5];
            const float keep = up ? a[i + 5] : a[i];
            a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool up = lane & 4;
        const float s0 = up ? a[0] : a[3], k0 = up ? a[3] : a[0];
        const float s1 = up ? a[1] : a[4], k1 = up ? a[4] : a[1];
        const float s2 = up ? a[2] : 0.0f, k2 = up ? 0.0f : a[2];
        a[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 4);
        a[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 4);
        a[2] = k2 + __shfl_xor_sync(0xffffffffu, s2, 4);
    }
    {
        const bool up = lane & 2;
        const float s0 = up ? a[0] : a[2], k0 = up ? a[2] : a[0];
        const float s1 = up ? a[1] : 0.0f, k1 = up ? 0.0f : a[1];
        a[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 2);
        a[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 2);
    }
    {
        const bool up = lane & 1;
        const float s0 = up ? a[0] : a[1], k0 = up ? a[1] : a[0];
        a[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 1);
    }
}
__device__ __forceinline__ int reduced_slot_pair(int lane)
{
    const int b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1, b0 = lane & 1;
    if (b1 && (b0 || b2)) return -1;
    return 5 * b3 + 3 * b2 + 2 * b1 + b0;
}

// value index held by lane L after warp_reduce10, or -1 for padding / duplicate lanes
__device__ __forceinline__ int reduced_slot(int lane)
{
    if (lane & 1) return -1;
    const int b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
    if (b2 && (b1 || b3)) return -1;
    return 5 * (lane >> 4) + 3 * b3 + 2 * b2 + b1;
}

// GS_BWD_STAGING: how a batch of 32 records reaches shared memory (compile-time; 0 is the product).
//   0  register prefetch: every lane loads its entry's 48 bytes one batch ahead (3 LDG.128) and stores them
//      (3 STS.128) at the top of the batch
//   1  bulk-async ring: every lane issues ONE cp.async.bulk of its entry's 48 bytes into the other half of a
//      2-deep ring, completion on an mbarrier per half (no record registers are carried through the visits)
#ifndef GS_BWD_STAGING
#define GS_BWD_STAGING 0
#endif
#if GS_BWD_STAGING == 1
__device__ __forceinline__ void bwd_mbar_init(unsigned long long *bar, unsigned count)
{
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// every lane arrives once per batch; a lane with an entry announces its 48 bytes and starts the copy
__device__ __forceinline__ void bwd_bulk_stage(void *dst_smem, const void *src, unsigned long long *bar, bool valid)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    if (valid) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 48;" ::"r"(b) : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 48, [%2];" ::"r"(d),
            "l"(src), "r"(b)
            : "memory");
    } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory");
    }
}
__device__ __forceinline__ void bwd_mbar_wait(unsigned long long *bar, unsigned parity)
{
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(b),
        "r"(parity)
        : "memory");
}
#endif

// GS_BWD_PAIR: 1 = the cross-lane sums of two contributing entries are formed together
// (warp_reduce10_pair): an entry with hits is held back until the next one of its batch arrives
#ifndef GS_BWD_PAIR
#define GS_BWD_PAIR 0
#endif

// resident warps per SM the kernel is compiled for: 20 (<= 96 registers; the product path, no
// magnitudes) or 16 when the |duv| magnitudes are accumulated too (they need 8 more registers)
#ifndef GS_BWD_MIN_WARPS
#define GS_BWD_MIN_WARPS 20
#endif
constexpr int bwd_min_ctas(int cells, bool magnitude)
{
    return (magnitude ? 16 : GS_BWD_MIN_WARPS) / (kTilePixels / (cells * 32));
}

template <int kCells, bool kMagnitude>
__global__ void __launch_bounds__(32 * (kTilePixels / (kCells * 32)), bwd_min_ctas(kCells, kMagnitude))
blend_bwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, const float *__restrict__ image_grad,
                 const float *__restrict__ acc_alpha, const int32_t *__restrict__ last_effective,
                 float *__restrict__ splat_grads, float *__restrict__ map_on_image,
                 const uint8_t *__restrict__ cell_masks, const long long mask_stride,
                 const uint8_t *__restrict__ quarter_flags, unsigned *__restrict__ fix_list,
                 GsWorkCounter *__restrict__ work, const unsigned num_units)
{
    constexpr int kParts = kTilePixels / (kCells * 32);  // warps per tile = units per tile
    constexpr int kPartsPerColumn = kParts / 2;
#if GS_BWD_STAGING == 1
    __shared__ __align__(128) float4 s_ring_all[kParts][2][kBatch * 3];
    __shared__ __align__(8) unsigned long long s_bar_all[kParts][2];
#else
    __shared__ float4 s_rec_all[kParts][kBatch * 3];
#endif
    __shared__ __align__(16) float s_out_all[kParts][kBatch * kGradSlots];

    const int lane = threadIdx.x & 31;
#if GS_BWD_STAGING == 1
    float4 (*s_ring)[kBatch * 3] = s_ring_all[threadIdx.x >> 5];
    unsigned long long *s_bar = s_bar_all[threadIdx.x >> 5];
    if (lane == 0) {
        bwd_mbar_init(s_bar, 32);
        bwd_mbar_init(s_bar + 1, 32);
    }
    __syncwarp();
    unsigned parity = 0u;  // bit h: the phase of half h to wait for next
#else
    float4 *s_rec = s_rec_all[threadIdx.x >> 5];
#endif
    float *s_out = s_out_all[threadIdx.x >> 5];
    const int tiles_per_row = width / kTile;
    const int my_slot = reduced_slot(lane);
#if GS_BWD_PAIR
    const int my_pair_slot = reduced_slot_pair(lane);
#endif
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
    const int start = tile_start[tile], end = tile_end[tile];
    const int q0 = 2 * (part / kPartsPerColumn) + (part % kPartsPerColumn) * (kCells / 2);  // first 8x8 quarter

    // a quarter with an irregular entry in reach is left, whole, to the literal pass
    unsigned whole = 0u;  // bit h: quarter q0 + h
#pragma unroll
    for (int h = 0; h < kCells / 2; ++h) whole |= (quarter_flags[4 * tile + q0 + h] ? 1u : 0u) << h;

    float T[kCells], wg[kCells], g0[kCells], g1[kCells], g2[kCells];
    float tot0[kCells], tot1[kCells];
    int last[kCells], cell_last[kCells];
    int hi_index = start;
#pragma unroll
    for (int j = 0; j < kCells; ++j) {
        const size_t px = (size_t)(pv0 + kCellH * j) * width + pu;
        last[j] = (end > start) ? min(last_effective[px], end) : start;
        if ((whole >> (j / 2)) & 1u) {
            fix_list[1u + atomicAdd(fix_list, 1u)] = (unsigned)px | kFixWholeBit;
            last[j] = start;  // nothing of this cell is processed here
        }
        T[j] = 1.0f - acc_alpha[px];
        g0[j] = image_grad[3 * px];
        g1[j] = image_grad[3 * px + 1];
        g2[j] = image_grad[3 * px + 2];
        wg[j] = tot0[j] = tot1[j] = 0.f;
        cell_last[j] = __reduce_max_sync(0xffffffffu, last[j]);
        hi_index = max(hi_index, cell_last[j]);
    }

    // software pipeline: records of the next batch in registers, payload two batches ahead
#if GS_BWD_STAGING == 1
    int g_cur = 0, g_next = 0;
    int half = 0;
    if (hi_index > start) {   // warp-uniform
        const int i = hi_index - 1 - lane;
        if (i >= start) g_cur = __ldg(payload + i);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the half was read / patched by this warp
        bwd_bulk_stage(&s_ring[0][3 * lane], records + (size_t)3 * g_cur, s_bar, i >= start);
        if (i - kBatch >= start) g_next = __ldg(payload + i - kBatch);
    }
#else
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0;
    int g_cur = 0, g_next = 0;
    {
        const int i = hi_index - 1 - lane;
        if (i >= start) {
            g_cur = __ldg(payload + i);
            const float4 *src = records + (size_t)3 * g_cur;
            n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
        }
        if (i - kBatch >= start) g_next = __ldg(payload + i - kBatch);
    }
#endif

    for (int top = hi_index; top > start; top -= kBatch) {
        // ---- stage: lane l holds entry top-1-l (reverse order) ---------------------------------
        const int my_i = top - 1 - lane;
        const int my_g = g_cur;
        unsigned my_mask = 0u;
        if (my_i >= start) {
            unsigned live = 0u;
#pragma unroll
            for (int j = 0; j < kCells; ++j) live |= (my_i < cell_last[j] ? 1u : 0u) << j;
            if (live) {
                // the forward's culling bits for these cells: plane q of the tile's 8x8 quarters, two
                // bits (upper / lower 8x4 cell) per entry
                // (a quarter none of whose cells still reaches this entry may never have had its byte
                // written by the forward: it is not read)
                unsigned m = 0u;
#pragma unroll
                for (int h = 0; h < kCells / 2; ++h)
                    if ((live >> (2 * h)) & 3u)
                        m |= ((unsigned)__ldg(cell_masks + (q0 + h) * mask_stride + my_i) & 3u) << (2 * h);
                my_mask = live & m;
            }
        }
#if GS_BWD_STAGING == 1
        float4 *s_rec = s_ring[half];
        s_out[kGradSlots * lane + 10] = 0.f;
        // ---- start the next batch's copies into the other half ------------------------------------
        {
            const int i = my_i - kBatch;
            g_cur = g_next;
            if (top - kBatch > start) {  // warp-uniform
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                bwd_bulk_stage(&s_ring[half ^ 1][3 * lane], records + (size_t)3 * g_cur, s_bar + (half ^ 1),
                               i >= start);
            }
            if (i - kBatch >= start) g_next = __ldg(payload + i - kBatch);
        }
        // ---- this batch: wait for its 32 copies, then the owner of an entry patches the band constants in
        bwd_mbar_wait(s_bar + half, (parity >> half) & 1u);
        parity ^= 1u << half;
        if (my_i >= start) {
            float4 q1 = s_rec[3 * lane + 1], q2 = s_rec[3 * lane + 2];
            const BandConsts band = band_consts(q1.y * q1.z, q2.w);
            q1.y = band.xlo;   // c xlo opacity red
            q2.z = band.hi;    // green blue hi s
            s_rec[3 * lane + 1] = q1;
            s_rec[3 * lane + 2] = q2;
        }
        half ^= 1;
        __syncwarp();
#else
        {
            const BandConsts band = band_consts(n1.y * n1.z, n2.w);
            s_rec[3 * lane] = n0;
            s_rec[3 * lane + 1] = make_float4(n1.x, band.xlo, n1.z, n1.w);  // c xlo opacity red
            s_rec[3 * lane + 2] = make_float4(n2.x, n2.y, band.hi, n2.w);   // green blue hi s
        }
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
#endif

        unsigned pending = __ballot_sync(0xffffffffu, my_mask != 0u);
#if GS_BWD_PAIR
        float held[10];
        int held_k = -1, held_count = 0;  // warp-uniform
#endif
        while (pending) {
            const int k = __ffs(pending) - 1;
            pending &= pending - 1u;
            const unsigned m = __shfl_sync(0xffffffffu, my_mask, k);
            const int i = top - 1 - k;
            const float4 r0 = s_rec[3 * k];      // u v a b
            const float4 r1 = s_rec[3 * k + 1];  // c xlo opacity red
            const float4 r2 = s_rec[3 * k + 2];  // green blue hi s   (see band_consts)
            const float dx = fx - r0.x;
            const float adx = r0.z * dx, bdx = r0.w * dx;
            float v[10];
#pragma unroll
            for (int q = 0; q < 10; ++q) v[q] = 0.f;
            int hits = 0;
#pragma unroll
            for (int j = 0; j < kCells; ++j) {
                if (!((m >> j) & 1u)) continue;  // warp-uniform
                const float dy = (fy0 + (float)(kCellH * j)) - r0.y;
                float v0, v1, xp;
                const float prod = fast_alpha(dx, dy, adx, bdx, r0.w, r1.x, r2.w, r2.z, v0, v1, xp);
                if ((i < last[j]) && (xp >= r1.y)) {
                    // Taken, but below the band's upper edge (x' < 0): the pixel is AMBIGUOUS.  The one
                    // instruction the hot block spends on it moves the sign bit of x' into last[j]; a
                    // negative last[j] also ends the pixel's traversal here (i < last[j] fails from now
                    // on), which the literal pass's replay of this kernel mirrors.
                    last[j] |= (int)(__float_as_uint(xp) & 0x80000000u);
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
#if GS_BWD_PAIR
            if (held_k < 0) {  // warp-uniform: wait for a partner
#pragma unroll
                for (int q = 0; q < 10; ++q) held[q] = v[q];
                held_k = k;
                held_count = count;
                continue;
            }
            warp_reduce10_pair(held, v, lane);
            {
                const int kk = (lane & 16) ? k : held_k;
                if (my_pair_slot >= 0) s_out[kGradSlots * kk + my_pair_slot] = held[0];
                if (lane == 1) s_out[kGradSlots * held_k + 10] = (float)held_count;
                if (lane == 17) s_out[kGradSlots * k + 10] = (float)count;
            }
            held_k = -1;
#else
            warp_reduce10(v, lane);
            if (my_slot >= 0) s_out[kGradSlots * k + my_slot] = v[0];
            if (lane == 1) s_out[kGradSlots * k + 10] = (float)count;
#endif
        }
#if GS_BWD_PAIR
        if (held_k >= 0) {  // the batch's odd one out
            warp_reduce10(held, lane);
            if (my_slot >= 0) s_out[kGradSlots * held_k + my_slot] = held[0];
            if (lane == 1) s_out[kGradSlots * held_k + 10] = (float)held_count;
        }
#endif
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
#pragma unroll
    for (int j = 0; j < kCells; ++j) {
        const size_t px = (size_t)(pv0 + kCellH * j) * width + pu;
        if (last[j] < 0) fix_list[1u + atomicAdd(fix_list, 1u)] = (unsigned)px;  // the literal pass re-does it
        if (kMagnitude) reinterpret_cast<float2 *>(map_on_image)[px] = make_float2(tot0[j], tot1[j]);
    }
    if (!work) return;
    __syncwarp();  // the next unit restages s_rec / s_out
    }
    gs_retire_warp(work, gridDim.x * kParts);
}

// ---- the literal pass over flagged pixels ---------------------------------------------------------
// One pixel, one traversal of its list from last-1 down to the tile start by the whole warp, 32 entries
// per step (the next step's records are fetched while this one is worked on): every lane forms its
// pair's alpha and decision independently (neither depends on T), then T = T_end / prod(1 - alpha) and
// w = sum(alpha T cg) over the entries behind follow from warp scans.  Two variants side by side on the
// same loads: the reference's arithmetic and decision (contributions ADDED) and, for a pixel the hot
// kernel did process, the hot kernel's arithmetic and (biased) decision (contributions SUBTRACTED).
struct FixVariant {
    bool take;
    float prod, v0, v1, inv, a;
    float T_carry, w_carry;
    float Tn, w_before, term;
};

__device__ __forceinline__ void fix_scan(FixVariant &x, const float cg, const int lane)
{
    // T after this entry = T_carry * (product of inv over lanes 0..l): inclusive scan
    float scan = x.inv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float o = __shfl_up_sync(0xffffffffu, scan, d);
        if (lane >= d) scan *= o;
    }
    x.Tn = x.T_carry * scan;
    // w before this entry = w_carry + sum of (alpha T cg) over lanes 0..l-1: exclusive scan
    x.term = x.take ? x.a * x.Tn * cg : 0.0f;
    float wsum = x.term;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float o = __shfl_up_sync(0xffffffffu, wsum, d);
        if (lane >= d) wsum += o;
    }
    x.w_before = x.w_carry + (wsum - x.term);
    x.T_carry = __shfl_sync(0xffffffffu, x.Tn, 31);
    x.w_carry = __shfl_sync(0xffffffffu, x.w_before + x.term, 31);
}

// the 11 per-pair terms (R.py:645-696) of one variant, times `sign`, into out[]
template <bool kMagnitude>
__device__ __forceinline__ void fix_terms(const FixVariant &x, const float sign, const float cg,
                                          const float opacity, const float g0, const float g1, const float g2,
                                          float (&out)[11], float &gm0, float &gm1)
{
    gm0 = gm1 = 0.f;
    if (!x.take) return;
    const float dcol = x.a * x.Tn;
    const float sa = (x.Tn * cg - x.inv * x.w_before) * x.prod;
    gm0 = sa * x.v0;
    gm1 = sa * x.v1;
    out[0] += sign * gm0;
    out[1] += sign * gm1;
    out[2] += sign * 0.5f * gm0 * x.v0;
    out[3] += sign * 0.5f * gm0 * x.v1;
    out[4] += sign * 0.5f * gm1 * x.v1;
    out[5] += sign * dcol * g0;
    out[6] += sign * dcol * g1;
    out[7] += sign * dcol * g2;
    out[8] += sign * sa * (1.0f - opacity);
    if (kMagnitude) out[9] += sign * sqrtf(fmaf(gm0, gm0, gm1 * gm1));
    out[10] += sign;
}

template <bool kMagnitude>
__device__ __forceinline__ void fixup_pixel(const int lane, const bool replay, const int start, const int last,
                                            const float fx, const float fy, const float T_end,
                                            const float g0, const float g1, const float g2,
                                            const int32_t *__restrict__ payload,
                                            const float4 *__restrict__ records,
                                            float *__restrict__ splat_grads, float &tot0, float &tot1)
{
    FixVariant lit, fast;
    bool stopped = false;  // warp-uniform: the replay has passed the hot kernel's last pair of this pixel
    lit.T_carry = fast.T_carry = T_end;
    lit.w_carry = fast.w_carry = 0.0f;
    // software pipeline as in the hot kernel: records one step ahead, payload two steps ahead, so that a
    // step waits for ONE load latency, not for the dependent payload -> record pair (the pass is a chain of
    // such steps along the longest flagged pixel's list)
    int g_next = 0, g_next2 = 0;
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0;
    {
        const int i = last - 1 - lane;
        if (i >= start) {
            g_next = __ldg(payload + i);
            const float4 *src = records + (size_t)3 * g_next;
            n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
        }
        if (i - kBatch >= start) g_next2 = __ldg(payload + i - kBatch);
    }
    for (int top = last; top > start; top -= kBatch) {
        const int i = top - 1 - lane;
        const int g = g_next;
        const float4 r0 = n0, r1 = n1, r2 = n2;
        {
            const int in = i - kBatch;
            g_next = g_next2;
            if (in >= start) {
                const float4 *src = records + (size_t)3 * g_next;
                n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
            }
            if (in - kBatch >= start) g_next2 = __ldg(payload + in - kBatch);
        }
        const bool valid = i >= start;
        const float dx = fx - r0.x, dy = fy - r0.y;
        const float opacity = r1.z;
        const float cg = fmaf(r2.y, g2, fmaf(r2.x, g1, r1.w * g0));
        // the reference: utils.py:331-345 in literal fp32, decision R.py:634 (false for NaN)
        lit.prod = literal_alpha(dx, dy, r0.z, r0.w, r1.x, r1.y, r1.z);
        lit.v0 = fmaf(r0.w, dy, r0.z * dx);
        lit.v1 = fmaf(r1.x, dy, r0.w * dx);
        lit.take = valid && (lit.prod >= kAlphaMin);
        lit.a = lit.take ? fminf(lit.prod, 0.99f) : 0.0f;
        lit.inv = lit.take ? 1.0f / (1.0f - lit.a) : 1.0f;
        fix_scan(lit, cg, lane);
        float out[11];
#pragma unroll
        for (int q = 0; q < 11; ++q) out[q] = 0.f;
        float gm0, gm1;
        fix_terms<kMagnitude>(lit, 1.0f, cg, opacity, g0, g1, g2, out, gm0, gm1);
        if (kMagnitude) {
            tot0 += fabsf(gm0);
            tot1 += fabsf(gm1);
        }
        if (replay) {  // warp-uniform: what the hot kernel added for this pixel, with the opposite sign
            const BandConsts band = band_consts(r1.y * r1.z, r2.w);
            float xp;
            fast.prod = fast_alpha(dx, dy, r0.z * dx, r0.w * dx, r0.w, r1.x, r2.w, band.hi, fast.v0, fast.v1, xp);
            fast.take = valid && !stopped && (xp >= band.xlo);
            // the hot kernel took the first ambiguous pair (x' < 0) of the pixel and nothing after it
            const unsigned amb = __ballot_sync(0xffffffffu, fast.take && xp < 0.0f);
            if (amb) {
                const int first = __ffs(amb) - 1;  // lanes are in traversal order
                fast.take = fast.take && lane <= first;
                stopped = true;
            }
            fast.a = fast.take ? fminf(fast.prod, 0.99f) : 0.0f;
            fast.inv = fast.take ? fast_rcp(1.0f - fast.a) : 1.0f;
            fix_scan(fast, cg, lane);
            fix_terms<kMagnitude>(fast, -1.0f, cg, opacity, g0, g1, g2, out, gm0, gm1);
        }
        if (lit.take || (replay && fast.take)) {
            float *dst = splat_grads + (size_t)kGradSlots * g;
#pragma unroll
            for (int q = 0; q < 11; ++q)
                if ((kMagnitude || q != 9) && out[q] != 0.0f) atomicAdd(dst + q, out[q]);
        }
    }
}

template <bool kMagnitude>
__global__ void __launch_bounds__(128)
blend_bwd_fixup_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                       const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                       const float4 *__restrict__ records, const float *__restrict__ image_grad,
                       const float *__restrict__ acc_alpha, const int32_t *__restrict__ last_effective,
                       float *__restrict__ splat_grads, float *__restrict__ map_on_image,
                       const unsigned *__restrict__ fix_list)
{
    const int lane = threadIdx.x & 31;
    const unsigned count = fix_list[0];
    const unsigned warps = gridDim.x * (blockDim.x >> 5);
    const int tiles_per_row = width / kTile;
    // the flagged pixels are dealt out one per warp, round robin (the list is in no particular order)
    for (unsigned idx = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); idx < count; idx += warps) {
        const unsigned entry = fix_list[1u + idx];
        const bool whole = (entry & kFixWholeBit) != 0u;
        const long long px = (long long)(entry & ~kFixWholeBit);
        const int pv = (int)(px / width), pu = (int)(px % width);
        const int tile = (pv / kTile) * tiles_per_row + pu / kTile;
        const int start = tile_start[tile], end = tile_end[tile];
        const int last = (end > start) ? min(last_effective[px], end) : start;
        const float T_end = 1.0f - acc_alpha[px];
        const float g0 = image_grad[3 * px], g1 = image_grad[3 * px + 1], g2 = image_grad[3 * px + 2];
        float tot0 = 0.f, tot1 = 0.f;
        fixup_pixel<kMagnitude>(lane, !whole, start, last, (float)pu + 0.5f, (float)pv + 0.5f, T_end, g0, g1, g2,
                                payload, records, splat_grads, tot0, tot1);
        if (kMagnitude) {
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                tot0 += __shfl_xor_sync(0xffffffffu, tot0, d);
                tot1 += __shfl_xor_sync(0xffffffffu, tot1, d);
            }
            if (lane == 0) reinterpret_cast<float2 *>(map_on_image)[px] = make_float2(tot0, tot1);
        }
    }
}

struct BwdLaunchInfo {
    int resident = 0;  // CTAs of the hot kernel resident at once on this device
    int sms = 0;
};

template <int kCells>
int launch_bwd(int tiles, cudaStream_t st, int height, int width, const int32_t *tile_start,
               const int32_t *tile_end, const int32_t *payload, const float4 *rec,
               const float *image_grad, const float *acc_alpha, const int32_t *last_effective,
               float *splat_grads, float *map_on_image, const uint8_t *cell_masks, long long mask_stride,
               const uint8_t *quarter_flags, unsigned *fix_list)
{
    constexpr int kParts = kTilePixels / (kCells * 32);
    constexpr int kThreads = 32 * kParts;
    // persistent launch: as many CTAs as are resident at once, warps pull (tile, part) units
    static BwdLaunchInfo info[64][2];
    int dev = 0;
    GS_CUDA_TRY(cudaGetDevice(&dev));
    BwdLaunchInfo &li = info[dev & 63][map_on_image ? 1 : 0];
    if (li.resident == 0) {
        int per_sm = 0;
        GS_CUDA_TRY(cudaDeviceGetAttribute(&li.sms, cudaDevAttrMultiProcessorCount, dev));
        if (map_on_image)
            GS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, blend_bwd_kernel<kCells, true>, kThreads, 0));
        else
            GS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, blend_bwd_kernel<kCells, false>, kThreads, 0));
        // GS_BWD_CTAS_PER_SM: run the persistent grid with fewer CTAs per SM than fit (read once per
        // device): registers and warp slots left over let another stream's small kernels (projection,
        // sorts, binning of the next view) run beside this one instead of behind it
        const char *ce = getenv("GS_BWD_CTAS_PER_SM");
        if (ce && atoi(ce) > 0 && atoi(ce) < per_sm) per_sm = atoi(ce);
        li.resident = li.sms * (per_sm > 0 ? per_sm : 1);
    }
    GS_CUDA_TRY(cudaMemsetAsync(fix_list, 0, sizeof(unsigned), st));  // the list's counter
    const char *pe = getenv("GS_BLEND_PERSISTENT");  // 0: one CTA per tile (A/B and tests; read per call)
    GsWorkCounter *work = (pe && atoi(pe) == 0) ? nullptr : gs_work_counter(st);
    const unsigned units = (unsigned)tiles * kParts;
    const int grid = work ? (tiles < li.resident ? tiles : li.resident) : tiles;
    // the literal pass: a fixed grid (the list's length is known on the device only), one flagged pixel
    // per warp and round; with an empty list -- the common case away from thresholds -- it costs a launch
    const long long fix_ctas = 4ll * li.sms;
    const char *fe = getenv("GS_BWD_FIXUP");  // 0: timing experiments only (results are then NOT exact)
    const bool fixup = !(fe && atoi(fe) == 0);
    {
        // tests only: behave like a persistent launch that died half way (its counter pair left at
        // arbitrary values, the call reporting an error)
        const char *te = getenv("GS_TEST_POISON_WORK_COUNTER");
        if (te && atoi(te) == 1 && work) {
            GS_CUDA_TRY(cudaMemsetAsync(work, 0x7f, sizeof(GsWorkCounter), st));
            gs_work_counter_failed(work);
            return GS_E_INVALID;
        }
    }
    GS_CUDA_TRY(gs_work_counter_arm(work, st));
#define GS_BWD_LAUNCH(MAG)                                                                              \
    do {                                                                                                \
        blend_bwd_kernel<kCells, MAG><<<grid, kThreads, 0, st>>>(                                       \
            height, width, tile_start, tile_end, payload, rec, image_grad, acc_alpha, last_effective,   \
            splat_grads, map_on_image, cell_masks, mask_stride, quarter_flags, fix_list, work, units);  \
        {                                                                                               \
            const cudaError_t e__ = cudaGetLastError();                                                 \
            if (e__ != cudaSuccess) {                                                                   \
                gs_work_counter_failed(work); /* its pair may be half way: the next launch re-arms it */ \
                return (int)e__;                                                                        \
            }                                                                                           \
        }                                                                                               \
        if (fixup) blend_bwd_fixup_kernel<MAG><<<(unsigned)fix_ctas, 128, 0, st>>>(                     \
            height, width, tile_start, tile_end, payload, rec, image_grad, acc_alpha, last_effective,   \
            splat_grads, map_on_image, fix_list);                                                       \
    } while (0)
    if (map_on_image) GS_BWD_LAUNCH(true); else GS_BWD_LAUNCH(false);
#undef GS_BWD_LAUNCH
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace

extern "C" int64_t gs_blend_bwd_scratch_bytes(int height, int width)
{
    // {count, one entry per pixel at most}: every pixel is appended at most once per call
    return (height > 0 && width > 0) ? 4 * ((int64_t)height * width + 1) : 0;
}

extern "C" int gs_blend_bwd(int height, int width, const int32_t *tile_start,
                            const int32_t *tile_end, const int32_t *sorted_payload,
                            const float *records, const float *image_grad, const float *acc_alpha,
                            const int32_t *last_effective, float *splat_grads, float *map_on_image,
                            const uint8_t *cell_masks, int64_t mask_stride, const uint8_t *quarter_flags,
                            void *scratch, void *stream)
{
    if (mask_stride < 0) return GS_E_INVALID;
    if (height <= 0 || width <= 0 || height % kTile || width % kTile) return GS_E_INVALID;
    if (!tile_start || !tile_end || !image_grad || !acc_alpha || !last_effective || !splat_grads)
        return GS_E_INVALID;
    // the culling bits and the irregular flags of the forward blend of the SAME lists are required: the
    // backward never redoes the culling arithmetic
    if (!cell_masks || !quarter_flags || !scratch) return GS_E_INVALID;
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
                             mask_stride, quarter_flags, (unsigned *)scratch);
    return launch_bwd<4>(tiles, st, height, width, tile_start, tile_end, sorted_payload, rec,
                         image_grad, acc_alpha, last_effective, splat_grads, map_on_image, cell_masks,
                         mask_stride, quarter_flags, (unsigned *)scratch);
}
