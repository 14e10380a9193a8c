// blend_fwd.cu -- per-tile front-to-back alpha blend (reference R.py:318-485, utils.py:275-284).
//
// Literal-fp32 profile: build with -fmad=false; the per-pair arithmetic follows
// gso_blend_fwd in oracle/gs_oracle.c operation for operation, so image, depth, accumulated
// alpha, last-effective index and per-pixel counts are bit-exact with the oracle.
//
// The kernel is instruction-issue bound (ncu: issue slots > 90 % busy, DRAM < 10 %), so the design
// minimises executed warp instructions rather than bytes:
//
//   * one WARP owns a column of kCells cells of 8x4 pixels (kCells = 4: an 8x16 half tile,
//     kCells = 2: an 8x8 quarter tile); lane l owns pixel (l % 8, l / 8) of every cell;
//   * the tile's sorted list is walked in batches of 32: lane l stages entry l (48-byte record,
//     three 16-byte loads, prefetched one batch ahead into registers) and computes a kCells-bit
//     mask -- bit j is set only if the splat can reach alpha >= 1/255 on some pixel centre of
//     cell j (exact minimum of the conic's quadratic form over every pixel row of the cell, with
//     a safety margin far above either profile's rounding).  About half of a tile's entries
//     reach no pixel of the tile at all (bounding-box corners, low opacity), and of the rest
//     most touch one or two cells;
//   * the warp then visits only the (entry, cell) combinations whose bit is set and whose cell
//     still has an unsaturated pixel -- the compositing arithmetic is skipped exactly where the
//     reference's own `alpha < 1/255` test would have discarded it, so results do not change;
//   * a cell retires when all its 32 pixels are saturated (one REDUX per entry), the warp stops when all its
//     cells have retired; the reference keeps looping without effect (R.py:387-394).
//
// Warps of one tile share nothing and never synchronise with each other.
#include <cstdlib>

#include "gs_common.cuh"
#include "gs_f32x2.cuh"

namespace {

constexpr int kBatch = 32;
constexpr int kPairH = 2 * kCellH;  // a pair cell: 8 x 8 pixels = two vertically stacked 8x4 cells

// exp of the literal-fp32 profile on two values at once; each half is operation for operation
// gs_expf_core (gs_common.cuh), with FMUL2 / FADD2 / FFMA2 in place of the scalar instructions
__device__ __forceinline__ f32x2 gs_expf_core2(f32x2 x)
{
    float t_lo, t_hi;
    unpack2(mul2(x, splat2(1.44269504088896341f)), t_lo, t_hi);
    // scalar adds on purpose: ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2
    // (it does not contract the scalar forms in a -fmad=false unit), which would change the bits
    const f32x2 tj = pack2(t_lo + 12582912.0f, t_hi + 12582912.0f);
    const f32x2 j = sub2(tj, splat2(12582912.0f));
    f32x2 f = fma2(j, splat2(-0.693145751953125f), x);
    f = fma2(j, splat2(-1.42860682030941723212e-6f), f);
    f32x2 p = splat2(1.9875691500e-4f);
    p = fma2(p, f, splat2(1.3981999507e-3f));
    p = fma2(p, f, splat2(8.3334519073e-3f));
    p = fma2(p, f, splat2(4.1665795894e-2f));
    p = fma2(p, f, splat2(1.6666665459e-1f));
    p = fma2(p, f, splat2(5.0000001201e-1f));
    const f32x2 f2 = mul2(f, f);
    const f32x2 r = add2(fma2(p, f2, f), splat2(1.0f));
    float rl, rh, tl, th;
    unpack2(r, rl, rh);
    unpack2(tj, tl, th);
    return pack2(__int_as_float(__float_as_int(rl) + (__float_as_int(tl) << 23)),
                 __int_as_float(__float_as_int(rh) + (__float_as_int(th) << 23)));
}

// Per-pixel compositing state of one lane: TWO pixels (4 rows apart, index 0 = lower, 1 = upper) in
// each of the warp's kPairs pair cells.  The products are formed two at a time (FMUL2); the sums
// they feed are scalar adds, because ptxas would otherwise contract mul2 + add2 into an FFMA2.
template <int kPairs>
struct PixelState {
    f32x2 T[kPairs];
    float C0[kPairs][2], C1[kPairs][2], C2[kPairs][2], D[kPairs][2], Wd[kPairs][2];
    int last[kPairs][2], cnt[kPairs][2];
    unsigned live;  // bit 2c / 2c+1: the lower / upper pixel of pair cell c is not saturated yet
};

// One (entry, pixel) step exactly as the reference writes it (R.py:437-472); used for entries with
// an irregular conic or non-finite fields, where the full gs_expf range handling matters and a
// masked "add zero" would not be a no-op.
template <bool kRgbOnly>
__device__ __forceinline__ void visit_scalar(float &T, float &C0, float &C1, float &C2, float &D,
                                             float &Wd, int &last, int &cnt, bool &live, const float dx,
                                             const float dxdxa, const float dy, const float4 &r0,
                                             const float4 &r1, const float4 &r2, const int idx1)
{
    const float e = -0.5f * (dxdxa + dy * dy * r1.x) - dx * dy * r0.w;
    float a = gs_expf(e) * r1.y * r1.z;
    if (!live || a < 1.0f / 255.0f) return;
    a = fminf(a, 0.99f);
    const float nT = T * (1.0f - a);
    if (nT < 0.0001f) {
        live = false;
        return;
    }
    last = idx1;
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

template <int kPairs, bool kRgbOnly>
__device__ __forceinline__ void visit_irregular(PixelState<kPairs> &st, const int c, const float dx,
                                             const float dxdxa, const float fy_lo, const float4 r0,
                                             const float4 r1, const float4 r2, const int idx1)
{
#pragma unroll
    for (int cc = 0; cc < kPairs; ++cc) {
        if (cc != c) continue;
        float T0, T1;
        unpack2(st.T[cc], T0, T1);
        bool l0 = (st.live >> (2 * cc)) & 1u, l1 = (st.live >> (2 * cc + 1)) & 1u;
        visit_scalar<kRgbOnly>(T0, st.C0[cc][0], st.C1[cc][0], st.C2[cc][0], st.D[cc][0], st.Wd[cc][0],
                               st.last[cc][0], st.cnt[cc][0], l0, dx, dxdxa, fy_lo - r0.y, r0, r1, r2, idx1);
        visit_scalar<kRgbOnly>(T1, st.C0[cc][1], st.C1[cc][1], st.C2[cc][1], st.D[cc][1], st.Wd[cc][1],
                               st.last[cc][1], st.cnt[cc][1], l1, dx, dxdxa, (fy_lo + (float)kCellH) - r0.y,
                               r0, r1, r2, idx1);
        st.T[cc] = pack2(T0, T1);
        st.live = (st.live & ~(3u << (2 * cc))) | ((l0 ? 1u : 0u) << (2 * cc)) | ((l1 ? 1u : 0u) << (2 * cc + 1));
    }
}

// One (entry, pair cell) visit: both pixels of the lane at once.  A pixel that does not take the
// entry (alpha < 1/255, already saturated, or saturating now) adds a literal zero, which leaves its
// sums bit-identical; every other operation is the reference's, in its order.
template <int kPairs, bool kRgbOnly>
__device__ __forceinline__ void visit_pair(PixelState<kPairs> &st, const int c, const float dx,
                                           const float dxdxa, const f32x2 fy, const float4 &r0,
                                           const float4 &r1, const float4 &r2, const int idx1)
{
    const f32x2 dy = sub2(fy, splat2(r0.y));
    float ql, qh;
    unpack2(mul2(mul2(dy, dy), splat2(r1.x)), ql, qh);
    // (scalar adds, see gs_expf_core2; the scaling by -0.5 is exact, so its contraction with the
    // subtraction that follows cannot change the result)
    const f32x2 t1 = mul2(splat2(-0.5f), pack2(dxdxa + ql, dxdxa + qh));
    const f32x2 t2 = mul2(mul2(splat2(dx), dy), splat2(r0.w));
    float el, eh;
    unpack2(sub2(t1, t2), el, eh);
    // e < -87 gives alpha < 1/255 with either exp, so the clamp does not change any decision
    const f32x2 ex = gs_expf_core2(pack2(fmaxf(el, -87.0f), fmaxf(eh, -87.0f)));
    float al, ah;
    unpack2(mul2(mul2(ex, splat2(r1.y)), splat2(r1.z)), al, ah);
    const bool pl = ((st.live >> (2 * c)) & 1u) && !(al < 1.0f / 255.0f);
    const bool ph = ((st.live >> (2 * c + 1)) & 1u) && !(ah < 1.0f / 255.0f);
    al = fminf(al, 0.99f);
    ah = fminf(ah, 0.99f);
    const f32x2 T = st.T[c];
    float nl, nh;
    unpack2(mul2(T, sub2(splat2(1.0f), pack2(al, ah))), nl, nh);
    const bool sl = pl && (nl < 0.0001f), sh = ph && (nh < 0.0001f);
    const bool ul = pl && !sl, uh = ph && !sh;
    if (sl) st.live &= ~(1u << (2 * c));
    if (sh) st.live &= ~(2u << (2 * c));
    const f32x2 a = pack2(ul ? al : 0.0f, uh ? ah : 0.0f);
    float w0, w1;
    unpack2(mul2(mul2(splat2(r1.w), a), T), w0, w1);
    st.C0[c][0] += w0;
    st.C0[c][1] += w1;
    unpack2(mul2(mul2(splat2(r2.x), a), T), w0, w1);
    st.C1[c][0] += w0;
    st.C1[c][1] += w1;
    unpack2(mul2(mul2(splat2(r2.y), a), T), w0, w1);
    st.C2[c][0] += w0;
    st.C2[c][1] += w1;
    if (!kRgbOnly) {
        unpack2(mul2(mul2(splat2(r2.z), a), T), w0, w1);
        st.D[c][0] += w0;
        st.D[c][1] += w1;
        unpack2(mul2(a, T), w0, w1);
        st.Wd[c][0] += w0;
        st.Wd[c][1] += w1;
        st.cnt[c][0] += ul ? 1 : 0;
        st.cnt[c][1] += uh ? 1 : 0;
    }
    if (ul) st.last[c][0] = idx1;
    if (uh) st.last[c][1] = idx1;
    float tl, th;
    unpack2(T, tl, th);
    st.T[c] = pack2(ul ? nl : tl, uh ? nh : th);
}

template <int kPairs, bool kRgbOnly>
__global__ void __launch_bounds__(32 * (kTilePixels / (kPairs * 64)))
blend_fwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, float *__restrict__ image,
                 float *__restrict__ depth, float *__restrict__ acc_alpha,
                 int32_t *__restrict__ last_effective, int32_t *__restrict__ valid_count,
                 uint8_t *__restrict__ cell_masks, const long long mask_stride,
                 uint8_t *__restrict__ quarter_flags)
{
    constexpr int kParts = kTilePixels / (kPairs * 64);  // warps per tile
    constexpr int kPartsPerColumn = kParts / 2;
    constexpr unsigned kAllPairs = (1u << kPairs) - 1u;
    __shared__ float4 s_rec_all[kParts][kBatch * 3];

    const int lane = threadIdx.x & 31, part = threadIdx.x >> 5;
    float4 *s_rec = s_rec_all[part];
    const int tiles_per_row = width / kTile;
    const int tile = blockIdx.x;
    const int tile_u = tile % tiles_per_row, tile_v = tile / tiles_per_row;
    const int col0 = tile_u * kTile + (part / kPartsPerColumn) * kCellW;          // first pixel column
    const int row0 = tile_v * kTile + (part % kPartsPerColumn) * (kPairs * kPairH);  // first pixel row
    const int pu = col0 + (lane & 7);
    const int pv0 = row0 + (lane >> 3);  // lower pixel of pair cell 0; upper = +4; pair cell c = +8c
    const float fx = (float)pu + 0.5f, fy0 = (float)pv0 + 0.5f;
    const float x_first = (float)col0 + 0.5f, y_first = (float)row0 + 0.5f;
    const int start = tile_start[tile], end = tile_end[tile];

    PixelState<kPairs> st;
#pragma unroll
    for (int c = 0; c < kPairs; ++c) {
        st.T[c] = splat2(1.0f);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            st.C0[c][h] = st.C1[c][h] = st.C2[c][h] = st.D[c][h] = st.Wd[c][h] = 0.0f;
            st.last[c][h] = start;
            st.cnt[c][h] = 0;
        }
    }
    st.live = (1u << (2 * kPairs)) - 1u;
    unsigned open = kAllPairs;  // pair cells that still have an unsaturated pixel (warp-uniform)
    unsigned seen_irregular = 0u;  // some entry this warp had to look at is irregular (warp-uniform)

    // software pipeline: records of the next batch in registers, payload two batches ahead
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0;
    int g_next = 0;
    {
        const int i = start + lane;
        if (i < end) {
            const float4 *src = records + (size_t)3 * __ldg(payload + i);
            n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
        }
        if (i + kBatch < end) g_next = __ldg(payload + i + kBatch);
    }

    for (int base = start; base < end && open; base += kBatch) {
        // ---- stage this batch (registers -> shared) and cull it against the open pair cells -----
        unsigned my_mask = 0u;
        if (base + lane < end) {
            // culling bits per 8x4 cell (two per pair cell); a pair cell is visited if either is set
            const unsigned fine = gs_entry_mask<2 * kPairs>(n0, n1, n2, x_first, y_first);
            unsigned raw = fine & kCellMaskIrregular;
#pragma unroll
            for (int c = 0; c < kPairs; ++c) raw |= ((fine >> (2 * c)) & 3u) ? (1u << c) : 0u;
            // hand the bits to the backward blend (same cells, same entries): plane q of the tile's
            // four 8x8 quarters, q = 2 * column + row half, one byte per list entry
            if (cell_masks) {
#pragma unroll
                for (int c = 0; c < kPairs; ++c) {
                    const int q = 2 * (part / kPartsPerColumn) + (part % kPartsPerColumn) * kPairs + c;
                    cell_masks[q * mask_stride + base + lane] = (uint8_t)((fine >> (2 * c)) & 3u);
                }
            }
            my_mask = (raw & open) ? (raw & (open | kCellMaskIrregular)) : 0u;
        }
        s_rec[3 * lane] = n0;
        s_rec[3 * lane + 1] = n1;
        s_rec[3 * lane + 2] = n2;
        // ---- prefetch the next batch ------------------------------------------------------------
        {
            const int i = base + kBatch + lane;
            if (i < end) {
                const float4 *src = records + (size_t)3 * g_next;
                n0 = __ldg(src); n1 = __ldg(src + 1); n2 = __ldg(src + 2);
            }
            if (i + kBatch < end) g_next = __ldg(payload + i + kBatch);
        }
        __syncwarp();

        unsigned pending = __ballot_sync(0xffffffffu, my_mask != 0u);
        const unsigned irregular = __ballot_sync(0xffffffffu, (my_mask & kCellMaskIrregular) != 0u);
        seen_irregular |= irregular;
        while (pending) {
            const int k = __ffs(pending) - 1;
            pending &= pending - 1u;
            unsigned m = 1u;  // one pair cell per warp: a pending entry is an entry of that cell
            if (kPairs > 1) {
                m = __shfl_sync(0xffffffffu, my_mask, k) & open;
                if (m == 0u) continue;  // warp-uniform
            }
            const float4 r0 = s_rec[3 * k];      // u v a b
            const float4 r1 = s_rec[3 * k + 1];  // c rescale alpha red
            const float4 r2 = s_rec[3 * k + 2];  // green blue z pad
            const float dx = fx - r0.x;
            const float dxdxa = dx * dx * r0.z;
            const int idx1 = base + k + 1;
            if ((irregular >> k) & 1u) {  // warp-uniform, rare
#pragma unroll
                for (int c = 0; c < kPairs; ++c)
                    if ((m >> c) & 1u)
                        visit_irregular<kPairs, kRgbOnly>(st, c, dx, dxdxa, fy0 + (float)(kPairH * c), r0, r1, r2, idx1);
            } else {
#pragma unroll
                for (int c = 0; c < kPairs; ++c) {
                    if (!((m >> c) & 1u)) continue;  // warp-uniform
                    const float fy = fy0 + (float)(kPairH * c);
                    visit_pair<kPairs, kRgbOnly>(st, c, dx, dxdxa, pack2(fy, fy + (float)kCellH), r0, r1, r2, idx1);
                }
            }
            const unsigned ob = __reduce_or_sync(0xffffffffu, st.live);  // a pair cell retires with its last live pixel
            open = 0u;
#pragma unroll
            for (int c = 0; c < kPairs; ++c) open |= ((ob >> (2 * c)) & 3u) ? (1u << c) : 0u;
            if (open == 0u) break;
        }
        __syncwarp();
    }

    // tells the backward blend that this 8x8 quarter's list holds an irregular entry among those its
    // pixels reach: the backward then leaves the quarter to its literal (guarded) pass
    if (quarter_flags && lane == 0) {
#pragma unroll
        for (int c = 0; c < kPairs; ++c)
            quarter_flags[4 * tile + 2 * (part / kPartsPerColumn) + (part % kPartsPerColumn) * kPairs + c] =
                seen_irregular ? 1 : 0;
    }
#pragma unroll
    for (int c = 0; c < kPairs; ++c) {
        float T[2];
        unpack2(st.T[c], T[0], T[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const size_t px = (size_t)(pv0 + kPairH * c + kCellH * h) * width + pu;
            image[3 * px] = st.C0[c][h];
            image[3 * px + 1] = st.C1[c][h];
            image[3 * px + 2] = st.C2[c][h];
            if (!kRgbOnly) {
                depth[px] = st.D[c][h] / fmaxf(st.Wd[c][h], 1e-6f);
                acc_alpha[px] = 1.0f - T[h];
                last_effective[px] = st.last[c][h];
                valid_count[px] = st.cnt[c][h];
            }
        }
    }
}

template <int kPairs>
int launch_fwd(int tiles, cudaStream_t st, int rgb_only, int height, int width,
               const int32_t *tile_start, const int32_t *tile_end, const int32_t *payload,
               const float4 *rec, float *image, float *depth, float *acc_alpha,
               int32_t *last_effective, int32_t *valid_count, uint8_t *cell_masks, long long mask_stride,
               uint8_t *quarter_flags)
{
    constexpr int kThreads = 32 * (kTilePixels / (kPairs * 64));
    if (rgb_only)
        blend_fwd_kernel<kPairs, true><<<tiles, kThreads, 0, st>>>(
            height, width, tile_start, tile_end, payload, rec, image, depth, acc_alpha,
            last_effective, valid_count, cell_masks, mask_stride, quarter_flags);
    else
        blend_fwd_kernel<kPairs, false><<<tiles, kThreads, 0, st>>>(
            height, width, tile_start, tile_end, payload, rec, image, depth, acc_alpha,
            last_effective, valid_count, cell_masks, mask_stride, quarter_flags);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

}  // namespace

extern "C" int gs_blend_fwd(int height, int width, const int32_t *tile_start,
                            const int32_t *tile_end, const int32_t *sorted_payload,
                            const float *records, float *image, float *depth, float *acc_alpha,
                            int32_t *last_effective, int32_t *valid_count, int rgb_only,
                            uint8_t *cell_masks, int64_t mask_stride, uint8_t *quarter_flags, void *stream)
{
    if (cell_masks && mask_stride < 0) return GS_E_INVALID;
    if ((cell_masks == nullptr) != (quarter_flags == nullptr)) return GS_E_INVALID;  // both or neither
    if (height <= 0 || width <= 0 || height % kTile || width % kTile) return GS_E_INVALID;
    if (!tile_start || !tile_end || !image) return GS_E_INVALID;
    if (!rgb_only && (!depth || !acc_alpha || !last_effective || !valid_count)) return GS_E_INVALID;
    int tiles = (width / kTile) * (height / kTile);
    static int pairs = 0;
    if (pairs == 0) {
        const char *e = getenv("GS_FWD_PAIRS");  // tuning knob: 8x8 pair cells per warp (1 or 2)
        pairs = e ? atoi(e) : 1;
        if (pairs != 1 && pairs != 2) pairs = 1;
    }
    const float4 *rec = reinterpret_cast<const float4 *>(records);
    cudaStream_t st = (cudaStream_t)stream;
    if (pairs == 1)
        return launch_fwd<1>(tiles, st, rgb_only, height, width, tile_start, tile_end, sorted_payload, rec,
                             image, depth, acc_alpha, last_effective, valid_count, cell_masks, mask_stride,
                             quarter_flags);
    return launch_fwd<2>(tiles, st, rgb_only, height, width, tile_start, tile_end, sorted_payload, rec,
                         image, depth, acc_alpha, last_effective, valid_count, cell_masks, mask_stride,
                         quarter_flags);
}
