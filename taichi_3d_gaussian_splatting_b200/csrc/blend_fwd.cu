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

namespace {

constexpr int kBatch = 32;

// Per-pixel compositing state of one lane: one pixel in each of the warp's kCells cells.
template <int kCells>
struct PixelState {
    float T[kCells], C0[kCells], C1[kCells], C2[kCells], D[kCells], Wd[kCells];
    int last[kCells], cnt[kCells];
    unsigned live;  // bit j: the pixel of cell j is not saturated yet
};

// One (entry, cell) visit for the lane's pixel of cell j.
// kGuard: the conic is irregular (not positive definite / finite), so the exponent may exceed 88
// or be NaN and the full gs_expf range handling is needed; regular entries have e <= ~0.
template <int kCells, bool kRgbOnly, bool kGuard>
__device__ __forceinline__ void visit(PixelState<kCells> &st, const int j, const float dx,
                                      const float dxdxa, const float fy, const float4 &r0,
                                      const float4 &r1, const float4 &r2, const int idx1)
{
    const float dy = fy - r0.y;
    const float e = -0.5f * (dxdxa + dy * dy * r1.x) - dx * dy * r0.w;
    // e < -87 gives alpha < 1/255 with either exp, so the clamp does not change any decision
    float ex = gs_expf_core(fmaxf(e, -87.0f));
    if (kGuard) {
        if (!(e <= 88.0f)) ex = (e != e) ? e : __int_as_float(0x7f800000);  // as gs_expf
    }
    float a = ex * r1.y * r1.z;
    const float T = st.T[j];
    const bool pass = ((st.live >> j) & 1u) && !(a < 1.0f / 255.0f);
    a = fminf(a, 0.99f);
    const float nT = T * (1.0f - a);
    const bool sat = pass && (nT < 0.0001f);
    if (sat) st.live &= ~(1u << j);
    if (pass && !sat) {
        st.last[j] = idx1;
        st.C0[j] += r1.w * a * T;
        st.C1[j] += r2.x * a * T;
        st.C2[j] += r2.y * a * T;
        if (!kRgbOnly) {
            st.D[j] += r2.z * a * T;
            st.Wd[j] += a * T;
            st.cnt[j] += 1;
        }
        st.T[j] = nT;
    }
}

template <int kCells, bool kRgbOnly, bool kSplit>
__global__ void __launch_bounds__(32 * (kTilePixels / (kCells * 32)))
blend_fwd_kernel(int height, int width, const int32_t *__restrict__ tile_start,
                 const int32_t *__restrict__ tile_end, const int32_t *__restrict__ payload,
                 const float4 *__restrict__ records, float *__restrict__ image,
                 float *__restrict__ depth, float *__restrict__ acc_alpha,
                 int32_t *__restrict__ last_effective, int32_t *__restrict__ valid_count)
{
    constexpr int kParts = kTilePixels / (kCells * 32);  // warps per tile
    constexpr int kPartsPerColumn = kParts / 2;
    constexpr unsigned kAll = (1u << kCells) - 1u;
    __shared__ float4 s_rec_all[kParts][kBatch * 3];

    const int lane = threadIdx.x & 31, part = threadIdx.x >> 5;
    float4 *s_rec = s_rec_all[part];
    const int tiles_per_row = width / kTile;
    const int tile = blockIdx.x;
    const int tile_u = tile % tiles_per_row, tile_v = tile / tiles_per_row;
    const int col0 = tile_u * kTile + (part / kPartsPerColumn) * kCellW;          // first pixel column
    const int row0 = tile_v * kTile + (part % kPartsPerColumn) * (kCells * kCellH);  // first pixel row
    const int pu = col0 + (lane & 7);
    const int pv0 = row0 + (lane >> 3);
    const float fx = (float)pu + 0.5f, fy0 = (float)pv0 + 0.5f;
    const float x_first = (float)col0 + 0.5f, y_first = (float)row0 + 0.5f;
    const int start = tile_start[tile], end = tile_end[tile];

    PixelState<kCells> st;
#pragma unroll
    for (int j = 0; j < kCells; ++j) {
        st.T[j] = 1.0f;
        st.C0[j] = st.C1[j] = st.C2[j] = st.D[j] = st.Wd[j] = 0.f;
        st.last[j] = start;
        st.cnt[j] = 0;
    }
    st.live = kAll;
    unsigned open = kAll;  // cells that still have an unsaturated pixel (warp-uniform)

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
        // ---- stage this batch (registers -> shared) and cull it against the open cells --------
        unsigned my_mask = 0u;
        if (base + lane < end) {
            const unsigned raw = gs_cell_mask<kCells>(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y * n1.z, x_first, y_first);
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
        while (pending) {
            const int k = __ffs(pending) - 1;
            pending &= pending - 1u;
            const unsigned mk = __shfl_sync(0xffffffffu, my_mask, k);
            const unsigned m = mk & open;
            if (m == 0u) continue;  // warp-uniform
            const float4 r0 = s_rec[3 * k];      // u v a b
            const float4 r1 = s_rec[3 * k + 1];  // c rescale alpha red
            const float4 r2 = s_rec[3 * k + 2];  // green blue z pad
            const float dx = fx - r0.x;
            const float dxdxa = dx * dx * r0.z;
            const int idx1 = base + k + 1;
            if (!kSplit || (mk & kCellMaskIrregular)) {  // warp-uniform, rare: guarded exp
#pragma unroll
                for (int j = 0; j < kCells; ++j) {
                    if (!((m >> j) & 1u)) continue;
                    visit<kCells, kRgbOnly, true>(st, j, dx, dxdxa, fy0 + (float)(kCellH * j), r0, r1, r2, idx1);
                }
            } else {
#pragma unroll
                for (int j = 0; j < kCells; ++j) {
                    if (!((m >> j) & 1u)) continue;  // warp-uniform
                    visit<kCells, kRgbOnly, false>(st, j, dx, dxdxa, fy0 + (float)(kCellH * j), r0, r1, r2, idx1);
                }
            }
            open = __reduce_or_sync(0xffffffffu, st.live);  // a cell retires with its last live pixel
            if (open == 0u) break;
        }
        __syncwarp();
    }

#pragma unroll
    for (int j = 0; j < kCells; ++j) {
        const size_t px = (size_t)(pv0 + kCellH * j) * width + pu;
        image[3 * px] = st.C0[j];
        image[3 * px + 1] = st.C1[j];
        image[3 * px + 2] = st.C2[j];
        if (!kRgbOnly) {
            depth[px] = st.D[j] / fmaxf(st.Wd[j], 1e-6f);
            acc_alpha[px] = 1.0f - st.T[j];
            last_effective[px] = st.last[j];
            valid_count[px] = st.cnt[j];
        }
    }
}

template <int kCells, bool kSplit>
int launch_fwd(int tiles, cudaStream_t st, int rgb_only, int height, int width,
               const int32_t *tile_start, const int32_t *tile_end, const int32_t *payload,
               const float4 *rec, float *image, float *depth, float *acc_alpha,
               int32_t *last_effective, int32_t *valid_count)
{
    constexpr int kThreads = 32 * (kTilePixels / (kCells * 32));
    if (rgb_only)
        blend_fwd_kernel<kCells, true, kSplit><<<tiles, kThreads, 0, st>>>(
            height, width, tile_start, tile_end, payload, rec, image, depth, acc_alpha,
            last_effective, valid_count);
    else
        blend_fwd_kernel<kCells, false, kSplit><<<tiles, kThreads, 0, st>>>(
            height, width, tile_start, tile_end, payload, rec, image, depth, acc_alpha,
            last_effective, valid_count);
    GS_LAUNCH_CHECK();
    return GS_OK;
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
    static int cells = 0, split = 1;
    if (cells == 0) {
        const char *e = getenv("GS_FWD_CELLS");  // tuning knob: 8x4 cells per warp (2 or 4)
        cells = e ? atoi(e) : 4;
        if (cells != 2 && cells != 4) cells = 4;
        const char *v = getenv("GS_FWD_SPLIT");  // tuning knob: separate loop for irregular conics
        if (v) split = atoi(v) != 0;
    }
    const float4 *rec = reinterpret_cast<const float4 *>(records);
    cudaStream_t st = (cudaStream_t)stream;
    if (cells == 2)
        return launch_fwd<2, true>(tiles, st, rgb_only, height, width, tile_start, tile_end, sorted_payload,
                                   rec, image, depth, acc_alpha, last_effective, valid_count);
    if (!split)
        return launch_fwd<4, false>(tiles, st, rgb_only, height, width, tile_start, tile_end,
                                    sorted_payload, rec, image, depth, acc_alpha, last_effective, valid_count);
    return launch_fwd<4, true>(tiles, st, rgb_only, height, width, tile_start, tile_end, sorted_payload, rec,
                               image, depth, acc_alpha, last_effective, valid_count);
}
