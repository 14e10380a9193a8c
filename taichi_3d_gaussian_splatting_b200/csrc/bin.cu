// bin.cu -- the tile lists without materialising a (tile id, payload) pair per list entry.
//
// Replaces, for the default pipeline, the chain  emit tile keys -> stable pair sort by tile id ->
// tile ranges  (reference R.py:131-193 and the Tensor.sort() of R.py:947-950) by two expanding,
// stable multisplits over the depth-ordered points:
//
//   pass 1 (columns):  point i (in depth order), covering tile columns [x0, x1) and rows [y0, y1),
//                      appends one 8-byte entry (payload, y0, y1-1) to the list of every column it
//                      covers.  The column lists are contiguous segments of one array, in column order;
//   pass 2 (rows):     every entry of that array appends its payload to the list of every row it
//                      covers.  Row r's list is then ordered by (column, depth order): the order of
//                      a stable sort of the pairs by tile id = r * cols + column.
//
// Both passes are one kernel template: a CTA takes a chunk of 1024 input entries, marks for every bin
// (column or row) which entries cover it in a 1024-bit mask in shared memory, ranks by popcount, takes
// its per-bin offset among all chunks by a decoupled look-back (as the onesweep sort does per
// digit) and writes.  The chunk of pass 2 that contains the first entry of a column also knows,
// for every row, where that (row, column) tile starts: the tile ranges fall out of the pass, no tile
// histogram and no scan over the K sorted ids is needed.
//
// HBM traffic at C2 (M = 0.93M points, K = 12.2M list entries, E = 3.2M column entries):
//   12 M + 8 E (pass 1) + 8 E + 4 K (pass 2)  ~ 0.11 GB   against
//   8 K (emit) + 2 x 16 K (two sort passes) + 4 K (ranges) ~ 0.54 GB for the pair-sort chain.
// The result (sorted payload, tile_start, tile_end) is bit-identical to that chain's.
#include <algorithm>
#include <cstdlib>

#include "gs_common.cuh"

namespace {

// GS_BIN_TIGHT_LOOPS: 1 (default) = the three short data-dependent loops of a pass (mark, scatter,
// copy-out: 2 to 4 trips each) are not unrolled: the unrolled forms pay their remainder handling on every
// entry (C2 0.148 -> 0.142 ms, C5 1.369 -> 1.357 ms for the whole call)
#ifndef GS_BIN_TIGHT_LOOPS
#define GS_BIN_TIGHT_LOOPS 1
#endif
#if GS_BIN_TIGHT_LOOPS
#define GS_BIN_SHORT_LOOP _Pragma("unroll 1")
#else
#define GS_BIN_SHORT_LOOP
#endif

constexpr int kBinBlock = 256;
#ifndef GS_BIN_CHUNK
#define GS_BIN_CHUNK 1024
#endif
constexpr int kBinChunk = GS_BIN_CHUNK;       // input entries per CTA (tuning switch; 2048 measured)
constexpr int kBinWords = kBinChunk / 32;     // mask words per bin
constexpr int kBinItems = kBinChunk / kBinBlock;
constexpr int kMaxBins = 256;                 // tile columns / rows: images up to 4096 x 4096
constexpr uint32_t kBinAggregate = 1u << 30;
constexpr uint32_t kBinPrefix = 2u << 30;
constexpr uint32_t kBinValue = (1u << 30) - 1;

// workspace, in 32-bit words: the two histograms, their exclusive scans (written by the last CTA of
// bin_count_kernel), the chunk tickets of the two passes, the totals and that kernel's CTA counter
constexpr int kWsColHist = 0, kWsRowHist = kMaxBins, kWsColExcl = 2 * kMaxBins, kWsRowExcl = 3 * kMaxBins;
constexpr int kWsTickets = 4 * kMaxBins, kWsColTotal = kWsTickets + 2, kWsCountDone = kWsTickets + 3;
constexpr int kWsHeader = 4 * kMaxBins + 8;

// exclusive scan of one value per thread over the 256 threads of the CTA; *total = the sum
__device__ __forceinline__ uint32_t block_exscan(uint32_t v, uint32_t *s_warp, uint32_t *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    __syncthreads();  // s_warp may still be read from a previous call
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0, sum = 0;
#pragma unroll
    for (int w = 0; w < kBinBlock / 32; ++w) {
        const uint32_t t = s_warp[w];
        if (w < warp) off += t;
        sum += t;
    }
    *total = sum;
    return off + incl - v;
}

// per-column entry counts and per-row list lengths of the M visible points (any order); the last CTA
// to finish scans both histograms once for every CTA of the two passes
__global__ void __launch_bounds__(256)
bin_count_kernel(const ushort4 *__restrict__ bbox, long long m, uint32_t *__restrict__ ws)
{
    __shared__ uint32_t s_col[kMaxBins], s_row[kMaxBins];
    __shared__ uint32_t s_warp[kBinBlock / 32];
    __shared__ bool s_last;
    s_col[threadIdx.x] = 0;
    s_row[threadIdx.x] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
        const ushort4 b = bbox[i];
        const int w = (int)b.y - (int)b.x, h = (int)b.w - (int)b.z;
        if (w <= 0 || h <= 0) continue;
        for (int c = b.x; c < (int)b.y; ++c) atomicAdd(&s_col[c], 1u);
        for (int r = b.z; r < (int)b.w; ++r) atomicAdd(&s_row[r], (uint32_t)w);
    }
    __syncthreads();
    if (s_col[threadIdx.x]) atomicAdd(ws + kWsColHist + threadIdx.x, s_col[threadIdx.x]);
    if (s_row[threadIdx.x]) atomicAdd(ws + kWsRowHist + threadIdx.x, s_row[threadIdx.x]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(ws + kWsCountDone, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    uint32_t total;
    const uint32_t col_excl = block_exscan(gs_ld_volatile_u32(ws + kWsColHist + threadIdx.x), s_warp, &total);
    ws[kWsColExcl + threadIdx.x] = col_excl;
    if (threadIdx.x == 0) ws[kWsColTotal] = total;
    ws[kWsRowExcl + threadIdx.x] = block_exscan(gs_ld_volatile_u32(ws + kWsRowHist + threadIdx.x), s_warp, &total);
}

constexpr int kBinStageBytes = 32768;  // shared-memory staging of a chunk's output
template <int kPass> struct BinOut { typedef unsigned long long type; };
template <> struct BinOut<2> { typedef int32_t type; };

// kPass 1: input = depth-ordered points, bins = tile columns, output = 8-byte entries
// kPass 2: input = those entries,        bins = tile rows,    output = payload + tile starts
template <int kPass>
__global__ void __launch_bounds__(kBinBlock)
bin_pass_kernel(const int32_t *__restrict__ order, const ushort4 *__restrict__ bbox, long long m,
                const unsigned long long *__restrict__ entries_in, unsigned long long *__restrict__ entries_out,
                int32_t *__restrict__ payload_out, uint32_t *__restrict__ ws, uint32_t *__restrict__ status,
                int cols, int rows, uint32_t *__restrict__ starts, const int staged)
{
    extern __shared__ __align__(16) unsigned long long smem[];
    const int nb = (kPass == 1) ? cols : rows;
    // [kBinWords][nb] pairs: .y = which of the word's 32 entries cover the bin, .x = output position of
    // the first of them (filled in after the look-back): one 8-byte load per (entry, bin) in the write loop
    uint2 *s_pair = reinterpret_cast<uint2 *>(smem);
    __shared__ uint32_t s_base[kMaxBins];     // output position of this chunk's first item of a bin
    __shared__ uint32_t s_total[kMaxBins];    // this chunk's items per bin
    __shared__ uint32_t s_colbase[kMaxBins];  // pass 2: first input entry of every column
    __shared__ int s_bound[kMaxBins];         // pass 2: columns that start inside this chunk
    __shared__ int s_nbound;
    __shared__ unsigned int s_ticket;
    __shared__ uint32_t s_loc[kMaxBins];      // chunk-local position of a bin's first item
    __shared__ uint32_t s_adj[kMaxBins];
    __shared__ uint32_t s_warp[kBinBlock / 32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // input entries of this pass (pass 2: known on the device only, so the host sizes the grid for what
    // fits the machine and every CTA takes chunks, in order, until they run out)
    const long long n = (kPass == 1) ? m : (long long)ws[kWsColTotal];
    // bin bases: the scans of bin_count_kernel's histograms
    const uint32_t bin_excl = ws[(kPass == 1 ? kWsColExcl : kWsRowExcl) + tid];
    if (kPass == 2) s_colbase[tid] = ws[kWsColExcl + tid];
    for (;;) {
    __syncthreads();  // the previous chunk's shared state is no longer read
    if (tid == 0) {
        s_ticket = atomicAdd(ws + kWsTickets + (kPass - 1), 1u);
        s_nbound = 0;
    }
    for (int i = tid; i < kBinWords * nb; i += kBinBlock) smem[i] = 0ull;
    __syncthreads();
    const long long chunk = s_ticket;
    const long long chunk_start = chunk * kBinChunk;
    if (chunk_start >= n) return;  // uniform: no chunk left
    const bool last = chunk_start + kBinChunk >= n;

    // ---- load: warp w of round j owns the 32 consecutive entries of mask word j * 8 + w ------------
    int lo[kBinItems], hi[kBinItems];
    unsigned long long value[kBinItems];
#pragma unroll
    for (int j = 0; j < kBinItems; ++j) {
        const long long gi = chunk_start + j * kBinBlock + tid;
        lo[j] = 1;
        hi[j] = 0;
        value[j] = 0;
        if (gi < n) {
            if (kPass == 1) {
                const int row = __ldg(order + gi);
                const ushort4 b = __ldg(bbox + row);
                if ((int)b.y > (int)b.x && (int)b.w > (int)b.z) {
                    lo[j] = b.x;
                    hi[j] = (int)b.y - 1;
                    value[j] = (unsigned long long)(uint32_t)row |
                               ((unsigned long long)((uint32_t)b.z | ((uint32_t)(b.w - 1) << 16)) << 32);
                }
            } else {
                const unsigned long long e = __ldg(entries_in + gi);
                lo[j] = (int)((e >> 32) & 0xffffu);
                hi[j] = (int)(e >> 48);
                value[j] = e;
            }
        }
    }

    // ---- mark ------------------------------------------------------------------------------------
    const uint32_t bit = 1u << lane, lt = bit - 1u;
#pragma unroll
    for (int j = 0; j < kBinItems; ++j) {
        uint2 *w = s_pair + (j * (kBinBlock / 32) + warp) * nb;
        GS_BIN_SHORT_LOOP
        for (int b = lo[j]; b <= hi[j]; ++b) atomicOr(&w[b].y, bit);
    }
    __syncthreads();

    // ---- per bin: prefix of the popcounts over the words, publish, look back ---------------------
    if (tid < nb) {
        uint32_t run = 0;
#pragma unroll 8
        for (int w = 0; w < kBinWords; ++w) run += (uint32_t)__popc(s_pair[w * nb + tid].y);
        s_total[tid] = run;
        uint32_t *my = status + chunk * nb + tid;
        gs_st_volatile_u32(my, (chunk == 0 ? kBinPrefix : kBinAggregate) | run);
        uint32_t excl = 0;
        if (chunk > 0) {
            constexpr int kLook = 8;
            long long b = chunk - 1;
            bool done = false;
            while (!done) {
                uint32_t w[kLook];
#pragma unroll
                for (int q = 0; q < kLook; ++q)
                    w[q] = (b - q >= 0) ? gs_ld_volatile_u32(status + (b - q) * nb + tid) : kBinPrefix;
#pragma unroll
                for (int q = 0; q < kLook; ++q) {
                    if (done) break;
                    uint32_t sw = w[q];
                    while ((sw >> 30) == 0) sw = gs_ld_volatile_u32(status + (b - q) * nb + tid);
                    excl += sw & kBinValue;
                    done = (sw >> 30) == 2;
                }
                b -= kLook;
            }
            gs_st_volatile_u32(my, kBinPrefix | (excl + run));
        }
        uint32_t off = bin_excl + excl;
        s_base[tid] = off;
#pragma unroll 8
        for (int w = 0; w < kBinWords; ++w) {
            s_pair[w * nb + tid].x = off;
            off += (uint32_t)__popc(s_pair[w * nb + tid].y);
        }
    }
    if (kPass == 2 && tid < cols) {
        const long long v = (long long)s_colbase[tid];
        if (v >= chunk_start && (last || v < chunk_start + kBinChunk)) s_bound[atomicAdd(&s_nbound, 1)] = tid;
    }
    __syncthreads();

    // ---- write -----------------------------------------------------------------------------------
    // A direct store per (entry, bin) would scatter 4 / 8 bytes to a different list per lane (one 32-byte
    // sector each: 2.5 of 3.6 ms at 5M points / 3840x2160).  The chunk's output is laid out bin by bin in
    // shared memory instead and every bin's run (~50 items) is copied out as one contiguous piece; when
    // the chunk's output exceeds the staging buffer the bins are taken in groups that fit.
    typedef typename BinOut<kPass>::type OutT;
    OutT *out = (kPass == 1) ? reinterpret_cast<OutT *>(entries_out) : reinterpret_cast<OutT *>(payload_out);
    if (!staged) {
        // direct stores: no staging buffer, so more chunks per SM; the better trade for narrow boxes
#pragma unroll
        for (int j = 0; j < kBinItems; ++j) {
            const int word = j * (kBinBlock / 32) + warp;
            GS_BIN_SHORT_LOOP
            for (int b = lo[j]; b <= hi[j]; ++b) {
                const uint2 pm = s_pair[word * nb + b];
                out[pm.x + (uint32_t)__popc(pm.y & lt)] = (OutT)value[j];
            }
        }
    } else {
    OutT *s_stage = reinterpret_cast<OutT *>(smem + kBinWords * nb);
    constexpr uint32_t kCap = kBinStageBytes / sizeof(OutT);
    uint32_t num_out;
    {
        const uint32_t loc = block_exscan(tid < nb ? s_total[tid] : 0u, s_warp, &num_out);
        if (tid < nb) {
            s_loc[tid] = loc;
            s_adj[tid] = loc - s_base[tid];  // chunk-local position = global position + s_adj (mod 2^32)
        }
    }
    __syncthreads();
    for (int gs = 0; gs < nb;) {
        const uint32_t first = s_loc[gs];
        int ge = nb;
        if (num_out - first > kCap) {  // the largest ge whose bins [gs, ge) fit (a bin holds <= kBinChunk items)
            int lo_g = gs + 1, hi_g = nb - 1;
            while (lo_g < hi_g) {
                const int mid = (lo_g + hi_g + 1) >> 1;
                if (s_loc[mid] - first <= kCap) lo_g = mid; else hi_g = mid - 1;
            }
            ge = lo_g;
        }
#pragma unroll
        for (int j = 0; j < kBinItems; ++j) {
            const int word = j * (kBinBlock / 32) + warp;
            const int b1 = min(hi[j], ge - 1);
            GS_BIN_SHORT_LOOP
            for (int b = max(lo[j], gs); b <= b1; ++b) {
                const uint2 pm = s_pair[word * nb + b];
                const uint32_t pos = pm.x + (uint32_t)__popc(pm.y & lt);
                s_stage[pos + s_adj[b] - first] = (OutT)value[j];
            }
        }
        __syncthreads();
        for (int bb = gs + warp; bb < ge; bb += kBinBlock / 32) {
            const OutT *src = s_stage + (s_loc[bb] - first);
            OutT *dst = out + s_base[bb];
            const uint32_t cnt = s_total[bb];
            GS_BIN_SHORT_LOOP
            for (uint32_t i = lane; i < cnt; i += 32) dst[i] = src[i];
        }
        gs = ge;
        if (gs < nb) __syncthreads();
    }
    }

    // ---- pass 2: where every (row, column) tile starts, for the columns that begin in this chunk ----
    if (kPass == 2) {
        const int nbound = s_nbound;
        for (int q = 0; q < nbound; ++q) {
            const int c = s_bound[q];
            const uint32_t eb = (uint32_t)((long long)s_colbase[c] - chunk_start);
            if (tid < rows) {
                uint32_t at;
                if (eb >= (uint32_t)kBinChunk) {
                    at = s_base[tid] + s_total[tid];
                } else {
                    const uint2 pm = s_pair[(int)(eb >> 5) * nb + tid];
                    at = pm.x + (uint32_t)__popc(pm.y & ((1u << (eb & 31u)) - 1u));
                }
                starts[(size_t)tid * cols + c] = at;
            }
        }
    }
    }  // next chunk
}

// tile t's list is [starts[t], starts[t + 1]); an empty tile reads (0, 0) as in the reference, whose
// ranges are zero-initialised and only written at key transitions (R.py:175-193)
__global__ void bin_ranges_kernel(const uint32_t *__restrict__ starts, int num_tiles, long long k,
                                  int32_t *__restrict__ tile_start, int32_t *__restrict__ tile_end)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_tiles) return;
    uint32_t s = starts[t];
    uint32_t e = (t + 1 < num_tiles) ? starts[t + 1] : (uint32_t)k;
    if (s == e) s = e = 0u;
    tile_start[t] = (int32_t)s;
    tile_end[t] = (int32_t)e;
}

inline int64_t bin_chunks(int64_t n) { return (n + kBinChunk - 1) / kBinChunk; }

inline int64_t bin_entry_capacity(int64_t m, int64_t k, int cols)
{
    const int64_t by_cols = m * (int64_t)cols;  // a point has at most one entry per column
    return k < by_cols ? k : by_cols;           // ... and at least one row per entry
}

inline size_t bin_smem(int nb, bool staged) { return (size_t)kBinWords * nb * 8 + (staged ? kBinStageBytes : 0); }

}  // namespace

extern "C" int64_t gs_bin_workspace_bytes(int64_t m, int64_t k, int width, int height)
{
    if (m < 0 || k < 0 || width <= 0 || height <= 0 || width % kTile || height % kTile) return GS_E_INVALID;
    const int cols = width / kTile, rows = height / kTile;
    if (cols > kMaxBins || rows > kMaxBins) return GS_E_CAPACITY;
    const int64_t words = kWsHeader + bin_chunks(m) * cols + bin_chunks(bin_entry_capacity(m, k, cols)) * rows +
                          (int64_t)cols * rows;
    return words * 4;
}

extern "C" int gs_bin_tiles(const int32_t *order, const uint16_t *bbox, int64_t m, int64_t k, int width,
                            int height, uint64_t *entries, int32_t *payload, int32_t *tile_start,
                            int32_t *tile_end, void *workspace, void *stream)
{
    if (m < 0 || k < 0 || width <= 0 || height <= 0 || width % kTile || height % kTile) return GS_E_INVALID;
    const int cols = width / kTile, rows = height / kTile, num_tiles = cols * rows;
    if (cols > kMaxBins || rows > kMaxBins) return GS_E_CAPACITY;
    if (k >= (1ll << 30)) return GS_E_CAPACITY;  // 30-bit counts in the look-back words
    if (!tile_start || !tile_end) return GS_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    if (m == 0 || k == 0) {
        GS_CUDA_TRY(cudaMemsetAsync(tile_start, 0, sizeof(int32_t) * (size_t)num_tiles, st));
        GS_CUDA_TRY(cudaMemsetAsync(tile_end, 0, sizeof(int32_t) * (size_t)num_tiles, st));
        return GS_OK;
    }
    if (!order || !bbox || !entries || !payload || !workspace) return GS_E_INVALID;

    const int64_t chunks1 = bin_chunks(m), chunks2 = bin_chunks(bin_entry_capacity(m, k, cols));
    uint32_t *ws = (uint32_t *)workspace;
    uint32_t *status1 = ws + kWsHeader;
    uint32_t *status2 = status1 + chunks1 * cols;
    uint32_t *starts = status2 + chunks2 * rows;   // fully written by pass 2: not cleared
    GS_CUDA_TRY(cudaMemsetAsync(ws, 0, (size_t)(kWsHeader + chunks1 * cols + chunks2 * rows) * 4, st));

    static GsPerDeviceOnce attr_set;
    int dev = 0;
    if (attr_set.need(dev)) {
        GS_CUDA_TRY(cudaFuncSetAttribute(bin_pass_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)bin_smem(kMaxBins, true)));
        GS_CUDA_TRY(cudaFuncSetAttribute(bin_pass_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)bin_smem(kMaxBins, true)));
        attr_set.mark(dev);
    }
    const ushort4 *boxes = reinterpret_cast<const ushort4 *>(bbox);
    int count_blocks = (int)((m + 255) / 256);
    if (count_blocks > 148 * 4) count_blocks = 148 * 4;
    bin_count_kernel<<<count_blocks, 256, 0, st>>>(boxes, m, ws);
    GS_LAUNCH_CHECK();
    // the staged write-out is for the row split, whose chunks give every bin a long run; the column
    // split measures the same or slightly better with direct stores and the occupancy the staging buffer
    // would cost (C2 0.164 against 0.169 ms for the whole call, C5 1.581 against 1.591 ms)
    const bool stage1 = false;
    // persistent grids: as many CTAs as are resident at once (a CTA that cannot start before others have
    // finished would only find the ticket counter exhausted); pass 2's chunk count is a capacity, most of
    // which is usually not there (C2: 3 125 chunks of 11 952, C5: 27 k of 181 k)
    struct Resident {
        int cols = -1, rows = -1, ctas1 = 0, ctas2 = 0;
    };
    static Resident cache[64];
    static std::mutex cache_mu;
    Resident res;
    {
        std::lock_guard<std::mutex> lock(cache_mu);
        Resident &c = cache[dev & 63];
        if (c.cols != cols || c.rows != rows) {
            int sms = 0, per_sm1 = 0, per_sm2 = 0;
            GS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            GS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm1, bin_pass_kernel<1>, kBinBlock,
                                                                      bin_smem(cols, stage1)));
            GS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm2, bin_pass_kernel<2>, kBinBlock,
                                                                      bin_smem(rows, true)));
            c.cols = cols; c.rows = rows;
            c.ctas1 = sms * std::max(per_sm1, 1);
            c.ctas2 = sms * std::max(per_sm2, 1);
        }
        res = c;
    }
    const int64_t grid1 = std::min<int64_t>(chunks1, res.ctas1);
    const int64_t grid2 = std::min<int64_t>(chunks2, res.ctas2);
    bin_pass_kernel<1><<<(unsigned)grid1, kBinBlock, bin_smem(cols, stage1), st>>>(
        order, boxes, m, nullptr, (unsigned long long *)entries, nullptr, ws, status1, cols, rows, nullptr,
        stage1 ? 1 : 0);
    GS_LAUNCH_CHECK();
    bin_pass_kernel<2><<<(unsigned)grid2, kBinBlock, bin_smem(rows, true), st>>>(
        nullptr, nullptr, m, (const unsigned long long *)entries, nullptr, payload, ws, status2, cols, rows,
        starts, 1);
    GS_LAUNCH_CHECK();
    bin_ranges_kernel<<<(num_tiles + 255) / 256, 256, 0, st>>>(starts, num_tiles, k, tile_start, tile_end);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int64_t gs_bin_entries_capacity(int64_t m, int64_t k, int width)
{
    if (m < 0 || k < 0 || width <= 0 || width % kTile) return GS_E_INVALID;
    return bin_entry_capacity(m, k, width / kTile);
}
