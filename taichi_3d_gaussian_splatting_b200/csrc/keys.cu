// keys.cu -- (tile, depth) key emission and tile-range extraction.  Integer work, HBM bound.
//
// Replaces generate_point_sort_key_by_num_overlap_tiles (R.py:131-172) and
// find_tile_start_and_end (R.py:175-193).
//
// Emission is block-cooperative: a block owns 256 consecutive visible points, whose key slots
// form one contiguous range; threads walk that range with stride 256 (coalesced 8-byte and
// 4-byte stores) and find the owning point by binary search over the block's offsets in
// shared memory, so a splat covering thousands of tiles costs no more per key than a small one.
#include "gs_common.cuh"

namespace {

constexpr int kEmitBlock = 256;

__global__ void __launch_bounds__(kEmitBlock)
emit_keys_kernel(const ushort4 *__restrict__ bbox, const int32_t *__restrict__ depth_key,
                 const long long *__restrict__ tile_offsets, long long m, long long k_total,
                 int tiles_per_row, long long *__restrict__ keys, int32_t *__restrict__ payload)
{
    __shared__ unsigned int s_end[kEmitBlock];  // inclusive end of each point's slots, block-relative
    __shared__ ushort4 s_box[kEmitBlock];
    __shared__ int32_t s_depth[kEmitBlock];
    __shared__ long long s_base;

    const int tid = threadIdx.x;
    const long long first = (long long)blockIdx.x * kEmitBlock;
    const long long i = first + tid;
    const int npts = (int)min((long long)kEmitBlock, m - first);
    if (tid == 0) s_base = tile_offsets[first];
    __syncthreads();
    const long long base = s_base;
    ushort4 box = make_ushort4(0, 0, 0, 0);
    unsigned int end_rel = 0;
    if (tid < npts) {
        box = bbox[i];
        int cnt = ((int)box.y - (int)box.x) * ((int)box.w - (int)box.z);
        end_rel = (unsigned int)(tile_offsets[i] - base) + (unsigned int)cnt;
        s_depth[tid] = depth_key[i];
    }
    s_box[tid] = box;
    s_end[tid] = end_rel;
    __syncthreads();
    const unsigned int total = s_end[npts - 1];
    for (unsigned int s = tid; s < total; s += kEmitBlock) {
        // first point whose inclusive end is > s (points with zero slots are skipped)
        int lo = 0, hi = npts - 1;
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (s_end[mid] > s) hi = mid; else lo = mid + 1;
        }
        ushort4 b = s_box[lo];
        int nv = (int)b.w - (int)b.z;
        int cnt = ((int)b.y - (int)b.x) * nv;
        unsigned int r = s - (s_end[lo] - (unsigned int)cnt);
        int tu = (int)b.x + (int)(r / (unsigned int)nv);
        int tv = (int)b.z + (int)(r % (unsigned int)nv);
        long long tile = (long long)(tu + tv * tiles_per_row);
        long long slot = base + s;
        if (slot < k_total) {
            keys[slot] = (long long)s_depth[lo] + (tile << 32);
            payload[slot] = (int32_t)(first + lo);
        }
    }
}

__global__ void tile_ranges_kernel(const long long *__restrict__ keys, long long k,
                                   int32_t *__restrict__ tile_start, int32_t *__restrict__ tile_end,
                                   int num_tiles)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    int t = (int)(keys[i] >> 32);
    if (i + 1 < k) {
        int tn = (int)(keys[i + 1] >> 32);
        if (t != tn) {
            if ((unsigned)tn < (unsigned)num_tiles) tile_start[tn] = (int32_t)(i + 1);
            if ((unsigned)t < (unsigned)num_tiles) tile_end[t] = (int32_t)(i + 1);
        }
    } else if ((unsigned)t < (unsigned)num_tiles) {
        tile_end[t] = (int32_t)k;
    }
}


// ---- depth-first pipeline -------------------------------------------------------------------------
// The visible points arrive as `order`: their row indices sorted by depth key (stable).  A block owns
// 256 consecutive entries of that order, counts their tiles, takes its exclusive offset among all
// blocks by a decoupled look-back (64-bit status words: flag << 62 | count), and emits its
// contiguous slot range exactly like emit_keys_kernel -- but the key is the tile id alone.  A
// stable sort of these pairs by tile id then yields, per tile, depth order with ties in ascending
// point order: the order a stable sort of the reference's 64-bit keys produces.
constexpr unsigned long long kEmitAggregate = 1ull << 62;
constexpr unsigned long long kEmitPrefix = 2ull << 62;
constexpr int kEmitChunk = 2048;  // slots per owner scan: 8 per thread
constexpr int kEmitHistPasses = 4;  // 32-bit tile ids: at most four 8-bit digits
constexpr unsigned long long kEmitValueMask = (1ull << 62) - 1;

__global__ void __launch_bounds__(kEmitBlock)
emit_tile_keys_kernel(const int32_t *__restrict__ order, const ushort4 *__restrict__ bbox, long long m,
                      long long k_total, int tiles_per_row, uint32_t *__restrict__ tile_ids,
                      int32_t *__restrict__ payload, unsigned long long *__restrict__ status,
                      unsigned int *__restrict__ ticket, uint32_t *__restrict__ digit_hist,
                      const int hist_passes)
{
    // digit histograms of the tile ids this CTA emits (what the pair sort's first kernel would
    // otherwise re-read all K ids for): hist_passes x 256 bins of 8-bit digits, low digit first
    __shared__ uint32_t s_hist[kEmitHistPasses * 256];
    if (digit_hist)
        for (int i = threadIdx.x; i < hist_passes * 256; i += kEmitBlock) s_hist[i] = 0;
    __shared__ unsigned int s_start[kEmitBlock];
    __shared__ float s_rcp[kEmitBlock];
    __shared__ __align__(16) int s_owner[kEmitChunk];
    __shared__ int s_warp_max[kEmitBlock / 32];
    __shared__ ushort4 s_box[kEmitBlock];
    __shared__ int32_t s_row[kEmitBlock];
    __shared__ unsigned int s_warp_total[kEmitBlock / 32];
    __shared__ unsigned long long s_base;
    __shared__ unsigned int s_ticket;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_ticket = atomicAdd(ticket, 1u);
    __syncthreads();
    const int block = (int)s_ticket;
    const long long first = (long long)block * kEmitBlock;
    const int npts = (int)min((long long)kEmitBlock, m - first);

    ushort4 box = make_ushort4(0, 0, 0, 0);
    int row = 0;
    unsigned int cnt = 0;
    if (tid < npts) {
        row = __ldg(order + first + tid);
        box = bbox[row];
        cnt = (unsigned int)(((int)box.y - (int)box.x) * ((int)box.w - (int)box.z));
    }
    // inclusive scan of the counts over the block
    unsigned int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp_total[warp] = incl;
    __syncthreads();
    unsigned int warp_excl = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < kEmitBlock / 32; ++w) {
        unsigned int t = s_warp_total[w];
        if (w < warp) warp_excl += t;
        block_total += t;
    }
    s_box[tid] = box;
    s_row[tid] = row;
    s_start[tid] = warp_excl + incl - cnt;
    s_rcp[tid] = 1.0f / (float)max((int)box.w - (int)box.z, 1);
    if (warp == 0) {
        if (lane == 0)
            gs_st_volatile_u64(status + block, (block == 0 ? kEmitPrefix : kEmitAggregate) | block_total);
        unsigned long long excl = 0;
        if (block > 0) {
            int base = block - 1;
            while (true) {
                int idx = base - lane;
                unsigned long long st = kEmitPrefix;  // positions before block 0: empty prefix
                if (idx >= 0) {
                    do { st = gs_ld_volatile_u64(status + idx); } while ((st >> 62) == 0);
                }
                unsigned int pmask = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                int firstp = pmask ? (__ffs(pmask) - 1) : 32;
                unsigned long long contrib = (lane <= firstp) ? (st & kEmitValueMask) : 0ull;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
                excl += contrib;
                if (pmask) break;
                base -= 32;
            }
            if (lane == 0) gs_st_volatile_u64(status + block, kEmitPrefix | (excl + block_total));
        }
        if (lane == 0) s_base = excl;
    }
    __syncthreads();
    const unsigned long long base = s_base;
    const unsigned int total = block_total;
    // Owner of every slot by a max-scan instead of a search per slot: in chunks of kEmitChunk slots,
    // each entry marks the slot its run starts at, an inclusive max-scan spreads the mark over the
    // run (8 consecutive slots per thread, then warp and block level), and the threads then walk the
    // chunk with stride kEmitBlock so that the stores are coalesced.
    const unsigned int my_start = warp_excl + incl - cnt;
    int4 *s_owner4 = reinterpret_cast<int4 *>(s_owner);
    for (unsigned int cb = 0; cb < total; cb += kEmitChunk) {
        s_owner4[2 * tid] = make_int4(-1, -1, -1, -1);
        s_owner4[2 * tid + 1] = make_int4(-1, -1, -1, -1);
        __syncthreads();
        if (cnt > 0) {
            if (my_start >= cb) {
                if (my_start < cb + kEmitChunk) s_owner[my_start - cb] = tid;
            } else if (my_start + cnt > cb) {
                s_owner[0] = tid;  // the run that straddles the chunk boundary
            }
        }
        __syncthreads();
        int4 a = s_owner4[2 * tid], b = s_owner4[2 * tid + 1];
        a.y = max(a.y, a.x);
        a.z = max(a.z, a.y);
        a.w = max(a.w, a.z);
        b.x = max(b.x, a.w);
        b.y = max(b.y, b.x);
        b.z = max(b.z, b.y);
        b.w = max(b.w, b.z);
        int run = b.w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, run, d);
            if (lane >= d) run = max(run, o);
        }
        int before = __shfl_up_sync(0xffffffffu, run, 1);
        if (lane == 0) before = -1;
        if (lane == 31) s_warp_max[warp] = run;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < kEmitBlock / 32; ++w)
            if (w < warp) before = max(before, s_warp_max[w]);
        a.x = max(a.x, before);
        a.y = max(a.y, before);
        a.z = max(a.z, before);
        a.w = max(a.w, before);
        b.x = max(b.x, before);
        b.y = max(b.y, before);
        b.z = max(b.z, before);
        b.w = max(b.w, before);
        s_owner4[2 * tid] = a;
        s_owner4[2 * tid + 1] = b;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < kEmitChunk / kEmitBlock; ++j) {
            const unsigned int ls = tid + kEmitBlock * j, sl = cb + ls;
            const long long slot = (long long)base + sl;
            if (sl < total && slot < k_total) {
                const int o = s_owner[ls];
                const ushort4 bx = s_box[o];
                const unsigned int nv = (unsigned int)((int)bx.w - (int)bx.z);
                const unsigned int r = sl - s_start[o];
                // r / nv from a float estimate (|error| <= 1 for r < 2^22) and one correction
                unsigned int q = (unsigned int)((float)r * s_rcp[o]);
                int rem = (int)(r - q * nv);
                if (rem < 0) {
                    rem += (int)nv;
                    --q;
                } else if (rem >= (int)nv) {
                    rem -= (int)nv;
                    ++q;
                }
                const uint32_t id = (uint32_t)(((int)bx.x + (int)q) + ((int)bx.z + rem) * tiles_per_row);
                tile_ids[slot] = id;
                payload[slot] = s_row[o];
                if (digit_hist)
                    for (int p = 0; p < hist_passes; ++p) atomicAdd(&s_hist[p * 256 + ((id >> (8 * p)) & 255u)], 1u);
            }
        }
        __syncthreads();
    }
    if (digit_hist)
        for (int i = tid; i < hist_passes * 256; i += kEmitBlock)
            if (s_hist[i]) atomicAdd(digit_hist + i, s_hist[i]);
}

// four ids per thread (one 16-byte load when the array is 16-byte aligned) plus the successor
__global__ void __launch_bounds__(256)
tile_ranges32_kernel(const uint32_t *__restrict__ tile_ids, long long k, int32_t *__restrict__ tile_start,
                     int32_t *__restrict__ tile_end, int num_tiles, int aligned)
{
    const long long i0 = 4 * ((long long)blockIdx.x * blockDim.x + threadIdx.x);
    if (i0 >= k) return;
    unsigned t[5];
    const int n = (int)min(4ll, k - i0);
    if (aligned && n == 4) {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(tile_ids + i0));
        t[0] = q.x; t[1] = q.y; t[2] = q.z; t[3] = q.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) t[j] = (j < n) ? __ldg(tile_ids + i0 + j) : 0u;
    }
    const bool has_next = i0 + n < k;
    t[4] = has_next ? __ldg(tile_ids + i0 + n) : 0u;
    if (n < 4) t[n] = t[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (j >= n) break;
        const long long i = i0 + j;
        if (j + 1 < n || has_next) {
            const unsigned a = t[j], b = t[j + 1];
            if (a != b) {
                if (b < (unsigned)num_tiles) tile_start[b] = (int32_t)(i + 1);
                if (a < (unsigned)num_tiles) tile_end[a] = (int32_t)(i + 1);
            }
        } else if (t[j] < (unsigned)num_tiles) {
            tile_end[t[j]] = (int32_t)k;
        }
    }
}

// the reference's 64-bit key of every sorted entry, rebuilt for inspection (R.py:159-160)
__global__ void sorted_keys_kernel(const uint32_t *__restrict__ tile_ids, const int32_t *__restrict__ payload,
                                   const int32_t *__restrict__ depth_key, long long k,
                                   long long *__restrict__ keys)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    keys[i] = (long long)depth_key[payload[i]] + ((long long)tile_ids[i] << 32);
}

}  // namespace

extern "C" int gs_emit_keys(const uint16_t *bbox, const int32_t *depth_key,
                            const int64_t *tile_offsets, int64_t m, int64_t k, int width,
                            int64_t *keys, int32_t *payload, void *stream)
{
    if (m < 0 || k < 0 || width <= 0 || width % kTile) return GS_E_INVALID;
    if (k >= (1ll << 31)) return GS_E_CAPACITY;  // int32 indices downstream (R.py:930, 954-957)
    if (m == 0 || k == 0) return GS_OK;
    if (!bbox || !depth_key || !tile_offsets || !keys || !payload) return GS_E_INVALID;
    int blocks = (int)((m + kEmitBlock - 1) / kEmitBlock);
    emit_keys_kernel<<<blocks, kEmitBlock, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const ushort4 *>(bbox), depth_key, (const long long *)tile_offsets, m, k,
        width / kTile, (long long *)keys, payload);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int gs_tile_ranges(const int64_t *sorted_keys, int64_t k, int32_t *tile_start,
                              int32_t *tile_end, int num_tiles, void *stream)
{
    if (k < 0 || num_tiles <= 0 || !tile_start || !tile_end) return GS_E_INVALID;
    if (k >= (1ll << 31)) return GS_E_CAPACITY;
    cudaStream_t st = (cudaStream_t)stream;
    GS_CUDA_TRY(cudaMemsetAsync(tile_start, 0, sizeof(int32_t) * (size_t)num_tiles, st));
    GS_CUDA_TRY(cudaMemsetAsync(tile_end, 0, sizeof(int32_t) * (size_t)num_tiles, st));
    if (k == 0) return GS_OK;
    if (!sorted_keys) return GS_E_INVALID;
    int threads = 256;
    long long blocks = (k + threads - 1) / threads;
    tile_ranges_kernel<<<(unsigned)blocks, threads, 0, st>>>((const long long *)sorted_keys, k,
                                                             tile_start, tile_end, num_tiles);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int64_t gs_emit_workspace_bytes(int64_t m)
{
    int64_t blocks = (m + kEmitBlock - 1) / kEmitBlock;
    return (blocks + 2) * 8;
}

extern "C" int gs_emit_tile_keys(const int32_t *order, const uint16_t *bbox, int64_t m, int64_t k,
                                 int width, uint32_t *tile_ids, int32_t *payload, void *workspace,
                                 uint32_t *digit_hist, int key_bits, void *stream)
{
    if (digit_hist && (key_bits < 1 || key_bits > 32)) return GS_E_INVALID;
    if (digit_hist)
        GS_CUDA_TRY(cudaMemsetAsync(digit_hist, 0, sizeof(uint32_t) * 256 * kEmitHistPasses, (cudaStream_t)stream));
    if (m < 0 || k < 0 || width <= 0 || width % kTile) return GS_E_INVALID;
    if (k >= (1ll << 31)) return GS_E_CAPACITY;  // int32 indices downstream (R.py:930, 954-957)
    if (m == 0 || k == 0) return GS_OK;
    if (!order || !bbox || !tile_ids || !payload || !workspace) return GS_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    int blocks = (int)((m + kEmitBlock - 1) / kEmitBlock);
    GS_CUDA_TRY(cudaMemsetAsync(workspace, 0, (size_t)gs_emit_workspace_bytes(m), st));
    unsigned long long *status = (unsigned long long *)workspace;
    emit_tile_keys_kernel<<<blocks, kEmitBlock, 0, st>>>(
        order, reinterpret_cast<const ushort4 *>(bbox), m, k, width / kTile, tile_ids, payload, status,
        (unsigned int *)(status + blocks), digit_hist, digit_hist ? (key_bits + 7) / 8 : 0);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int gs_tile_ranges32(const uint32_t *sorted_tile_ids, int64_t k, int32_t *tile_start,
                                int32_t *tile_end, int num_tiles, void *stream)
{
    if (k < 0 || num_tiles <= 0 || !tile_start || !tile_end) return GS_E_INVALID;
    if (k >= (1ll << 31)) return GS_E_CAPACITY;
    cudaStream_t st = (cudaStream_t)stream;
    GS_CUDA_TRY(cudaMemsetAsync(tile_start, 0, sizeof(int32_t) * (size_t)num_tiles, st));
    GS_CUDA_TRY(cudaMemsetAsync(tile_end, 0, sizeof(int32_t) * (size_t)num_tiles, st));
    if (k == 0) return GS_OK;
    if (!sorted_tile_ids) return GS_E_INVALID;
    int threads = 256;
    long long blocks = ((k + 3) / 4 + threads - 1) / threads;
    int aligned = ((uintptr_t)sorted_tile_ids % 16) == 0;
    tile_ranges32_kernel<<<(unsigned)blocks, threads, 0, st>>>(sorted_tile_ids, k, tile_start, tile_end,
                                                               num_tiles, aligned);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int gs_sorted_keys(const uint32_t *sorted_tile_ids, const int32_t *sorted_payload,
                              const int32_t *depth_key, int64_t k, int64_t *keys, void *stream)
{
    if (k < 0) return GS_E_INVALID;
    if (k == 0) return GS_OK;
    if (!sorted_tile_ids || !sorted_payload || !depth_key || !keys) return GS_E_INVALID;
    int threads = 256;
    long long blocks = (k + threads - 1) / threads;
    sorted_keys_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(
        sorted_tile_ids, sorted_payload, depth_key, k, (long long *)keys);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
