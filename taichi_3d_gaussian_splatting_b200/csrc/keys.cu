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
