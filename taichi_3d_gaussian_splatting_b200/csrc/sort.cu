// sort.cu -- stable LSD radix sorts, onesweep style: one histogram pass over the keys, then one
// read + one scattered write per 8-bit digit with a decoupled look-back over the per-tile digit
// counts (no separate scan pass).  Ties keep their input order.
//
// Two users:
//   * the depth-first pipeline (default): the M visible points are sorted by their 32-bit depth
//     key (gs_sort_points_by_depth, values = row indices generated on the fly), tile keys are
//     emitted in that order, and the K (tile id, payload) pairs are sorted by tile id only
//     (gs_sort_pairs32) -- 8 bytes per pair and two passes at C2 instead of 12 bytes and four;
//   * the literal pipeline: (int64 key, int32 payload) pairs as the reference builds them
//     (gs_sort_pairs; replaces Tensor.sort() + gather, reference R.py:947-950).  Only the
//     significant bits are sorted: a key is depth + (tile << 32); the host passes how many low
//     bits the depth needs and how many bits the tile id needs, and the digits are taken from
//     the compacted value (tile << depth_bits) | depth.  depth_bits = tile_bits = 32 selects a
//     full signed 64-bit sort (sign bit flipped), used when depths may be negative.
#include <cstdlib>

#include "gs_common.cuh"

namespace {

constexpr int kRadix = 256;
constexpr int kSortBlock = 256;
constexpr int kSortWarps = kSortBlock / 32;
constexpr int kItems = 16;
constexpr int kSortTile = kSortBlock * kItems;
constexpr int kMaxPasses = 8;
constexpr uint32_t kStatAggregate = 1u << 30;
constexpr uint32_t kStatPrefix = 2u << 30;
constexpr uint32_t kStatMask = (1u << 30) - 1;

struct KeySpec {
    int depth_bits;
    unsigned long long depth_mask;
    unsigned long long xor_mask;
};

__device__ __forceinline__ unsigned long long compact_key(long long key, const KeySpec &s)
{
    unsigned long long u = (unsigned long long)key;
    unsigned long long hi = u >> 32, lo = u & 0xffffffffull;
    unsigned long long c = (s.depth_bits >= 32) ? u : ((hi << s.depth_bits) | (lo & s.depth_mask));
    return c ^ s.xor_mask;
}
// 32-bit keys are sorted as they are (the spec is ignored)
__device__ __forceinline__ unsigned int compact_key(unsigned int key, const KeySpec &) { return key; }

template <typename KeyT>
__global__ void __launch_bounds__(256)
histogram_kernel(const KeyT *__restrict__ keys, long long k, const long long *__restrict__ k_dev,
                 KeySpec spec, int num_passes, uint32_t *__restrict__ ghist)
{
    if (k_dev) k = *k_dev;  // the count lives on the device; the host only knows a capacity
    __shared__ uint32_t sh[kMaxPasses * kRadix];
    for (int i = threadIdx.x; i < num_passes * kRadix; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    // four independent loads in flight per thread (the loop is latency bound otherwise)
    long long i = tid0;
    for (; i + 3 * stride < k; i += 4 * stride) {
        KeyT q[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) q[j] = keys[i + j * stride];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            auto c = compact_key(q[j], spec);
            for (int p = 0; p < num_passes; ++p)
                atomicAdd(&sh[p * kRadix + (int)((c >> (8 * p)) & 0xff)], 1u);
        }
    }
    for (; i < k; i += stride) {
        auto c = compact_key(keys[i], spec);
        for (int p = 0; p < num_passes; ++p)
            atomicAdd(&sh[p * kRadix + (int)((c >> (8 * p)) & 0xff)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < num_passes * kRadix; i += blockDim.x)
        if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// exclusive scan of each pass's 256-bin histogram; one block of 256 threads per pass
__global__ void __launch_bounds__(256)
scan_histogram_kernel(const uint32_t *__restrict__ ghist, uint32_t *__restrict__ gbase, long long k,
                      const long long *__restrict__ k_dev, unsigned int *__restrict__ trivial)
{
    __shared__ uint32_t s_warp[kSortWarps];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (k_dev) k = *k_dev;
    uint32_t v = ghist[blockIdx.x * kRadix + tid];
    // every key has the same digit in this pass: the pass is the identity permutation and the
    // onesweep kernel only copies (e.g. the top pass of depth keys far below the far plane)
    if (k > 0 && (long long)v == k) trivial[blockIdx.x] = 1u;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    gbase[blockIdx.x * kRadix + tid] = off + incl - v;
}

// vals_in == nullptr: the value of an element is its index (first pass of an argsort)
// digit_bits (1..8): how many bits of this pass's digit can be non-zero (the last pass of a key
// whose width is not a multiple of 8 has fewer).  kBallot: build the peer masks from digit_bits
// warp ballots instead of match.any.
template <typename KeyT, bool kBallot>
__global__ void __launch_bounds__(kSortBlock, sizeof(KeyT) == 4 ? 3 : 2)
onesweep_kernel(const KeyT *__restrict__ keys_in, const int32_t *__restrict__ vals_in,
                KeyT *__restrict__ keys_out, int32_t *__restrict__ vals_out, long long k,
                const long long *__restrict__ k_dev, KeySpec spec, int shift, int digit_bits,
                const uint32_t *__restrict__ gbase, uint32_t *__restrict__ status,
                unsigned int *__restrict__ ticket, const unsigned int *__restrict__ trivial)
{
    if (k_dev) k = *k_dev;  // launched for a capacity: tiles past the real count leave at once
    extern __shared__ __align__(16) unsigned char smem[];
    KeyT *s_keys = reinterpret_cast<KeyT *>(smem);
    int32_t *s_vals = reinterpret_cast<int32_t *>(s_keys + kSortTile);
    uint32_t *s_warp_hist = reinterpret_cast<uint32_t *>(s_vals + kSortTile);
    uint32_t *s_digit_base = s_warp_hist + kSortWarps * kRadix;
    uint32_t *s_global_off = s_digit_base + kRadix;
    uint32_t *s_scan = s_global_off + kRadix;  // kSortWarps entries
    __shared__ unsigned int s_ticket;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_ticket = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) s_warp_hist[w * kRadix + tid] = 0;
    __syncthreads();
    const long long tile = s_ticket;
    const long long tile_base = tile * kSortTile;
    if (tile_base >= k) return;
    const int valid = (int)min((long long)kSortTile, k - tile_base);

    KeyT key[kItems];
    int32_t val[kItems];
    const int warp_first = warp * (kItems * 32);
    if (valid == kSortTile) {  // every tile but the last: no per-item bounds checks
        const KeyT *kp = keys_in + tile_base + warp_first + lane;
#pragma unroll
        for (int j = 0; j < kItems; ++j) key[j] = kp[j * 32];
        if (vals_in) {
            const int32_t *vp = vals_in + tile_base + warp_first + lane;
#pragma unroll
            for (int j = 0; j < kItems; ++j) val[j] = vp[j * 32];
        } else {
#pragma unroll
            for (int j = 0; j < kItems; ++j) val[j] = (int32_t)(tile_base + warp_first + j * 32 + lane);
        }
    } else {
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            int local = warp_first + j * 32 + lane;
            if (local < valid) {
                key[j] = keys_in[tile_base + local];
                val[j] = vals_in ? vals_in[tile_base + local] : (int32_t)(tile_base + local);
            } else {
                key[j] = 0;
                val[j] = 0;
            }
        }
    }

    if (*trivial) {  // identity pass (see scan_histogram_kernel): element i stays at i
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            const int local = warp_first + j * 32 + lane;
            if (local < valid) {
                keys_out[tile_base + local] = key[j];
                vals_out[tile_base + local] = val[j];
            }
        }
        return;
    }

    // ---- warp-level stable ranking per digit (match.any) ---------------------------------
    // Three unrolled phases so that the 16 items' match -> shared atomic -> shuffle chains
    // overlap instead of running back to back: (1) peer masks, (2) the lowest peer adds the group
    // size to the warp's digit counter and keeps the old value (shared atomics of one warp to
    // one address retire in issue order, so item j is ranked before item j+1), (3) broadcast.
    uint32_t rank[kItems];
    unsigned int peers[kItems];
    const unsigned int lanemask_lt = (1u << lane) - 1u;
    uint32_t *my_hist = s_warp_hist + warp * kRadix;
    const int pad_digit = (1 << digit_bits) - 1;  // padding of the last tile: the largest digit
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        int local = warp_first + j * 32 + lane;
        int d = (local < valid) ? (int)((compact_key(key[j], spec) >> shift) & 0xff) : pad_digit;
        if (kBallot) {
            unsigned int p = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                if (b < digit_bits) {  // warp-uniform
                    const bool bit = d & (1 << b);
                    const unsigned int vote = __ballot_sync(0xffffffffu, bit);
                    p &= bit ? vote : ~vote;
                }
            }
            peers[j] = p;
        } else {
            peers[j] = __match_any_sync(0xffffffffu, d);
        }
    }
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        int local = warp_first + j * 32 + lane;
        int d = (local < valid) ? (int)((compact_key(key[j], spec) >> shift) & 0xff) : pad_digit;
        rank[j] = 0;
        if ((peers[j] & lanemask_lt) == 0u) rank[j] = atomicAdd(&my_hist[d], (uint32_t)__popc(peers[j]));
    }
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        rank[j] = __shfl_sync(0xffffffffu, rank[j], __ffs(peers[j]) - 1) +
                  (uint32_t)__popc(peers[j] & lanemask_lt);
    }
    __syncthreads();

    // ---- per-digit exclusive scan over warps; thread `tid` owns digit `tid` -----------------
    uint32_t block_count = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
        uint32_t c = s_warp_hist[w * kRadix + tid];
        s_warp_hist[w * kRadix + tid] = block_count;
        block_count += c;
    }
    // the padding items of the last tile sit in the largest digit and must not be published
    uint32_t publish_count = block_count;
    if (tid == pad_digit) publish_count -= (uint32_t)(kSortTile - valid);
    gs_st_volatile_u32(status + tile * kRadix + tid,
                       (tile == 0 ? kStatPrefix : kStatAggregate) | publish_count);

    // exclusive scan of block_count over the 256 digits
    uint32_t incl = block_count;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w)
        if (w < warp) warp_off += s_scan[w];
    const uint32_t digit_base = warp_off + incl - block_count;
    s_digit_base[tid] = digit_base;
    __syncthreads();

    // ---- scatter the tile into shared memory in digit order -----------------------------------
    // positions first (independent shared loads), stores second: the warp issues in order, so a
    // load -> store chain per item would serialise the 16 items
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        int local = warp_first + j * 32 + lane;
        int d = (local < valid) ? (int)((compact_key(key[j], spec) >> shift) & 0xff) : pad_digit;
        rank[j] += s_digit_base[d] + my_hist[d];
    }
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        s_keys[rank[j]] = key[j];
        s_vals[rank[j]] = val[j];
    }

    // ---- decoupled look-back for digit `tid` -----------------------------------------------
    uint32_t excl = 0;
    if (tile > 0) {
        // kLook predecessors are read at once (independent loads in flight) and then consumed
        // in order; a word that is not published yet is polled on its own
        constexpr int kLook = 8;
        long long b = tile - 1;
        bool done = false;
        while (!done) {
            uint32_t w[kLook];
#pragma unroll
            for (int q = 0; q < kLook; ++q)
                w[q] = (b - q >= 0) ? gs_ld_volatile_u32(status + (b - q) * kRadix + tid) : kStatPrefix;
#pragma unroll
            for (int q = 0; q < kLook; ++q) {
                if (done) break;
                uint32_t sw = w[q];
                while ((sw >> 30) == 0) sw = gs_ld_volatile_u32(status + (b - q) * kRadix + tid);
                excl += sw & kStatMask;
                done = (sw >> 30) == 2;
            }
            b -= kLook;
        }
        gs_st_volatile_u32(status + tile * kRadix + tid, kStatPrefix | (excl + publish_count));
    }
    s_global_off[tid] = gbase[tid] + excl - digit_base;  // unsigned wrap-around is intended
    __syncthreads();

    // ---- coalesced write-out: consecutive smem slots of one digit are consecutive in global -----
    // (loads of all 16 slots first, then the offsets, then the stores -- same reason as above)
    {
        KeyT kk[kItems];
        int32_t vv[kItems];
        uint32_t off[kItems];
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            const int i = tid + j * kSortBlock;
            kk[j] = (i < valid) ? s_keys[i] : (KeyT)0;
            vv[j] = (i < valid) ? s_vals[i] : 0;
        }
#pragma unroll
        for (int j = 0; j < kItems; ++j)
            off[j] = s_global_off[(int)((compact_key(kk[j], spec) >> shift) & 0xff)];
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            const int i = tid + j * kSortBlock;
            if (i < valid) {
                const uint32_t pos = (uint32_t)i + off[j];
                keys_out[pos] = kk[j];
                vals_out[pos] = vv[j];
            }
        }
    }
}

constexpr size_t sort_smem(size_t key_bytes)
{
    return (size_t)kSortTile * (key_bytes + 4) + (size_t)kSortWarps * kRadix * 4 +
           (size_t)kRadix * 8 + (size_t)kSortWarps * 4;
}

inline int64_t sort_tiles(int64_t k) { return (k + kSortTile - 1) / kSortTile; }

// Runs `num_passes` onesweep passes over k elements.  The first pass reads (keys_first,
// vals_first) -- vals_first may be null (values = indices) -- and writes (keys_a, vals_a); later
// passes ping-pong between the a and b buffers.  Workspace layout:
// ghist | gbase | tickets (64 B) | status[passes][tiles][256].
template <typename KeyT>
int sort_impl(const KeyT *keys_first, const int32_t *vals_first, KeyT *keys_a, int32_t *vals_a,
              KeyT *keys_b, int32_t *vals_b, int64_t k, const KeySpec &spec, int total_bits,
              void *workspace, cudaStream_t st, const long long *k_dev = nullptr,
              const uint32_t *ready_hist = nullptr)
{
    const int num_passes = (total_bits + 7) / 8;
    const int64_t tiles = sort_tiles(k);
    uint32_t *ghist = (uint32_t *)workspace;
    uint32_t *gbase = ghist + kMaxPasses * kRadix;
    unsigned int *tickets = (unsigned int *)(gbase + kMaxPasses * kRadix);
    uint32_t *status = (uint32_t *)(tickets + 16);
    size_t used = (size_t)kMaxPasses * kRadix * 4 * 2 + 64 + (size_t)num_passes * tiles * kRadix * 4;
    GS_CUDA_TRY(cudaMemsetAsync(workspace, 0, used, st));

    static GsPerDeviceOnce attr_set;  // one per KeyT instantiation
    static bool ballot = true;
    int dev = 0;
    if (attr_set.need(dev)) {
        GS_CUDA_TRY(cudaFuncSetAttribute(onesweep_kernel<KeyT, false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sort_smem(sizeof(KeyT))));
        GS_CUDA_TRY(cudaFuncSetAttribute(onesweep_kernel<KeyT, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sort_smem(sizeof(KeyT))));
        const char *e = getenv("GS_SORT_RANK");  // tuning knob: "match" or "ballot"
        if (e) ballot = (e[0] == 'b');
        attr_set.mark(dev);
    }

    if (!ready_hist) {  // else: the producer of the keys counted the digits while writing them
        int hist_blocks = (int)min((int64_t)148 * 8, (k + 256 * 8 - 1) / (256 * 8));
        if (hist_blocks < 1) hist_blocks = 1;
        histogram_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(keys_first, k, k_dev, spec, num_passes, ghist);
        GS_LAUNCH_CHECK();
    }
    scan_histogram_kernel<<<num_passes, 256, 0, st>>>(ready_hist ? ready_hist : ghist, gbase, k, k_dev,
                                                      tickets + kMaxPasses);
    GS_LAUNCH_CHECK();

    const KeyT *kin = keys_first;
    const int32_t *vin = vals_first;
    for (int p = 0; p < num_passes; ++p) {
        KeyT *kout = (p & 1) ? keys_b : keys_a;
        int32_t *vout = (p & 1) ? vals_b : vals_a;
        const int digit_bits = min(8, total_bits - 8 * p);
        if (ballot)
            onesweep_kernel<KeyT, true><<<(unsigned)tiles, kSortBlock, sort_smem(sizeof(KeyT)), st>>>(
                kin, vin, kout, vout, k, k_dev, spec, 8 * p, digit_bits, gbase + p * kRadix,
                status + (size_t)p * tiles * kRadix, tickets + p, tickets + kMaxPasses + p);
        else
            onesweep_kernel<KeyT, false><<<(unsigned)tiles, kSortBlock, sort_smem(sizeof(KeyT)), st>>>(
                kin, vin, kout, vout, k, k_dev, spec, 8 * p, digit_bits, gbase + p * kRadix,
                status + (size_t)p * tiles * kRadix, tickets + p, tickets + kMaxPasses + p);
        GS_LAUNCH_CHECK();
        kin = kout;
        vin = vout;
    }
    return GS_OK;
}

inline int64_t sort_scratch_bytes(int64_t k)
{
    return (int64_t)kMaxPasses * kRadix * 4 * 2 + 64 + (int64_t)kMaxPasses * sort_tiles(k) * kRadix * 4;
}

inline int64_t align256(int64_t x) { return (x + 255) / 256 * 256; }

}  // namespace

extern "C" int64_t gs_sort_workspace_bytes(int64_t k) { return sort_scratch_bytes(k); }

extern "C" int64_t gs_point_sort_workspace_bytes(int64_t m)
{
    // scratch | keys a | keys b | values tmp
    return align256(sort_scratch_bytes(m)) + 3 * align256(4 * (m > 0 ? m : 1));
}

extern "C" int gs_sort_pairs(int64_t *keys, int32_t *payload, int64_t *keys_alt,
                             int32_t *payload_alt, int64_t k, int depth_bits, int tile_bits,
                             void *workspace, int *result_in_alt, void *stream)
{
    if (result_in_alt) *result_in_alt = 0;
    if (k < 0 || depth_bits < 0 || depth_bits > 32 || tile_bits < 0 || tile_bits > 32)
        return GS_E_INVALID;
    if (k >= (1ll << 30)) return GS_E_CAPACITY;  // 30-bit counts in the look-back words
    if (k <= 1) return GS_OK;
    if (!keys || !payload || !keys_alt || !payload_alt || !workspace) return GS_E_INVALID;

    KeySpec spec;
    spec.depth_bits = depth_bits;
    spec.depth_mask = depth_bits >= 32 ? 0xffffffffull : ((1ull << depth_bits) - 1ull);
    const bool full = (depth_bits == 32 && tile_bits == 32);
    spec.xor_mask = full ? 0x8000000000000000ull : 0ull;
    int num_passes = (depth_bits + tile_bits + 7) / 8;
    if (num_passes == 0) return GS_OK;
    // pass 0 writes the alt buffers, pass 1 the primary ones, ...
    int rc = sort_impl<long long>((const long long *)keys, payload, (long long *)keys_alt, payload_alt,
                                  (long long *)keys, payload, k, spec, depth_bits + tile_bits, workspace,
                                  (cudaStream_t)stream);
    if (rc != GS_OK) return rc;
    if (result_in_alt) *result_in_alt = (num_passes & 1);
    return GS_OK;
}

extern "C" int gs_sort_pairs32(uint32_t *keys, int32_t *payload, uint32_t *keys_alt,
                               int32_t *payload_alt, int64_t k, int key_bits, void *workspace,
                               const uint32_t *digit_hist, int *result_in_alt, void *stream)
{
    if (result_in_alt) *result_in_alt = 0;
    if (k < 0 || key_bits < 0 || key_bits > 32) return GS_E_INVALID;
    if (k >= (1ll << 30)) return GS_E_CAPACITY;
    if (k <= 1 || key_bits == 0) return GS_OK;
    if (!keys || !payload || !keys_alt || !payload_alt || !workspace) return GS_E_INVALID;
    KeySpec spec = {32, 0xffffffffull, 0ull};
    int num_passes = (key_bits + 7) / 8;
    int rc = sort_impl<unsigned int>(keys, payload, keys_alt, payload_alt, keys, payload, k, spec,
                                     key_bits, workspace, (cudaStream_t)stream, nullptr, digit_hist);
    if (rc != GS_OK) return rc;
    if (result_in_alt) *result_in_alt = (num_passes & 1);
    return GS_OK;
}

extern "C" int gs_sort_points_by_depth(const int32_t *depth_key, int64_t m, const int64_t *m_dev,
                                       int depth_bits, int32_t *order, void *workspace, void *stream)
{
    if (m < 0 || depth_bits < 1 || depth_bits > 31) return GS_E_INVALID;
    if (m >= (1ll << 30)) return GS_E_CAPACITY;
    if (m == 0) return GS_OK;
    if (!depth_key || !order || !workspace) return GS_E_INVALID;
    unsigned char *ws = (unsigned char *)workspace;
    const int64_t scratch = align256(sort_scratch_bytes(m)), col = align256(4 * m);
    unsigned int *keys_a = (unsigned int *)(ws + scratch);
    unsigned int *keys_b = (unsigned int *)(ws + scratch + col);
    int32_t *vals_tmp = (int32_t *)(ws + scratch + 2 * col);
    KeySpec spec = {32, 0xffffffffull, 0ull};
    const int num_passes = (depth_bits + 7) / 8;
    // the last pass must land in `order`: pass p writes the a buffers when p is even
    int32_t *vals_a = (num_passes & 1) ? order : vals_tmp;
    int32_t *vals_b = (num_passes & 1) ? vals_tmp : order;
    return sort_impl<unsigned int>((const unsigned int *)depth_key, nullptr, keys_a, vals_a, keys_b,
                                   vals_b, m, spec, depth_bits, workspace, (cudaStream_t)stream,
                                   (const long long *)m_dev);
}
