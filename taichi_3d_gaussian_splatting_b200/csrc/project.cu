// project.cu -- pose inverse and the fused cull / compaction / projection / SH / tile-count pass.
//
// Literal-fp32 profile: build this file with -fmad=false.  Every expression below follows the
// operation order of oracle/gs_oracle.c (which follows the reference's Python source), with the
// structural zeros of J and S skipped (x + 0*y == x).
//
// Replaces: inverse_SE3_qt_torch (reference utils.py:426-432), filter_point_in_camera
// (R.py:31-78), mask compaction (R.py:861-865), generate_point_attributes_in_camera_plane
// (R.py:239-315), generate_num_overlap_tiles (R.py:106-128), cumsum/cat (R.py:913-922).
#include "gs_common.cuh"

namespace {

// points per CTA.  The staged feature rows (224 B per point) bound the CTAs per SM: 128-point CTAs keep 28
// warps resident against 24 with 256, and a CTA-wide barrier holds up four warps instead of eight
// (C2 0.160 -> 0.155 ms, C5 0.541 -> 0.516 ms for pose inverse + projection)
#ifndef GS_PROJ_BLOCK
#define GS_PROJ_BLOCK 128
#endif
constexpr int kProjBlock = GS_PROJ_BLOCK;
// GS_PROJ_WALKER: 1 = a fifth warp without points does the CTA's look-back over the preceding blocks, so
// that the walk runs beside the second half of the four point warps' arithmetic instead of after warp 0's
#ifndef GS_PROJ_WALKER
#define GS_PROJ_WALKER 0
#endif
constexpr int kProjThreads = kProjBlock + (GS_PROJ_WALKER ? 32 : 0);
constexpr int kWalkWarp = GS_PROJ_WALKER ? kProjBlock / 32 : 0;
constexpr unsigned long long kFlagAggregate = 1ull << 62;
constexpr unsigned long long kFlagPrefix = 2ull << 62;
constexpr unsigned long long kValueMask = (1ull << 62) - 1;
// Look-back word.  Default: the number of visible points alone (n < 2^30 cannot overflow 62 bits; the
// total tile count K is summed separately, 64-bit, with one atomic per block).  With per-point tile
// OFFSETS requested (the literal key pipeline), the word packs (visible << 35 | tiles): 27 bits of
// points and 35 bits of tiles, and the entry point refuses inputs whose worst case n * tiles-per-image
// does not fit (GS_E_CAPACITY) instead of letting the tile sum carry into the point count.
constexpr int kPackShift = 35;
constexpr unsigned long long kPackTileMask = (1ull << kPackShift) - 1;

__global__ void inverse_pose_kernel(const float *__restrict__ q_pc_cam,
                                    const float *__restrict__ t_pc_cam, int n_obj,
                                    float *__restrict__ q_cam_pc, float *__restrict__ t_cam_pc)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_obj) return;
    const float *q = q_pc_cam + 4 * k, *t = t_pc_cam + 3 * k;
    float qi0 = -q[0], qi1 = -q[1], qi2 = -q[2], qi3 = q[3];
    float n = sqrtf(((qi0 * qi0 + qi1 * qi1) + qi2 * qi2) + qi3 * qi3);
    float x0 = qi0 / n, y0 = qi1 / n, z0 = qi2 / n, w0 = qi3 / n;
    // (qn * [t,0]) * conj(qn), reference utils.py:402-423
    float x1 = t[0], y1 = t[1], z1 = t[2], w1 = 0.0f;
    float ax = w0 * x1 + x0 * w1 + y0 * z1 - z0 * y1;
    float ay = w0 * y1 - x0 * z1 + y0 * w1 + z0 * x1;
    float az = w0 * z1 + x0 * y1 - y0 * x1 + z0 * w1;
    float aw = w0 * w1 - x0 * x1 - y0 * y1 - z0 * z1;
    float cx = -x0, cy = -y0, cz = -z0, cw = w0;
    float rx = aw * cx + ax * cw + ay * cz - az * cy;
    float ry = aw * cy - ax * cz + ay * cw + az * cx;
    float rz = aw * cz + ax * cy - ay * cx + az * cw;
    q_cam_pc[4 * k + 0] = qi0; q_cam_pc[4 * k + 1] = qi1;
    q_cam_pc[4 * k + 2] = qi2; q_cam_pc[4 * k + 3] = qi3;
    t_cam_pc[3 * k + 0] = -rx; t_cam_pc[3 * k + 1] = -ry; t_cam_pc[3 * k + 2] = -rz;
}

struct ProjectParams {
    const float *xyz;
    float *features;
    const int8_t *invalid_mask;
    const int32_t *object_id;
    long long n;
    const float *intrinsics;
    const float *q_cam_pc;
    const float *t_cam_pc;
    float near_plane, far_plane;
    int width, height;
    float key_scale;
    int32_t *ids;
    float *point_uv;
    float *point_in_camera;
    float *conic;
    float *alpha;
    float *color;
    float *radii;
    int32_t *num_tiles;
    long long *tile_offsets;
    float *records;
    uint16_t *bbox;
    int32_t *depth_key;
    int32_t *slot_of_point;  // nullable: [n] position of point i in the visible list, or -1
    long long *totals;
    unsigned long long *status;  // one word per block of 256 points, zero on entry
    unsigned int *ticket;        // zero on entry
    int num_blocks;
};

__device__ __forceinline__ float dot16(const float *f, const float *b)
{
    float s = f[0] * b[0];
#pragma unroll
    for (int i = 1; i < 16; ++i) s = s + f[i] * b[i];
    return s;
}

// ---- bulk (TMA) staging of the feature rows ---------------------------------------------------
// The 32 points of a warp own 32 consecutive 224-byte rows = one contiguous 7168-byte run of
// point_cloud_features: one elected lane fetches it with a single cp.async.bulk (UBLKCP) into
// shared memory, completion tracked by an mbarrier, while the warp does the cull arithmetic.
constexpr int kRowBytes = GS_FEATURES * 4;          // 224
constexpr int kWarpRowBytes = 32 * kRowBytes;       // 7168
constexpr int kProjWarps = kProjBlock / 32;
constexpr size_t kProjSmem = (size_t)kProjWarps * kWarpRowBytes + kProjWarps * 8;

__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void *dst_smem, const void *src, unsigned bytes,
                                          unsigned long long *bar)
{
    unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d),
        "l"(src), "r"(bytes), "r"(b)
        : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    unsigned b = (unsigned)__cvta_generic_to_shared(bar);
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

// Feature rows come through shared memory by cp.async.bulk (the features pointer must be 16-byte
// aligned; the entry point checks).  kOffsets: also produce the exclusive scan of the tile counts
// (tile_offsets), which only the literal key pipeline reads.
template <bool kOffsets>
__global__ void __launch_bounds__(kProjThreads) project_cull_sh_kernel(ProjectParams p)
{
    constexpr bool kBulk = true;
    extern __shared__ __align__(128) unsigned char s_dyn[];
    __shared__ unsigned int s_ticket;
    __shared__ unsigned long long s_warp_total[kProjBlock / 32];
    __shared__ unsigned long long s_warp_tiles[kProjBlock / 32];
    __shared__ unsigned long long s_block_excl;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_ticket = atomicAdd(p.ticket, 1u);
    __syncthreads();
    const int block = (int)s_ticket;
    const bool walker = GS_PROJ_WALKER && warp == kWalkWarp;   // owns no points: its lanes read as "past the end"
    const long long i = walker ? p.n + lane : (long long)block * kProjBlock + tid;
    const int srow = walker ? 0 : warp;
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(s_dyn + (size_t)kProjWarps * kWarpRowBytes) + srow;
    const float4 *s_row = reinterpret_cast<const float4 *>(s_dyn + (size_t)srow * kWarpRowBytes) + 14 * lane;
    if (kBulk && !walker) {
        const long long i0 = i - lane;
        const long long rows = min(32ll, p.n - i0);
        if (lane == 0) {
            mbar_init(bar, 1);
            if (rows > 0)
                bulk_load(s_dyn + (size_t)warp * kWarpRowBytes, p.features + (size_t)GS_FEATURES * i0,
                          (unsigned)(rows * kRowBytes), bar);
        }
        __syncwarp();
    }

    // ---- cull (R.py:50-78) -------------------------------------------------------------
    bool visible = false;
    float W[9], tx = 0.f, ty = 0.f, tz = 0.f;   // rotation / translation of T_camera_pointcloud
    float K[9];
    float px = 0.f, py = 0.f, pz = 0.f;          // point in pointcloud frame
    float cx = 0.f, cy = 0.f, cz = 0.f, u = 0.f, v = 0.f;
    int obj = 0;
#pragma unroll
    for (int k = 0; k < 9; ++k) K[k] = __ldg(p.intrinsics + k);
    if (i < p.n && p.invalid_mask[i] != 1) {
        px = __ldg(p.xyz + 3 * i); py = __ldg(p.xyz + 3 * i + 1); pz = __ldg(p.xyz + 3 * i + 2);
        obj = __ldg(p.object_id + i);
        const float *q = p.q_cam_pc + 4 * obj, *t = p.t_cam_pc + 3 * obj;
        gs_rotation_from_quaternion(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3), W);
        tx = __ldg(t); ty = __ldg(t + 1); tz = __ldg(t + 2);
        cx = ((W[0] * px + W[1] * py) + W[2] * pz) + tx;
        cy = ((W[3] * px + W[4] * py) + W[5] * pz) + ty;
        cz = ((W[6] * px + W[7] * py) + W[8] * pz) + tz;
        float k0 = (K[0] * cx + K[1] * cy) + K[2] * cz;
        float k1 = (K[3] * cx + K[4] * cy) + K[5] * cz;
        u = k0 / cz;
        v = k1 / cz;
        const float margin = (float)(kTile * GS_BOUNDARY_TILES);
        visible = cz > p.near_plane && cz < p.far_plane && u >= -margin &&
                  u < (float)(p.width + kTile * GS_BOUNDARY_TILES) && v >= -margin &&
                  v < (float)(p.height + kTile * GS_BOUNDARY_TILES);
    }

    // ---- per visible point attributes (R.py:260-315) -----------------------------------
    float conic0 = 0.f, conic1 = 0.f, conic2 = 0.f, rescale = 0.f, alpha = 0.f;
    float col0 = 0.f, col1 = 0.f, col2 = 0.f, radius = 0.f;
    int tu0 = 0, tu1 = 0, tv0 = 0, tv1 = 0, count = 0;
    if (kBulk && i - lane < p.n) mbar_wait(bar, 0);  // every warp that issued a copy waits for it
    // Phase 1: what the tile count needs (covariance -> radius -> bounding box).  The block's
    // aggregate is published right after it, so the look-back of this and of later blocks overlaps
    // phase 2 (conic, opacity, SH colour) instead of stalling the block at a barrier.
    float4 *grow = reinterpret_cast<float4 *>(p.features + (size_t)GS_FEATURES * i);
    const float4 *row = kBulk ? s_row : grow;
    float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float c00 = 0.f, c01 = 0.f, c10 = 0.f, c11 = 0.f;
    if (visible) {
        float4 q4 = row[0];
        s4 = row[1];
        // normalise the quaternion in place (R.py:196-205)
        float n2 = ((q4.x * q4.x + q4.y * q4.y) + q4.z * q4.z) + q4.w * q4.w;
        float invlen = 1.0f / sqrtf(n2);
        q4.x = invlen * q4.x; q4.y = invlen * q4.y; q4.z = invlen * q4.z; q4.w = invlen * q4.w;
        grow[0] = q4;

        // Sigma' = J W Sigma W^T J^T (reference GaussianPoint3D.py:162-191)
        float fx = K[0], fy = K[4];
        float J0 = fx / cz, J2 = -(fx * cx) / (cz * cz);
        float J4 = fy / cz, J5 = -(fy * cy) / (cz * cz);
        float R[9];
        gs_rotation_from_quaternion(q4.x, q4.y, q4.z, q4.w, R);
        float es[3] = {gs_expf(s4.x), gs_expf(s4.y), gs_expf(s4.z)};
        float B[9], Sg[9];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) B[3 * r + c] = (R[3 * r + c] * es[c]) * es[c];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                Sg[3 * r + c] = (B[3 * r] * R[3 * c] + B[3 * r + 1] * R[3 * c + 1]) +
                                B[3 * r + 2] * R[3 * c + 2];
        float U[6], V[6], X[6];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            U[c] = J0 * W[c] + J2 * W[6 + c];
            U[3 + c] = J4 * W[3 + c] + J5 * W[6 + c];
        }
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                V[3 * r + c] = (U[3 * r] * Sg[c] + U[3 * r + 1] * Sg[3 + c]) + U[3 * r + 2] * Sg[6 + c];
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                X[3 * r + c] = (V[3 * r] * W[3 * c] + V[3 * r + 1] * W[3 * c + 1]) +
                               V[3 * r + 2] * W[3 * c + 2];
        c00 = X[0] * J0 + X[2] * J2;
        c01 = X[1] * J4 + X[2] * J5;
        c10 = X[3] * J0 + X[5] * J2;
        c11 = X[4] * J4 + X[5] * J5;

        // radius from the unfiltered covariance (R.py:311-315)
        float lam = (c00 + c11 + sqrtf((c00 - c11) * (c00 - c11) + 4.0f * c01 * c10)) / 2.0f;
        radius = sqrtf(lam) * 3.0f;

        // tile bounding box (R.py:81-103); non-finite splats get an empty box
        if (isfinite(u) && isfinite(v) && isfinite(radius)) {
            float r = fmaxf(radius, 1.0f);
            float min_u = fmaxf(0.0f, u - r), max_u = u + r;
            float min_v = fmaxf(0.0f, v - r), max_v = v + r;
            int tw = p.width / kTile, th = p.height / kTile;
            const float cap = 1.0e9f;
            tu0 = min((int)fminf(floorf(min_u / kTile), cap), tw);
            tu1 = min(max((int)fminf(floorf(max_u / kTile), cap) + 1, tu0 + 1), tw);
            tv0 = min((int)fminf(floorf(min_v / kTile), cap), th);
            tv1 = min(max((int)fminf(floorf(max_v / kTile), cap) + 1, tv0 + 1), th);
        }
        count = (tu1 - tu0) * (tv1 - tv0);
    }


    // ---- ordered compaction, part 1: block scan, publish the block aggregate ----------------
    unsigned long long mine, incl;
    if (kOffsets) {
        mine = ((unsigned long long)(visible ? 1u : 0u) << kPackShift) | (unsigned int)count;
        incl = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31 && !walker) s_warp_total[warp] = incl;
    } else {
        // rank among the warp's visible points by ballot; the tiles are only summed
        const unsigned ballot = __ballot_sync(0xffffffffu, visible);
        mine = visible ? 1ull : 0ull;
        incl = (unsigned long long)__popc(ballot & (0xffffffffu >> (31 - lane)));
        const unsigned tiles = __reduce_add_sync(0xffffffffu, (unsigned)count);
        if (lane == 31 && !walker) {
            s_warp_total[warp] = incl;
            s_warp_tiles[warp] = tiles;
        }
    }
    __syncthreads();
    unsigned long long warp_excl = 0, block_total = 0, block_tiles = 0;
#pragma unroll
    for (int w = 0; w < kProjBlock / 32; ++w) {
        unsigned long long t = s_warp_total[w];
        if (w < warp) warp_excl += t;
        block_total += t;
        if (!kOffsets) block_tiles += s_warp_tiles[w];
    }
    if (warp == kWalkWarp && lane == 0) {
        gs_st_volatile_u64(p.status + block, (block == 0 ? kFlagPrefix : kFlagAggregate) | block_total);
        if (!kOffsets && block_tiles) atomicAdd(reinterpret_cast<unsigned long long *>(p.totals) + 1, block_tiles);
    }

    // Phase 2: the rest of the per-point attributes (R.py:286-310)
    if (visible) {
        // conic + rescale on a by-value copy (reference utils.py:257-272)
        {
            float det0 = c00 * c11 - c01 * c10;
            float a = c00 + (float)0.3, d = c11 + (float)0.3;
            float det = a * d - c01 * c10;
            rescale = sqrtf(fmaxf(0.0f, det0 / det));
            float id = 1.0f / det;
            conic0 = id * d;
            conic1 = id * (-c01);
            conic2 = id * a;
        }
        alpha = 1.0f / (1.0f + gs_expf(-s4.w));

        // SH colour on the ray from the camera centre (R.py:280-310)
        float ox = ((-W[0]) * tx + (-W[3]) * ty) + (-W[6]) * tz;
        float oy = ((-W[1]) * tx + (-W[4]) * ty) + (-W[7]) * tz;
        float oz = ((-W[2]) * tx + (-W[5]) * ty) + (-W[8]) * tz;
        float basis[16];
        gs_sh_basis(px - ox, py - oy, pz - oz, basis);
        float coef[16];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                float4 t4 = row[2 + 4 * ch + k4];
                coef[4 * k4] = t4.x; coef[4 * k4 + 1] = t4.y;
                coef[4 * k4 + 2] = t4.z; coef[4 * k4 + 3] = t4.w;
            }
            float c = gs_sigmoid(dot16(coef, basis));
            if (ch == 0) col0 = c; else if (ch == 1) col1 = c; else col2 = c;
        }

    }

    // ---- ordered compaction, part 2: decoupled look-back over blocks -------------------------
    if (warp == kWalkWarp) {
        unsigned long long excl = 0;
        if (block > 0) {
            int base = block - 1;
            while (true) {
                int idx = base - lane;
                unsigned long long s = kFlagPrefix;  // positions before block 0: empty prefix
                if (idx >= 0) {
                    do { s = gs_ld_volatile_u64(p.status + idx); } while ((s >> 62) == 0);
                }
                unsigned int pmask = __ballot_sync(0xffffffffu, (s >> 62) == 2);
                int first = pmask ? (__ffs(pmask) - 1) : 32;
                unsigned long long contrib = (lane <= first) ? (s & kValueMask) : 0ull;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
                excl += contrib;
                if (pmask) break;
                base -= 32;
            }
            if (lane == 0) gs_st_volatile_u64(p.status + block, kFlagPrefix | (excl + block_total));
        }
        if (lane == 0) {
            s_block_excl = excl;
            if (block == p.num_blocks - 1) {
                unsigned long long tot = excl + block_total;
                if (kOffsets) {
                    p.totals[0] = (long long)(tot >> kPackShift);
                    p.totals[1] = (long long)(tot & kPackTileMask);
                } else {
                    p.totals[0] = (long long)tot;  // totals[1] = K: summed by the blocks' atomics
                }
            }
        }
    }
    __syncthreads();
    if (!visible) {
        if (p.slot_of_point && i < p.n) p.slot_of_point[i] = -1;
        return;
    }

    unsigned long long my_excl = s_block_excl + warp_excl + (incl - mine);
    const long long idx = kOffsets ? (long long)(my_excl >> kPackShift) : (long long)my_excl;

    // what the path itself reads: ids, camera-space position, the packed record, bbox, depth key
    p.ids[idx] = (int32_t)i;
    if (p.slot_of_point) p.slot_of_point[i] = (int32_t)idx;
    p.point_in_camera[3 * idx] = cx; p.point_in_camera[3 * idx + 1] = cy;
    p.point_in_camera[3 * idx + 2] = cz;
    // the reference-named per-point buffers (R.py:873-884), for callers that look at them (the
    // backward hook's payload: uv and tile counts; tests: all of them)
    if (p.point_uv) reinterpret_cast<float2 *>(p.point_uv)[idx] = make_float2(u, v);
    if (p.conic) reinterpret_cast<float4 *>(p.conic)[idx] = make_float4(conic0, conic1, conic2, rescale);
    if (p.alpha) p.alpha[idx] = alpha;
    if (p.color) { p.color[3 * idx] = col0; p.color[3 * idx + 1] = col1; p.color[3 * idx + 2] = col2; }
    if (p.radii) p.radii[idx] = radius;
    if (p.num_tiles) p.num_tiles[idx] = count;
    if (kOffsets) p.tile_offsets[idx] = (long long)(my_excl & kPackTileMask);
    float4 *rec = reinterpret_cast<float4 *>(p.records + (size_t)kRecFloats * idx);
    rec[0] = make_float4(u, v, conic0, conic1);
    rec[1] = make_float4(conic2, rescale, alpha, col0);
    rec[2] = make_float4(col1, col2, cz, gs_decision_band_log2(conic0, conic1, conic2, rescale * alpha));
    reinterpret_cast<ushort4 *>(p.bbox)[idx] =
        make_ushort4((unsigned short)tu0, (unsigned short)tu1, (unsigned short)tv0, (unsigned short)tv1);
    p.depth_key[idx] = (int32_t)(cz * p.key_scale);
}

}  // namespace

extern "C" int64_t gs_project_workspace_bytes(int64_t n)
{
    int64_t blocks = (n + kProjBlock - 1) / kProjBlock;
    return (blocks + 2) * 8;
}

extern "C" int gs_inverse_pose(const float *q_pc_cam, const float *t_pc_cam, int n_obj,
                               float *q_cam_pc, float *t_cam_pc, void *stream)
{
    if (!q_pc_cam || !t_pc_cam || !q_cam_pc || !t_cam_pc || n_obj <= 0) return GS_E_INVALID;
    inverse_pose_kernel<<<(n_obj + 63) / 64, 64, 0, (cudaStream_t)stream>>>(q_pc_cam, t_pc_cam, n_obj,
                                                                          q_cam_pc, t_cam_pc);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int gs_project_cull_sh(const float *xyz, float *features, const int8_t *invalid_mask,
                                  const int32_t *object_id, int64_t n, const float *intrinsics,
                                  const float *q_cam_pc, const float *t_cam_pc, int n_obj,
                                  float near_plane, float far_plane, int width, int height,
                                  float depth_to_sort_key_scale, int32_t *ids, float *point_uv,
                                  float *point_in_camera, float *conic, float *alpha, float *color,
                                  float *radii, int32_t *num_tiles, int64_t *tile_offsets,
                                  float *records, uint16_t *bbox, int32_t *depth_key,
                                  int32_t *slot_of_point, int64_t *totals, void *workspace, void *stream)
{
    if (n < 0 || width <= 0 || height <= 0 || width % kTile || height % kTile || n_obj <= 0)
        return GS_E_INVALID;
    if (n >= (1ll << 30)) return GS_E_CAPACITY;
    if (!totals || !workspace) return GS_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    GS_CUDA_TRY(cudaMemsetAsync(totals, 0, 16, st));
    if (n == 0) return GS_OK;
    if (!xyz || !features || !invalid_mask || !object_id || !intrinsics || !q_cam_pc || !t_cam_pc ||
        !ids || !point_in_camera || !records || !bbox || !depth_key)
        return GS_E_INVALID;
    // the feature rows are fetched with cp.async.bulk and written back as 16-byte vectors
    if (((uintptr_t)features % 16) != 0 || ((uintptr_t)records % 16) != 0) return GS_E_INVALID;
    const bool offsets = tile_offsets != nullptr;
    if (offsets) {
        // worst case of the packed look-back word: every point covers the whole image
        const long long tiles_per_image = (long long)(width / kTile) * (height / kTile);
        if (n >= (1ll << (62 - kPackShift)) || n * tiles_per_image > (long long)kPackTileMask)
            return GS_E_CAPACITY;
    }
    ProjectParams p;
    p.xyz = xyz; p.features = features; p.invalid_mask = invalid_mask; p.object_id = object_id;
    p.n = n; p.intrinsics = intrinsics; p.q_cam_pc = q_cam_pc; p.t_cam_pc = t_cam_pc;
    p.near_plane = near_plane; p.far_plane = far_plane; p.width = width; p.height = height;
    p.key_scale = depth_to_sort_key_scale;
    p.ids = ids; p.point_uv = point_uv; p.point_in_camera = point_in_camera; p.conic = conic;
    p.alpha = alpha; p.color = color; p.radii = radii; p.num_tiles = num_tiles;
    p.tile_offsets = (long long *)tile_offsets; p.records = records; p.bbox = bbox;
    p.depth_key = depth_key; p.slot_of_point = slot_of_point; p.totals = (long long *)totals;
    int blocks = (int)((n + kProjBlock - 1) / kProjBlock);
    p.num_blocks = blocks;
    p.status = (unsigned long long *)workspace;
    p.ticket = (unsigned int *)((unsigned long long *)workspace + blocks);
    GS_CUDA_TRY(cudaMemsetAsync(workspace, 0, (size_t)gs_project_workspace_bytes(n), st));
    static GsPerDeviceOnce attr_set;
    int dev = 0;
    if (attr_set.need(dev)) {
        GS_CUDA_TRY(cudaFuncSetAttribute(project_cull_sh_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kProjSmem));
        GS_CUDA_TRY(cudaFuncSetAttribute(project_cull_sh_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kProjSmem));
        attr_set.mark(dev);
    }
    if (offsets)
        project_cull_sh_kernel<true><<<blocks, kProjThreads, kProjSmem, st>>>(p);
    else
        project_cull_sh_kernel<false><<<blocks, kProjThreads, kProjSmem, st>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

// ---- no-allocation variant ---------------------------------------------------------------------------
// gs_project_cull_sh_static: the same pass with its look-back scratch in a file-scope device array
// instead of a caller-allocated workspace, for callers that cannot allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the array (1 MiB) until exit.  The
// array's name is a host-side stand-in in host code, not a device address: the launcher takes its
// address with cudaGetSymbolAddress once per device, with that device current.  All calls on a device
// share the array: gs_project_cull_sh clears the words a call uses before its kernel (stream-ordered
// memset), so nothing of the previous call is read, but two calls that OVERLAP on the GPU (different
// streams, no dependency) race on it.
static constexpr long long kStaticMaxPoints = GS_PROJECT_STATIC_MAX_POINTS;
static constexpr long long kStaticWords = (kStaticMaxPoints + kProjBlock - 1) / kProjBlock + 2;
static __device__ __align__(256) unsigned long long g_project_static_scratch[kStaticWords];

extern "C" int gs_project_cull_sh_static(const float *xyz, float *features, const int8_t *invalid_mask,
                                         const int32_t *object_id, int64_t n, const float *intrinsics,
                                         const float *q_cam_pc, const float *t_cam_pc, int n_obj,
                                         float near_plane, float far_plane, int width, int height,
                                         float depth_to_sort_key_scale, int32_t *ids, float *point_uv,
                                         float *point_in_camera, float *conic, float *alpha, float *color,
                                         float *radii, int32_t *num_tiles, int64_t *tile_offsets,
                                         float *records, uint16_t *bbox, int32_t *depth_key,
                                         int32_t *slot_of_point, int64_t *totals, void *stream)
{
    if (n > kStaticMaxPoints) return GS_E_STATIC_LIMIT;
    static_assert(sizeof(g_project_static_scratch) >= (size_t)(kStaticWords * 8), "scratch size");
    static void *address[64];
    static GsPerDeviceOnce resolved;
    int dev = 0;
    if (resolved.need(dev)) {
        if (dev < 0 || dev >= 64) return GS_E_INVALID;
        void *ptr = nullptr;
        GS_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_project_static_scratch));  // of the CURRENT device
        if (!ptr) return GS_E_INVALID;
        address[dev] = ptr;
        resolved.mark(dev);
    }
    return gs_project_cull_sh(xyz, features, invalid_mask, object_id, n, intrinsics, q_cam_pc, t_cam_pc,
                              n_obj, near_plane, far_plane, width, height, depth_to_sort_key_scale, ids,
                              point_uv, point_in_camera, conic, alpha, color, radii, num_tiles,
                              tile_offsets, records, bbox, depth_key, slot_of_point, totals, address[dev],
                              stream);
}
