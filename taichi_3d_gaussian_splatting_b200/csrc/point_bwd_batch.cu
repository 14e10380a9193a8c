// point_bwd_batch.cu -- the per-point chain rule of a whole BATCH of views in one pass over the points.
//
// The view-sharded step renders V views against one cloud and wants the SUM of
// their point gradients.  gs_point_bwd in accumulate mode does that view by view: every view reads the
// 224-byte feature row of each of its visible points again and adds its 236-byte gradient row into the
// N x 59 buffer with RED (a read-modify-write in L2 / HBM), and the buffer needs a zero fill per step:
// per step 4 x 59 N + V x (280 + 2 x 236) M bytes.  Here ONE thread owns a point for the whole batch:
//
//   * its feature row is fetched once (cp.async into shared memory, as in point_bwd.cu);
//   * for every view in which the point is visible (slot_of_point[v][p] >= 0, written by
//     gs_project_cull_sh) it reads the view's 48-byte reduced splat gradients and its camera-space
//     position, and accumulates what depends on the view IN REGISTERS: d xyz (3), the symmetric
//     dL/dSigma = U^T G U (6 -- the chain from dL/dSigma to the quaternion and the scales is linear in
//     it and does not depend on the view, so it runs once per point, not once per view), the opacity
//     logit (1) and the 48 SH coefficients (the basis depends on the view direction);
//   * the finished 236-byte row is STORED, once, for every point of the range whether it was visible
//     or not: no zero fill of the buffer, no RED.
//
// Traffic per step: (224 + 236 + 4 V) N + 60 V M bytes (C2, V = 8: 0.93 GB against 4.1 GB).  A call
// covers a range of point ids, so that a step can be cut into chunks whose results are exchanged
// (all-reduce over the ranks) while the next chunk is computed.
//
// Same arithmetic as point_bwd.cu (fast-fp32 profile; reference R.py:707-772, 1102-1125, 1167-1182,
// GaussianPoint3D.py:132-159, 237-331, 351-373); the sums over views are formed in view order.
#include "gs_common.cuh"

namespace {

constexpr int kBatchBlock = 128;
constexpr int kRowStride4 = 15;  // float4 per staged row: 14 used + 1 pad (conflict-free 16-byte lanes)

struct BatchParams {
    const float *xyz;
    const float *features;
    const int32_t *object_id;
    long long first, count;       // point ids [first, first + count)
    const gs_view_t *views;
    int num_views;
    int keep;
    float q_factor, s_factor, alpha_factor, color_factor, high_order_color_factor;
    float *grad_xyz;
    float *grad_features;
    gs_view_t one;      // inline_one != 0: the batch is this one view (no device table)
    int inline_one;
};

typedef float4 BatchRows[32 * kRowStride4];

// one block of 128 consecutive points (block index `block` of the range), all views
__device__ __forceinline__ void batch_block(const BatchParams &p, BatchRows *s_rows, const long long block)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long local = block * kBatchBlock + threadIdx.x;
    const bool valid = local < p.count;
    const long long pid = p.first + local;
    float4 *rows = s_rows[warp];
    const int half = lane >> 4, col = lane & 15;
    const long long pid0 = pid - lane;  // the warp's 32 points are consecutive ids

    // the warp's 32 feature rows: 448 asynchronous 16-byte copies, all in flight at once
#pragma unroll 8
    for (int r = 0; r < 32; r += 2) {
        const long long rp = pid0 + r + half;
        if (col < 14 && rp < p.first + p.count) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(rows + (r + half) * kRowStride4 + col);
            const float4 *src = reinterpret_cast<const float4 *>(p.features + (size_t)GS_FEATURES * rp) + col;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");

    float px = 0.f, py = 0.f, pz = 0.f;
    int obj = 0;
    if (valid) {
        px = __ldg(p.xyz + 3 * pid); py = __ldg(p.xyz + 3 * pid + 1); pz = __ldg(p.xyz + 3 * pid + 2);
        obj = __ldg(p.object_id + pid);
    }
    float gxyz[3] = {0.f, 0.f, 0.f};
    float A[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // dL/dSigma: 00 01 02 11 12 22
    float glogit = 0.f;
    float sh[48];
#pragma unroll
    for (int k = 0; k < 48; ++k) sh[k] = 0.f;

    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    float4 *row = rows + lane * kRowStride4;

    // the slot of the NEXT view is fetched while this view is worked on: an iteration then waits for the
    // view's 60 bytes only, not for the dependent pair slot -> gradients
    int slot_next = -1;
    if (valid && p.num_views > 0) slot_next = __ldg((p.inline_one ? p.one : p.views[0]).slot_of_point + pid);
    for (int v = 0; v < p.num_views; ++v) {
        const gs_view_t &view = p.inline_one ? p.one : p.views[v];
        const int slot = slot_next;
        slot_next = -1;
        if (valid && v + 1 < p.num_views) slot_next = __ldg(p.views[v + 1].slot_of_point + pid);
        if (slot < 0) continue;
        const float4 *sg4 = reinterpret_cast<const float4 *>(view.splat_grads + (size_t)kGradSlots * slot);
        const float4 ga = __ldg(sg4), gb = __ldg(sg4 + 1), gc = __ldg(sg4 + 2);
        const float gu = ga.x, gv = ga.y, c00 = ga.z, c01 = ga.w, c11 = gb.x;
        const float gcol[3] = {gb.y, gb.z, gb.w};
        glogit += gc.x;
        float K[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) K[k] = __ldg(view.intrinsics + k);
        float W[9];
        {
            const float *q = view.q_cam_pc + 4 * obj;
            gs_rotation_from_quaternion(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3), W);
        }
        const float *pic = view.point_in_camera + 3 * (size_t)slot;
        const float cx = __ldg(pic), cy = __ldg(pic + 1), cz = __ldg(pic + 2);
        const float iz = 1.0f / cz, iz2 = iz * iz;
        // d uv / d xyz = (d uv / d p_cam) W   (GaussianPoint3D.py:149-159, full K)
        const float d[6] = {K[0] * iz, K[1] * iz, (-K[0] * cx - K[1] * cy) * iz2,
                            K[3] * iz, K[4] * iz, (-K[3] * cx - K[4] * cy) * iz2};
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const float d0 = d[0] * W[j] + d[1] * W[3 + j] + d[2] * W[6 + j];
            const float d1 = d[3] * W[j] + d[4] * W[3 + j] + d[5] * W[6 + j];
            gxyz[j] += gu * d0 + gv * d1;
        }
        // U = J W;  dL/dSigma += U^T G U
        const float J0 = K[0] * iz, J2 = -(K[0] * cx) * iz2, J4 = K[4] * iz, J5 = -(K[4] * cy) * iz2;
        float U[6], GU[6];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            U[j] = J0 * W[j] + J2 * W[6 + j];
            U[3 + j] = J4 * W[3 + j] + J5 * W[6 + j];
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            GU[j] = c00 * U[j] + c01 * U[3 + j];
            GU[3 + j] = c01 * U[j] + c11 * U[3 + j];
        }
        A[0] += U[0] * GU[0] + U[3] * GU[3];
        A[1] += U[0] * GU[1] + U[3] * GU[4];
        A[2] += U[0] * GU[2] + U[3] * GU[5];
        A[3] += U[1] * GU[1] + U[4] * GU[4];
        A[4] += U[1] * GU[2] + U[4] * GU[5];
        A[5] += U[2] * GU[2] + U[5] * GU[5];
        // SH colour Jacobian: sigmoid'(raw) * basis (GaussianPoint3D.py:351-373); the ray starts at the
        // input translation t_pointcloud_camera (R.py:731-732,749)
        const float *org = view.t_pc_cam + 3 * obj;
        float basis[16];
        gs_sh_basis(px - __ldg(org), py - __ldg(org + 1), pz - __ldg(org + 2), basis);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            float raw = 0.f;
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                const float4 c = row[2 + 4 * ch + k4];
                raw += c.x * basis[4 * k4] + c.y * basis[4 * k4 + 1] + c.z * basis[4 * k4 + 2] +
                       c.w * basis[4 * k4 + 3];
            }
            const float sg = 1.0f / (1.0f + __expf(-raw));
            const float scale = gcol[ch] * sg * (1.0f - sg);
#pragma unroll
            for (int k = 0; k < 16; ++k) sh[16 * ch + k] = fmaf(scale, basis[k], sh[16 * ch + k]);
        }
    }

    if (valid) {
        // once per point: dL/dM = 2 (dL/dSigma) M, then to the scales and the (normalised) quaternion
        const float4 q4 = row[0], s4 = row[1];
        const float qx = q4.x, qy = q4.y, qz = q4.z, qw = q4.w;
        float R[9];
        gs_rotation_from_quaternion(qx, qy, qz, qw, R);
        const float es[3] = {__expf(s4.x), __expf(s4.y), __expf(s4.z)};
        const float Af[9] = {A[0], A[1], A[2], A[1], A[3], A[4], A[2], A[4], A[5]};
        float D[9];
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int l = 0; l < 3; ++l) {
                const float am = Af[3 * k] * R[l] + Af[3 * k + 1] * R[3 + l] + Af[3 * k + 2] * R[6 + l];
                D[3 * k + l] = 2.0f * am * es[l] * es[l];
            }
        float gs3[3];
#pragma unroll
        for (int l = 0; l < 3; ++l) gs3[l] = D[l] * R[l] + D[3 + l] * R[3 + l] + D[6 + l] * R[6 + l];
        float gq[4];
        gq[0] = 2.0f * (D[1] * qy + D[2] * qz + D[3] * qy - D[5] * qw + D[6] * qz + D[7] * qw) -
                4.0f * (D[4] * qx + D[8] * qx);
        gq[1] = 2.0f * (D[1] * qx + D[2] * qw + D[3] * qx + D[5] * qz - D[6] * qw + D[7] * qz) -
                4.0f * (D[0] * qy + D[8] * qy);
        gq[2] = 2.0f * (-D[1] * qw + D[2] * qx + D[3] * qw + D[5] * qy + D[6] * qx + D[7] * qy) -
                4.0f * (D[0] * qz + D[4] * qz);
        gq[3] = 2.0f * (-D[1] * qz + D[2] * qy + D[3] * qz - D[5] * qx - D[6] * qy + D[7] * qx);
        // the gradient row replaces the feature row in shared memory (band mask and factors applied)
        row[0] = make_float4(gq[0] * p.q_factor, gq[1] * p.q_factor, gq[2] * p.q_factor, gq[3] * p.q_factor);
        row[1] = make_float4(gs3[0] * p.s_factor, gs3[1] * p.s_factor, gs3[2] * p.s_factor,
                             glogit * p.alpha_factor);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                float o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int k = 4 * k4 + e;
                    const float f = (k == 0) ? p.color_factor : p.high_order_color_factor;
                    o[e] = (k < p.keep) ? sh[16 * ch + k] * f : 0.0f;
                }
                row[2 + 4 * ch + k4] = make_float4(o[0], o[1], o[2], o[3]);
            }
        p.grad_xyz[3 * pid] = gxyz[0];
        p.grad_xyz[3 * pid + 1] = gxyz[1];
        p.grad_xyz[3 * pid + 2] = gxyz[2];
    }
    __syncwarp();
#pragma unroll 8
    for (int r = 0; r < 32; r += 2) {
        const long long rp = pid0 + r + half;
        if (col < 14 && rp < p.first + p.count)
            reinterpret_cast<float4 *>(p.grad_features + (size_t)GS_FEATURES * rp)[col] =
                rows[(r + half) * kRowStride4 + col];
    }
}

__global__ void __launch_bounds__(kBatchBlock) point_bwd_batch_kernel(BatchParams p)
{
    __shared__ BatchRows s_rows[kBatchBlock / 32];
    batch_block(p, s_rows, (long long)blockIdx.x);
}

// ---- the pass fused with the step's exchange over NVLink / NVSwitch -------------------------------------
// One launch does the per-point pass AND the sum of the N x 59 gradient over the ranks.  The gradient
// buffer is SYMMETRIC memory (the same allocation on every rank, mapped into every peer, with a multicast
// address: a load from it is reduced in the switch over all ranks' copies, a store to it lands in all of
// them).  The points are cut into `chunks` ranges.  CTAs take one of two roles:
//   compute CTAs  pull blocks of 128 points by ticket (in order) and write their rows into the LOCAL copy;
//                 the CTA that completes a range tells every rank so (a flag in that rank's flag array);
//   reducer CTAs  for each range in turn: wait until every rank has completed it, sum THIS rank's share of
//                 it over all ranks with multimem.ld_reduce and write the sum to all ranks with multimem.st
//                 (two-shot all-reduce), then tell every rank; finally wait until every rank has reduced
//                 every range, so that the buffer is final on this rank when the kernel ends.
// The reduction of a range thus runs in the switch while the following ranges are computed, with no
// collective launch per range.  Flags carry the caller's strictly increasing `epoch`, so nothing is reset
// between steps.  Waits are bounded (~10 s of clock): on expiry the kernel records an error word and goes
// on instead of hanging the GPU.
constexpr int kMaxExchangeRanks = 16;
constexpr long long kSpinLimit = 20000000000ll;  // ~10 s at 1.9 GHz

struct ExchangeParams {
    float *mc_base;      // multicast mapping of the gradient buffer
    uint32_t *flags[kMaxExchangeRanks];  // every rank's flag array: [2][chunks][world]
    int rank, world, chunks;
    uint32_t epoch;
    uint32_t *scratch;   // local: ticket, error, chunk_done[chunks], reduce_done[chunks]
    int reducers;        // CTAs in the reducer role
    long long n;         // points (the buffer holds 56 n feature floats, then 3 n position floats)
    long long blocks, blocks_per_chunk;
};

__device__ __forceinline__ void st_release_sys(uint32_t *addr, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *addr)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce(const float *mc)
{
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st(float *mc, const float4 v)
{
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}
// wait until *addr >= epoch (flags only grow); false on expiry
__device__ __forceinline__ bool wait_flag(const uint32_t *addr, uint32_t epoch)
{
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(addr) - epoch) < 0) {
        if (clock64() - t0 > kSpinLimit) return false;
        __nanosleep(256);  // system-scope loads of one word by every waiting CTA: keep them rare
    }
    return true;
}
// this rank's share [lo, hi) of a segment of len4 float4s
__device__ __forceinline__ void my_share(long long len4, int rank, int world, long long &lo, long long &hi)
{
    const long long per = (len4 + world - 1) / world;
    lo = min(per * rank, len4);
    hi = min(lo + per, len4);
}

__global__ void __launch_bounds__(kBatchBlock) point_bwd_batch_exchange_kernel(BatchParams p, ExchangeParams x)
{
    __shared__ BatchRows s_rows[kBatchBlock / 32];
    __shared__ long long s_block;
    __shared__ bool s_ok;
    uint32_t *ticket = x.scratch, *error = x.scratch + 1;
    uint32_t *chunk_done = x.scratch + 2, *reduce_done = x.scratch + 2 + x.chunks;
    const int tid = threadIdx.x;

    if ((int)blockIdx.x >= x.reducers) {
        // ---- compute role ----
        for (;;) {
            __syncthreads();  // s_block and the staged rows of the previous block are no longer read
            if (tid == 0) s_block = (long long)atomicAdd(ticket, 1u);
            __syncthreads();
            const long long block = s_block;
            if (block >= x.blocks) return;
            batch_block(p, s_rows, block);
            __syncthreads();  // every row of the block has been stored by this CTA
            if (tid == 0) {
                __threadfence_system();  // ... and is visible, to the peers too, before the count
                const int c = (int)(block / x.blocks_per_chunk);
                const long long in_chunk = min(x.blocks_per_chunk, x.blocks - (long long)c * x.blocks_per_chunk);
                if (atomicAdd(chunk_done + c, 1u) == (uint32_t)(in_chunk - 1)) {
                    __threadfence_system();  // the whole range, to every peer, before its flag
                    for (int r = 0; r < x.world; ++r)
                        st_release_sys(x.flags[r] + (size_t)c * x.world + x.rank, x.epoch);
                }
            }
        }
    }

    // ---- reducer role ----
    const long long threads = (long long)x.reducers * kBatchBlock;
    const long long me = (long long)blockIdx.x * kBatchBlock + tid;
    uint32_t *my_flags = x.flags[x.rank];
    for (int c = 0; c < x.chunks; ++c) {
        if (tid == 0) {  // one poller per CTA
            bool ok = true;
            for (int r = 0; r < x.world; ++r) ok = wait_flag(my_flags + (size_t)c * x.world + r, x.epoch) && ok;
            s_ok = ok;
        }
        __syncthreads();
        if (!s_ok && tid == 0) atomicExch(error, 1u);
        // rows [r0, r1) of this range: 56 floats of features each, then 3 floats of positions each
        const long long r0 = min((long long)c * x.blocks_per_chunk * kBatchBlock, x.n);
        const long long r1 = min(r0 + x.blocks_per_chunk * kBatchBlock, x.n);
        const long long seg_base[2] = {14 * r0, 14 * x.n + (3 * r0) / 4};          // in float4s (r0 % 128 == 0)
        const long long seg_len[2] = {14 * (r1 - r0), (3 * (r1 - r0) + 3) / 4};     // the tail float4 may be padding
#pragma unroll
        for (int sgm = 0; sgm < 2; ++sgm) {
            long long lo, hi;
            my_share(seg_len[sgm], x.rank, x.world, lo, hi);
            float4 *mc = reinterpret_cast<float4 *>(x.mc_base) + seg_base[sgm];
            constexpr int kUnroll = 8;
            for (long long i = lo + me; i < hi; i += threads * kUnroll) {
                float4 v[kUnroll];
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    const long long j = i + u * threads;
                    if (j < hi) v[u] = multimem_ld_reduce(reinterpret_cast<const float *>(mc + j));
                }
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    const long long j = i + u * threads;
                    if (j < hi) multimem_st(reinterpret_cast<float *>(mc + j), v[u]);
                }
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();
            if (atomicAdd(reduce_done + c, 1u) == (uint32_t)(x.reducers - 1)) {
                __threadfence_system();
                for (int r = 0; r < x.world; ++r)
                    st_release_sys(x.flags[r] + (size_t)(x.chunks + c) * x.world + x.rank, x.epoch);
            }
        }
    }
    // the buffer is final here once every rank has reduced its share of every range
    if (blockIdx.x == 0) {
        bool ok = true;
        for (int f = tid; f < x.chunks * x.world; f += kBatchBlock)
            ok = wait_flag(my_flags + (size_t)x.chunks * x.world + f, x.epoch) && ok;
        if (!ok) atomicExch(error, 2u);
    }
}

}  // namespace

extern "C" int64_t gs_exchange_flag_words(int chunks, int world)
{
    return (chunks > 0 && world > 0) ? 2ll * chunks * world : GS_E_INVALID;
}
extern "C" int64_t gs_exchange_scratch_bytes(int chunks) { return chunks > 0 ? 4ll * (2 + 2 * chunks) : GS_E_INVALID; }

extern "C" int gs_point_bwd_batch_exchange(const float *xyz, const float *features, const int32_t *object_id,
                                           int64_t n, const gs_view_t *views_device, int num_views,
                                           int color_max_sh_band, float q_factor, float s_factor,
                                           float alpha_factor, float color_factor, float high_order_color_factor,
                                           const gs_exchange_t *exchange, void *stream)
{
    if (n <= 0 || num_views < 0 || !exchange) return GS_E_INVALID;
    const gs_exchange_t &e = *exchange;
    if (e.world < 2 || e.world > kMaxExchangeRanks || e.rank < 0 || e.rank >= e.world || e.chunks < 1 ||
        e.epoch == 0 || !e.multicast_base || !e.local_base || !e.scratch)
        return GS_E_INVALID;
    if (!xyz || !features || !object_id || (num_views > 0 && !views_device)) return GS_E_INVALID;
    if (((uintptr_t)features | (uintptr_t)e.local_base | (uintptr_t)e.multicast_base | (uintptr_t)views_device) & 15u)
        return GS_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    BatchParams p;
    p.xyz = xyz; p.features = features; p.object_id = object_id;
    p.first = 0; p.count = n; p.views = views_device; p.num_views = num_views;
    p.keep = color_max_sh_band <= 0 ? 1 : color_max_sh_band == 1 ? 4 : color_max_sh_band == 2 ? 9 : 16;
    p.q_factor = q_factor; p.s_factor = s_factor; p.alpha_factor = alpha_factor;
    p.color_factor = color_factor; p.high_order_color_factor = high_order_color_factor;
    p.inline_one = 0;
    p.grad_features = e.local_base;                    // features first (16-byte aligned rows for every n)
    p.grad_xyz = e.local_base + (size_t)GS_FEATURES * n;
    ExchangeParams x;
    x.mc_base = (float *)e.multicast_base;
    for (int r = 0; r < kMaxExchangeRanks; ++r) x.flags[r] = r < e.world ? (uint32_t *)e.flags[r] : nullptr;
    for (int r = 0; r < e.world; ++r)
        if (!x.flags[r]) return GS_E_INVALID;
    x.rank = e.rank; x.world = e.world; x.chunks = e.chunks; x.epoch = e.epoch;
    x.scratch = (uint32_t *)e.scratch;
    x.n = n;
    x.blocks = (n + kBatchBlock - 1) / kBatchBlock;
    x.blocks_per_chunk = (x.blocks + e.chunks - 1) / e.chunks;
    // every range must hold at least one block: fewer ranges for a tiny cloud
    x.chunks = (int)((x.blocks + x.blocks_per_chunk - 1) / x.blocks_per_chunk);
    static int resident[64];
    int dev = 0;
    GS_CUDA_TRY(cudaGetDevice(&dev));
    if (resident[dev & 63] == 0) {
        int sms = 0, per_sm = 0;
        GS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        GS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, point_bwd_batch_exchange_kernel,
                                                                  kBatchBlock, 0));
        resident[dev & 63] = sms * (per_sm > 0 ? per_sm : 1);
    }
    // measured on B200s at 1M points (tools/exp_peer_exchange.py, pass + exchange after 8 views): more
    // reducer CTAs than this only add contention in the switch (8 ranks: 16 / 32 / 74 / 148 CTAs -> 0.93 /
    // 0.83 / 0.88 / 1.01 ms against 1.07 ms for the pass followed by NCCL's all_reduce; 2 ranks: 32 / 48 / 74 /
    // 148 -> 1.03 / 0.94 / 0.91 / 1.08 ms against 0.90 ms)
    // (4 ranks: 48 / 64 -> 0.85 / 0.89 ms against 1.02 ms)
    int reducers = e.reducer_ctas > 0 ? e.reducer_ctas : (e.world <= 2 ? 72 : e.world <= 4 ? 48 : 32);
    if (reducers > resident[dev & 63] / 2) reducers = resident[dev & 63] / 2;
    if (reducers < 1) reducers = 1;
    x.reducers = reducers;
    long long compute = resident[dev & 63] - reducers;
    if (compute > x.blocks) compute = x.blocks;
    if (compute < 1) compute = 1;
    // NOTE: the flag words of a launch use the caller's chunk count for their layout
    if (x.chunks != e.chunks) return GS_E_INVALID;
    GS_CUDA_TRY(cudaMemsetAsync(x.scratch, 0, (size_t)gs_exchange_scratch_bytes(x.chunks), st));
    point_bwd_batch_exchange_kernel<<<(unsigned)(reducers + compute), kBatchBlock, 0, st>>>(p, x);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int gs_point_bwd_batch(const float *xyz, const float *features, const int32_t *object_id,
                                  int64_t first_point, int64_t num_points, const gs_view_t *views_device,
                                  int num_views, int color_max_sh_band, float q_factor, float s_factor,
                                  float alpha_factor, float color_factor, float high_order_color_factor,
                                  float *grad_xyz, float *grad_features, void *stream)
{
    if (first_point < 0 || num_points < 0 || num_views < 0) return GS_E_INVALID;
    if (num_points == 0) return GS_OK;
    if (!xyz || !features || !object_id || !grad_xyz || !grad_features || (num_views > 0 && !views_device))
        return GS_E_INVALID;
    if (((uintptr_t)features | (uintptr_t)grad_features | (uintptr_t)views_device) & 15u) return GS_E_INVALID;
    BatchParams p;
    p.xyz = xyz; p.features = features; p.object_id = object_id;
    p.first = first_point; p.count = num_points; p.views = views_device; p.num_views = num_views;
    p.keep = color_max_sh_band <= 0 ? 1 : color_max_sh_band == 1 ? 4 : color_max_sh_band == 2 ? 9 : 16;
    p.q_factor = q_factor; p.s_factor = s_factor; p.alpha_factor = alpha_factor;
    p.color_factor = color_factor; p.high_order_color_factor = high_order_color_factor;
    p.grad_xyz = grad_xyz; p.grad_features = grad_features;
    p.inline_one = 0;
    const long long blocks = (num_points + kBatchBlock - 1) / kBatchBlock;
    point_bwd_batch_kernel<<<(unsigned)blocks, kBatchBlock, 0, (cudaStream_t)stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}

extern "C" int gs_point_bwd_all_points(const float *xyz, const float *features, const int32_t *object_id,
                                       int64_t n, const gs_view_t *view, int color_max_sh_band,
                                       float q_factor, float s_factor, float alpha_factor, float color_factor,
                                       float high_order_color_factor, float *grad_xyz, float *grad_features,
                                       void *stream)
{
    if (n < 0 || !view) return GS_E_INVALID;
    if (n == 0) return GS_OK;
    if (!xyz || !features || !object_id || !grad_xyz || !grad_features) return GS_E_INVALID;
    if (!view->slot_of_point || !view->point_in_camera || !view->splat_grads || !view->intrinsics ||
        !view->q_cam_pc || !view->t_pc_cam)
        return GS_E_INVALID;
    if (((uintptr_t)features | (uintptr_t)grad_features | (uintptr_t)view->splat_grads) & 15u) return GS_E_INVALID;
    BatchParams p;
    p.xyz = xyz; p.features = features; p.object_id = object_id;
    p.first = 0; p.count = n; p.views = nullptr; p.num_views = 1;
    p.keep = color_max_sh_band <= 0 ? 1 : color_max_sh_band == 1 ? 4 : color_max_sh_band == 2 ? 9 : 16;
    p.q_factor = q_factor; p.s_factor = s_factor; p.alpha_factor = alpha_factor;
    p.color_factor = color_factor; p.high_order_color_factor = high_order_color_factor;
    p.grad_xyz = grad_xyz; p.grad_features = grad_features;
    p.one = *view;
    p.inline_one = 1;
    const long long blocks = (n + kBatchBlock - 1) / kBatchBlock;
    point_bwd_batch_kernel<<<(unsigned)blocks, kBatchBlock, 0, (cudaStream_t)stream>>>(p);
    GS_LAUNCH_CHECK();
    return GS_OK;
}
