// Shared device helpers for the sm_100a rasteriser kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <map>
#include <set>
#include <mutex>
#include <utility>

#include "gs_b200.h"

#define GS_LAUNCH_CHECK()                          \
    do {                                           \
        cudaError_t e__ = cudaGetLastError();      \
        if (e__ != cudaSuccess) return (int)e__;   \
    } while (0)

#define GS_CUDA_TRY(expr)                          \
    do {                                           \
        cudaError_t e__ = (expr);                  \
        if (e__ != cudaSuccess) return (int)e__;   \
    } while (0)

// Function attributes (opt-in dynamic shared memory) and occupancy numbers are PER DEVICE: a process
// that drives several GPUs must set / query them once on each.  Bit d of `done` = done on device d.
struct GsPerDeviceOnce {
    std::atomic<unsigned long long> done{0ull};
    bool need(int &dev)
    {
        if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
        return dev < 0 || dev >= 64 || !((done.load(std::memory_order_acquire) >> dev) & 1ull);
    }
    void mark(int dev)
    {
        if (dev >= 0 && dev < 64) done.fetch_or(1ull << dev, std::memory_order_release);
    }
};

constexpr int kTile = GS_TILE;
constexpr int kTilePixels = GS_TILE * GS_TILE;
constexpr int kRecFloats = GS_RECORD_FLOATS;
constexpr int kGradSlots = GS_GRAD_SLOTS;

// exp of the literal-fp32 profile: Cody-Waite reduction + Cephes polynomial with fma at
// fixed places.  Must stay operation-for-operation identical to gso_expf in
// oracle/gs_oracle.c.  Written with the _rn intrinsics, which the compiler never contracts, so the
// result has the same bits in a unit built with -fmad=false (project.cu, blend_fwd.cu) and in a
// fast-fp32 unit (blend_bwd.cu, which re-decides pairs on the alpha = 1/255 edge with it).
__device__ __forceinline__ float gs_expf_core(float x)
{
    float t = __fmul_rn(x, 1.44269504088896341f);
    float tj = __fadd_rn(t, 12582912.0f);  // 1.5 * 2^23: the low mantissa bits of tj hold rint(t)
    float j = __fsub_rn(tj, 12582912.0f);
    float f = fmaf(j, -0.693145751953125f, x);
    f = fmaf(j, -1.42860682030941723212e-6f, f);
    float p = 1.9875691500e-4f;
    p = fmaf(p, f, 1.3981999507e-3f);
    p = fmaf(p, f, 8.3334519073e-3f);
    p = fmaf(p, f, 4.1665795894e-2f);
    p = fmaf(p, f, 1.6666665459e-1f);
    p = fmaf(p, f, 5.0000001201e-1f);
    float f2 = __fmul_rn(f, f);
    float r = __fadd_rn(fmaf(p, f2, f), 1.0f);
    // (int)j << 23 without the conversion: bits(tj) = 0x4B400000 + j and 0x4B400000 << 23 == 0 (mod 2^32)
    return __int_as_float(__float_as_int(r) + (__float_as_int(tj) << 23));
}

__device__ __forceinline__ float gs_expf(float x)
{
    if (!(x >= -87.0f)) return (x != x) ? x : 0.0f;
    if (x > 88.0f) return __int_as_float(0x7f800000);
    return gs_expf_core(x);
}

__device__ __forceinline__ float gs_sigmoid(float x) { return 1.0f / (1.0f + gs_expf(-x)); }

// reference GaussianPoint3D.py:30-48
__device__ __forceinline__ void gs_rotation_from_quaternion(float x, float y, float z, float w,
                                                            float R[9])
{
    float xx = x * x, yy = y * y, zz = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.0f - 2.0f * (yy + zz); R[1] = 2.0f * (xy - wz);        R[2] = 2.0f * (xz + wy);
    R[3] = 2.0f * (xy + wz);        R[4] = 1.0f - 2.0f * (xx + zz); R[5] = 2.0f * (yz - wx);
    R[6] = 2.0f * (xz - wy);        R[7] = 2.0f * (yz + wx);        R[8] = 1.0f - 2.0f * (xx + yy);
}

// reference SphericalHarmonics.py:10-32 (direction is normalised first)
__device__ __forceinline__ void gs_sh_basis(float dx, float dy, float dz, float b[16])
{
    float s = (dx * dx + dy * dy) + dz * dz;
    float inv = 1.0f / sqrtf(s);
    float x = inv * dx, y = inv * dy, z = inv * dz;
    b[0] = (float)0.28209479177387814;
    b[1] = (float)-0.48860251190291987 * y;
    b[2] = (float)0.48860251190291987 * z;
    b[3] = (float)-0.48860251190291987 * x;
    b[4] = (float)1.0925484305920792 * x * y;
    b[5] = (float)-1.0925484305920792 * y * z;
    b[6] = (float)0.94617469575755997 * z * z - (float)0.31539156525251999;
    b[7] = (float)-1.0925484305920792 * x * z;
    b[8] = (float)0.54627421529603959 * x * x - (float)0.54627421529603959 * y * y;
    b[9] = (float)0.59004358992664352 * y * (-3.0f * x * x + y * y);
    b[10] = (float)2.8906114426405538 * x * y * z;
    b[11] = (float)0.45704579946446572 * y * (1.0f - 5.0f * z * z);
    b[12] = (float)0.3731763325901154 * z * (5.0f * z * z - 3.0f);
    b[13] = (float)0.45704579946446572 * x * (1.0f - 5.0f * z * z);
    b[14] = (float)1.4453057213202769 * z * (x * x - y * y);
    b[15] = (float)0.59004358992664352 * x * (-x * x + 3.0f * y * y);
}

// ---- decision band of the backward blend ------------------------------------------------------------
// The backward blend forms alpha with fast arithmetic (FMA-contracted quadratic form, ex2.approx) but
// must take the reference's decision alpha >= 1/255, which the reference takes on ITS fp32 arithmetic
// (R.py:634, utils.py:331-345).  band = a bound of the relative difference between the two alphas of a
// pair of this splat NEAR the threshold: the quadratic forms differ by <= 7 u P (u = 2^-24, P = sum of
// the magnitudes of the form's terms <= quad (1 + rho) / (1 - rho), rho = |b| / sqrt(a c), and
// quad = tau = 2 ln(255 peak) at the threshold), i.e. 2.1e-7 tau (1 + rho) / (1 - rho) in alpha; the
// exponentials, the constants and the products add 1.5e-6.  Both with a factor >= 2 of slack; capped at
// 2 % (the forward's culling bits guarantee alpha < 0.9754 / 255 for the pairs they drop, so a dropped
// pair is never one the band could reach; the cap is met beyond an axis ratio of ~50 : 1 at full
// opacity, where the bound stops being rigorous).
// Returned, and stored as the 12th float of the splat's record: s = log2(peak / hi) with
// hi = (1 + band) / 255 -- the form the backward's inner loop consumes (blend_bwd.cu, band_consts).
__device__ __forceinline__ float gs_decision_band_log2(float a, float b, float c, float peak)
{
    if (!(peak > 0.0f) || !(a > 0.0f) || !(c > 0.0f) || !(a * c > b * b) || !isfinite(a + b + c + peak))
        return 0.0f;  // never reaches the threshold, or irregular (decided by the literal pass alone)
    const float rho = fminf(sqrtf((b * b) / (a * c)), 0.9999f);
    const float tau = fmaxf(2.0f * logf(255.0f * peak), 0.0f);
    const float band = fminf(3e-6f + 6e-7f * tau * ((1.0f + rho) / (1.0f - rho)), 0.02f);
    return log2f((255.0f * peak) / (1.0f + band));
}

// ---- per-cell culling shared by the two blend kernels --------------------------------------------
// A warp owns a column of kCells cells of 8x4 pixels.  Bit j of the result is set <=> the splat
// may reach alpha >= 1/255 at a pixel centre of cell j (rows [kRows j, kRows (j+1)) of the column, first
// pixel centre (x_first, y_first)).  q(d) = a dx^2 + 2 b dx dy + c dy^2 must stay <= tau =
// 2 ln(255 * peak), peak = rescale * opacity, there; per pixel row the minimum over the row's 8
// centres is taken at dx clamped to the row, so the test is exact up to the margin (0.05 absolute
// + 4e-6 relative to the positive terms, far above the rounding of either arithmetic profile).
// Conics that are not positive definite / finite keep every bit (plus kCellMaskIrregular).
constexpr int kCellW = 8, kCellH = 4;
// set together with every cell bit for a conic that is not positive definite / finite: only such
// an entry can produce an exponent above zero or NaN, so only it needs the guarded exp
constexpr unsigned kCellMaskIrregular = 1u << 8;

template <int kCells, int kRows = kCellH>
__device__ __forceinline__ unsigned gs_cell_mask(float u, float v, float a, float b, float c,
                                                 float peak, float x_first, float y_first)
{
    constexpr unsigned kAll = (1u << kCells) - 1u;
    if (!(a > 0.0f) || !(c > 0.0f) || !(a * c > b * b) || !isfinite(a + b + c + u + v + peak))
        return kAll | kCellMaskIrregular;
    if (!(peak * 255.0f >= 0.97f)) return 0u;  // alpha <= peak < 1/255 everywhere (e <= 0)
    const float tau = fmaf(2.0f, __logf(peak * 255.0f), 0.05f);
    const float slope = -b / a;  // unconstrained minimiser of a row: dx = slope * dy
    const float lo = x_first - u, hi = lo + (float)(kCellW - 1);
    const float b2 = 2.0f * b;
    unsigned mask = 0u;
#pragma unroll
    for (int r = 0; r < kCells * kRows; ++r) {
        const float dy = (y_first + (float)r) - v;
        const float dx = fminf(fmaxf(slope * dy, lo), hi);
        const float pos = fmaf(a * dx, dx, (c * dy) * dy);
        const float q = fmaf(b2 * dx, dy, pos);
        if (q <= fmaf(4e-6f, pos, tau)) mask |= 1u << (r / kRows);
    }
    return mask;
}

// The culling bits of one staged list entry (48-byte record r0 = u v a b, r1 = c rescale opacity red,
// r2 = green blue depth pad) for a warp's kCells cells, as BOTH blend kernels must see them: entries
// with a conic that is not positive definite / finite, or with a non-finite colour / depth / opacity,
// keep every bit and carry kCellMaskIrregular -- they take the guarded literal arithmetic in the forward
// (the masked "add zero" of the packed path is not a no-op for them); the forward also flags the 8x8
// quarters whose pixels reach such an entry, and the backward leaves those quarters to its literal pass
// (the decision alpha >= 1/255 is then taken with the literal exp, never with ex2.approx of a positive
// / NaN exponent).
template <int kCells>
__device__ __forceinline__ unsigned gs_entry_mask(const float4 &r0, const float4 &r1, const float4 &r2,
                                                  float x_first, float y_first)
{
    unsigned fine = gs_cell_mask<kCells, kCellH>(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y * r1.z, x_first, y_first);
    if (!isfinite(((r1.y + r1.z) + (r1.w + r2.x)) + (r2.y + r2.z)))
        fine = ((1u << kCells) - 1u) | kCellMaskIrregular;
    return fine;
}
__device__ __forceinline__ unsigned long long gs_ld_volatile_u64(const unsigned long long *p)
{
    return *reinterpret_cast<const volatile unsigned long long *>(p);
}
__device__ __forceinline__ void gs_st_volatile_u64(unsigned long long *p, unsigned long long v)
{
    *reinterpret_cast<volatile unsigned long long *>(p) = v;
}
__device__ __forceinline__ uint32_t gs_ld_volatile_u32(const uint32_t *p)
{
    return *reinterpret_cast<const volatile uint32_t *>(p);
}
__device__ __forceinline__ void gs_st_volatile_u32(uint32_t *p, uint32_t v)
{
    *reinterpret_cast<volatile uint32_t *>(p) = v;
}

// ---- work counters of the persistent blend kernels -------------------------------------------------
// The blend kernels run as many CTAs as fit the machine and every WARP pulls its next unit of work (a
// quarter / half tile) from a counter, so a warp whose pixels saturate early moves on instead of
// idling until the slowest warp of its CTA retires.  One {next, done} pair per (device, stream): a
// stream runs its launches in order and the last warp of a launch resets the pair, so no memset is
// needed between launches.  nullptr (table full, allocation failed) = the caller launches one CTA per
// tile instead.
// A launch that FAILED (the launch call or the call before it returned an error) may not have re-armed
// its pair: the launcher reports it with gs_work_counter_failed() and the next gs_work_counter_arm()
// of that pair zeroes it in stream order before the kernel; a pair that was never reported costs the
// arm call one lock and a lookup in an empty set.
struct GsWorkCounter {
    unsigned int next, done;
};

inline GsWorkCounter *gs_work_counter(cudaStream_t st)
{
    constexpr int kSlots = 4096;
    struct PerDevice {
        GsWorkCounter *base = nullptr;
        int used = 0;
        bool failed = false;
    };
    static std::mutex mu;
    static std::map<int, PerDevice> devices;
    static std::map<std::pair<int, cudaStream_t>, GsWorkCounter *> table;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    auto it = table.find({dev, st});
    if (it != table.end()) return it->second;
    PerDevice &d = devices[dev];
    if (d.failed) return nullptr;
    if (!d.base) {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) return nullptr;
        if (cudaMalloc(&d.base, sizeof(GsWorkCounter) * kSlots) != cudaSuccess ||
            cudaMemset(d.base, 0, sizeof(GsWorkCounter) * kSlots) != cudaSuccess) {
            (void)cudaGetLastError();
            d.failed = true;
            d.base = nullptr;
            return nullptr;
        }
    }
    if (d.used >= kSlots) return nullptr;
    GsWorkCounter *c = d.base + d.used++;
    table[{dev, st}] = c;
    return c;
}

inline std::mutex &gs_work_counter_dirty_lock()
{
    static std::mutex mu;
    return mu;
}
inline std::set<GsWorkCounter *> &gs_work_counter_dirty_set()
{
    static std::set<GsWorkCounter *> dirty;
    return dirty;
}
// the launch that used `c` did not (or may not) run to its end
inline void gs_work_counter_failed(GsWorkCounter *c)
{
    if (!c) return;
    std::lock_guard<std::mutex> lock(gs_work_counter_dirty_lock());
    gs_work_counter_dirty_set().insert(c);
}
// before a persistent launch on `st`: re-arms a pair a failed launch left behind
inline cudaError_t gs_work_counter_arm(GsWorkCounter *c, cudaStream_t st)
{
    if (!c) return cudaSuccess;
    {
        std::lock_guard<std::mutex> lock(gs_work_counter_dirty_lock());
        std::set<GsWorkCounter *> &dirty = gs_work_counter_dirty_set();
        if (dirty.empty() || !dirty.erase(c)) return cudaSuccess;
    }
    return cudaMemsetAsync(c, 0, sizeof(GsWorkCounter), st);
}

// next unit of the calling warp (all lanes get it); units >= total mean "done"
__device__ __forceinline__ unsigned gs_next_unit(GsWorkCounter *ctr)
{
    unsigned u = 0;
    if ((threadIdx.x & 31) == 0) u = atomicAdd(&ctr->next, 1u);
    return __shfl_sync(0xffffffffu, u, 0);
}

// called once by every warp of the launch when it has run out of units; the last one re-arms the pair
__device__ __forceinline__ void gs_retire_warp(GsWorkCounter *ctr, unsigned total_warps)
{
    if ((threadIdx.x & 31) == 0) {
        const unsigned d = atomicAdd(&ctr->done, 1u);
        if (d == total_warps - 1u) {
            ctr->next = 0u;
            ctr->done = 0u;
        }
    }
}
