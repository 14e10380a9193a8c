// Loss front-end: the step between the rasteriser's image and its
// cotangent.  One forward pass produces L1, SSIM and the three SSIM derivative maps; one backward
// pass turns the maps into dL/d(rasterised image), clamp and HWC layout included.
//
// Replaces, per training iteration (reference GaussianPointTrainer.py:172-182 and
// LossFunction.py:20-52): torch.clamp, permute, abs().mean(), pytorch_msssim.ssim (ten grouped
// conv2d launches + ~25 elementwise ones) and the autograd backward of all of them.
//
// Arithmetic profile: fast fp32 (FMA contraction); compared with oracle/loss_oracle.py in fp64 at
// the tolerances written in tests/test_loss_gpu.py.
#include <cuda_runtime.h>
#include <stdint.h>

#include "gs_b200.h"
#include "gs_common.cuh"

namespace {

constexpr int kWin = 11;            // pytorch_msssim window (win_size=11, win_sigma=1.5)
constexpr int kLossTile = 32;       // output tile edge
constexpr int kHalo = kLossTile + kWin - 1;  // 42
constexpr int kInStride = kHalo + 1;         // 43: odd stride, conflict-free row-parallel reads
constexpr int kMidStride = kLossTile + 1;    // 33
constexpr int kLossThreads = 128;
constexpr int kRowsPerThread = kLossTile / (kLossThreads / 32);  // 8
constexpr float kC1 = 0.01f * 0.01f, kC2 = 0.03f * 0.03f;

// `_fspecial_gauss_1d(11, 1.5)` as torch evaluates it in fp32 (oracle/loss_oracle.py holds the same table)
__constant__ float c_win[kWin] = {0x1.0d9570p-10f, 0x1.f1fe02p-8f, 0x1.26eb18p-5f, 0x1.bff0fep-4f,
                                  0x1.b43c3ep-3f,  0x1.106560p-2f, 0x1.b43c3ep-3f, 0x1.bff0fep-4f,
                                  0x1.26eb18p-5f,  0x1.f1fe02p-8f, 0x1.0d9570p-10f};

struct Strides {
    long long h, w, c;  // in elements
};

__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// 4-byte asynchronous global -> shared copy (LDGSTS); !valid writes a zero without reading.
__device__ __forceinline__ void cp_async4(float *dst, const float *src, const bool valid)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int bytes = valid ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// Starts the copy of the 42x42 halo of all three channels into s[c][row][kInStride]; positions
// outside the image become 0.  One warp per row; the lanes run along the tensor's unit stride
// (channel-interleaved HWC rows of 126 floats, or planar rows of 42).  No register staging: every
// copy of the block is in flight at once (cp_async_wait_all + __syncthreads before use).
__device__ __forceinline__ void load_halo_async(float *s, const float *src, const Strides st, const int y0,
                                                const int x0, const int height, const int width)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (st.c == 1) {
        // everything that depends on the position in the row is the same for every row: hoisted
        int dst_off[4], src_off[4];
        bool col_ok[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int j = lane + 32 * t, col = j / 3, ch = j - col * 3;
            col_ok[t] = j < 3 * kHalo && (x0 + col) < width;
            dst_off[t] = ch * kHalo * kInStride + col;
            src_off[t] = (int)((x0 + col) * st.w) + ch;
        }
        for (int r = warp; r < kHalo; r += kLossThreads / 32) {
            const bool row_ok = (y0 + r) < height;
            const float *row = src + (row_ok ? (y0 + r) * st.h : 0);
            float *drow = s + r * kInStride;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                if (lane + 32 * t < 3 * kHalo) {
                    const bool ok = row_ok && col_ok[t];
                    cp_async4(drow + dst_off[t], ok ? row + src_off[t] : src, ok);
                }
            }
        }
    } else {
        for (int cr = warp; cr < 3 * kHalo; cr += kLossThreads / 32) {
            const int ch = cr / kHalo, r = cr - ch * kHalo;
            const int y = y0 + r;
            const bool row_ok = y < height;
            const float *row = src + (row_ok ? y * st.h + ch * st.c : 0);
#pragma unroll
            for (int col = lane; col < kHalo; col += 32) {
                const bool ok = row_ok && (x0 + col) < width;
                cp_async4(s + cr * kInStride + col, ok ? row + (x0 + col) * st.w : src, ok);
            }
        }
    }
}

__device__ __forceinline__ float block_sum(float v, float *s_red)
{
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < kLossThreads / 32; ++i) t += s_red[i];
    return t;
}

// ---------------------------------------------------------------------------------------------
// Forward: per 32x32 tile of the (H-10)x(W-10) SSIM map.
//   maps [3 kinds][3 channels][H-10][W-10]: d ssim / d filtered X (at fixed filtered XX, XY),
//   d ssim / d filtered XX, d ssim / d filtered XY.
//   partials [blocks][2]: sum of the tile's ssim values, sum of |x - y| over the pixels it owns.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kLossThreads)
loss_fwd_kernel(const float *__restrict__ image, const Strides ist, const float *__restrict__ gt,
                const Strides gst, const int height, const int width, float *__restrict__ maps,
                float *__restrict__ partials)
{
    extern __shared__ float smem[];
    float *sx = smem;                                // [3][42][43]
    float *sy = sx + 3 * kHalo * kInStride;          // [3][42][43]
    float *sm = sy + 3 * kHalo * kInStride;          // [5][42][33]
    __shared__ float s_red[kLossThreads / 32];

    const int oh = height - (kWin - 1), ow = width - (kWin - 1);
    const int y0 = blockIdx.y * kLossTile, x0 = blockIdx.x * kLossTile;
    load_halo_async(sx, image, ist, y0, x0, height, width);
    load_halo_async(sy, gt, gst, y0, x0, height, width);
    cp_async_wait_all();
    __syncthreads();

    // One pass over the halo: clamp the prediction in place (GaussianPointTrainer.py:173) and sum
    // |x - y| over the pixels this block owns: its 32x32 input square, the last row / column of
    // blocks also taking what is left of the image (at most 10 more rows / columns, inside the halo).
    float l1 = 0.0f;
    {
        const int own_h = (blockIdx.y == gridDim.y - 1) ? min(kHalo, height - y0) : kLossTile;
        const int own_w = (blockIdx.x == gridDim.x - 1) ? min(kHalo, width - x0) : kLossTile;
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            for (int r = warp; r < kHalo; r += kLossThreads / 32) {
#pragma unroll
                for (int col = lane; col < kHalo; col += 32) {
                    const int i = (ch * kHalo + r) * kInStride + col;
                    const float x = clamp01(sx[i]);
                    sx[i] = x;
                    if (r < own_h && col < own_w) l1 += fabsf(x - sy[i]);
                }
            }
        }
    }
    __syncthreads();

    float w[kWin];
#pragma unroll
    for (int k = 0; k < kWin; ++k) w[k] = c_win[k];

    float ssim_sum = 0.0f;
    const int vc = threadIdx.x & 31, vg = threadIdx.x >> 5;  // vertical pass: column, group of 8 rows
    for (int ch = 0; ch < 3; ++ch) {
        // horizontal pass: (halo row, group of 4 columns) per thread, 5 filtered quantities
        for (int task = threadIdx.x; task < kHalo * (kLossTile / 4); task += kLossThreads) {
            const int r = task >> 3, g = task & 7;
            const float *px = sx + (ch * kHalo + r) * kInStride + 4 * g;
            const float *py = sy + (ch * kHalo + r) * kInStride + 4 * g;
            float a[5][4];
#pragma unroll
            for (int q = 0; q < 5; ++q)
#pragma unroll
                for (int o = 0; o < 4; ++o) a[q][o] = 0.0f;
#pragma unroll
            for (int j = 0; j < 4 + kWin - 1; ++j) {
                const float x = px[j], y = py[j];
                const float xx = x * x, yy = y * y, xy = x * y;
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    const int k = j - o;
                    if (k >= 0 && k < kWin) {
                        a[0][o] = fmaf(w[k], x, a[0][o]);
                        a[1][o] = fmaf(w[k], y, a[1][o]);
                        a[2][o] = fmaf(w[k], xx, a[2][o]);
                        a[3][o] = fmaf(w[k], yy, a[3][o]);
                        a[4][o] = fmaf(w[k], xy, a[4][o]);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < 5; ++q)
#pragma unroll
                for (int o = 0; o < 4; ++o) sm[(q * kHalo + r) * kMidStride + 4 * g + o] = a[q][o];
        }
        __syncthreads();

        // vertical pass: 8 consecutive output rows of one column per thread
        float f[5][kRowsPerThread];
#pragma unroll
        for (int q = 0; q < 5; ++q) {
#pragma unroll
            for (int o = 0; o < kRowsPerThread; ++o) f[q][o] = 0.0f;
#pragma unroll
            for (int j = 0; j < kRowsPerThread + kWin - 1; ++j) {
                const float v = sm[(q * kHalo + vg * kRowsPerThread + j) * kMidStride + vc];
#pragma unroll
                for (int o = 0; o < kRowsPerThread; ++o) {
                    const int k = j - o;
                    if (k >= 0 && k < kWin) f[q][o] = fmaf(w[k], v, f[q][o]);
                }
            }
        }
        const int ox = x0 + vc;
#pragma unroll
        for (int o = 0; o < kRowsPerThread; ++o) {
            const int oy = y0 + vg * kRowsPerThread + o;
            if (oy < oh && ox < ow) {
                const float mu1 = f[0][o], mu2 = f[1][o];
                const float s1 = f[2][o] - mu1 * mu1, s2 = f[3][o] - mu2 * mu2, s12 = f[4][o] - mu1 * mu2;
                const float a1 = 2.0f * mu1 * mu2 + kC1, a2 = 2.0f * s12 + kC2;
                const float b1 = mu1 * mu1 + mu2 * mu2 + kC1, b2 = s1 + s2 + kC2;
                const float rb1 = __frcp_rn(b1), rb2 = __frcp_rn(b2);
                const float lum = a1 * rb1, cs = a2 * rb2;
                ssim_sum += lum * cs;
                const float d_s1 = -(lum * cs) * rb2;
                const float d_s12 = 2.0f * lum * rb2;
                const float d_mu1 = 2.0f * cs * rb1 * (mu2 - mu1 * lum);
                const float d_mu = d_mu1 - 2.0f * mu1 * d_s1 - mu2 * d_s12;
                const size_t plane = (size_t)oh * ow, at = (size_t)oy * ow + ox;
                maps[(0 * 3 + ch) * plane + at] = d_mu;
                maps[(1 * 3 + ch) * plane + at] = d_s1;
                maps[(2 * 3 + ch) * plane + at] = d_s12;
            }
        }
        __syncthreads();
    }
    const float t_ssim = block_sum(ssim_sum, s_red);
    const float t_l1 = block_sum(l1, s_red);
    if (threadIdx.x == 0) {
        const int b = blockIdx.y * gridDim.x + blockIdx.x;
        partials[2 * b] = t_ssim;
        partials[2 * b + 1] = t_l1;
    }
}

// ---------------------------------------------------------------------------------------------
// Scale regulariser, forward partials (LossFunction.py:40-52): per block, the sum over its valid
// points of ||exp(s)||_2 and their count.  reg_partials [blocks][2] (sum, count as float).
// ---------------------------------------------------------------------------------------------
constexpr int kRegThreads = 256;
constexpr int kRegBlocks = 148 * 4;

__global__ void __launch_bounds__(kRegThreads)
scale_reg_fwd_kernel(const float *__restrict__ features, const int8_t *__restrict__ invalid_mask,
                     const long long n, float *__restrict__ reg_partials)
{
    __shared__ float s_sum[kRegThreads / 32], s_cnt[kRegThreads / 32];
    float sum = 0.0f, cnt = 0.0f;
    for (long long i = (long long)blockIdx.x * kRegThreads + threadIdx.x; i < n;
         i += (long long)gridDim.x * kRegThreads) {
        // the row is fetched whether or not the point is valid: the loads are independent of the mask
        const float *row = features + i * 56 + 4;
        const float s0 = __ldg(row), s1 = __ldg(row + 1), s2 = __ldg(row + 2);
        if (invalid_mask[i] != 0) continue;
        const float e0 = expf(s0), e1 = expf(s1), e2 = expf(s2);
        sum += sqrtf(e0 * e0 + e1 * e1 + e2 * e2);
        cnt += 1.0f;
    }
    sum = warp_sum(sum);
    cnt = warp_sum(cnt);
    if ((threadIdx.x & 31) == 0) {
        s_sum[threadIdx.x >> 5] = sum;
        s_cnt[threadIdx.x >> 5] = cnt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.0f, b = 0.0f;
        for (int i = 0; i < kRegThreads / 32; ++i) {
            a += s_sum[i];
            b += s_cnt[i];
        }
        reg_partials[2 * blockIdx.x] = a;
        reg_partials[2 * blockIdx.x + 1] = b;
    }
}

// One block: fixed-order fp64 sums of the partials (deterministic), then the scalars
//   0 loss  1 L1  2 1-SSIM  3 regulariser value (0 if absent)  4 number of valid points  5 SSIM
__global__ void __launch_bounds__(256)
loss_finalize_kernel(const float *__restrict__ partials, const int blocks,
                     const float *__restrict__ reg_partials, const int reg_blocks, const int height,
                     const int width, const float lambda, const float reg_weight,
                     float *__restrict__ scalars)
{
    __shared__ double s[4][256];
    double a = 0.0, b = 0.0, c = 0.0, d = 0.0;
    for (int i = threadIdx.x; i < blocks; i += 256) {
        a += (double)partials[2 * i];
        b += (double)partials[2 * i + 1];
    }
    for (int i = threadIdx.x; i < reg_blocks; i += 256) {
        c += (double)reg_partials[2 * i];
        d += (double)reg_partials[2 * i + 1];
    }
    s[0][threadIdx.x] = a;
    s[1][threadIdx.x] = b;
    s[2][threadIdx.x] = c;
    s[3][threadIdx.x] = d;
    __syncthreads();
    for (int half = 128; half > 0; half >>= 1) {
        if (threadIdx.x < half)
            for (int q = 0; q < 4; ++q) s[q][threadIdx.x] += s[q][threadIdx.x + half];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double oh = height - (kWin - 1), ow = width - (kWin - 1);
        const double ssim = s[0][0] / (3.0 * oh * ow);
        const double l1 = s[1][0] / (3.0 * (double)height * (double)width);
        double loss = (1.0 - (double)lambda) * l1 + (double)lambda * (1.0 - ssim);
        double reg = 0.0;
        if (reg_blocks > 0) {
            reg = s[2][0] / s[3][0];  // NaN without valid points, as the mean of an empty tensor is
            loss += (double)reg_weight * reg;
        }
        scalars[0] = (float)loss;
        scalars[1] = (float)l1;
        scalars[2] = (float)(1.0 - ssim);
        scalars[3] = (float)reg;
        scalars[4] = (float)s[3][0];
        scalars[5] = (float)ssim;
    }
}

// ---------------------------------------------------------------------------------------------
// Backward: per 32x32 tile of the image.  dL/dx(p) = upstream * [ g_ssim * (Ga + 2 x Gb + y Gc)
// + g_l1 * sign(x - y) ] where G* is the full correlation of a derivative map with the window
// (the adjoint of the valid filter), masked by the clamp (0 <= raw <= 1).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kLossThreads)
loss_bwd_kernel(const float *__restrict__ image, const Strides ist, const float *__restrict__ gt,
                const Strides gst, const int height, const int width, const float g_ssim,
                const float g_l1, const float *__restrict__ maps, const float *__restrict__ upstream,
                float *__restrict__ grad_image)
{
    extern __shared__ float smem[];
    float *sin = smem;                              // [3 kinds][42][43]
    float *sm = sin + 3 * kHalo * kInStride;        // [3 kinds][42][33]
    float *sout = sm + 3 * kHalo * kMidStride;      // [32][97]: the tile's gradient, all channels

    const int oh = height - (kWin - 1), ow = width - (kWin - 1);
    const size_t plane = (size_t)oh * ow;
    const int y0 = blockIdx.y * kLossTile, x0 = blockIdx.x * kLossTile;
    const float up = upstream ? upstream[0] : 1.0f;
    float w[kWin];
#pragma unroll
    for (int k = 0; k < kWin; ++k) w[k] = c_win[k];

    const int vc = threadIdx.x & 31, vg = threadIdx.x >> 5;
    // halo of the three maps of one channel: map positions (y0-10 .. y0+31, x0-10 .. x0+31), zero
    // outside; one warp per row, all copies in flight at once; the column part is row-invariant
    int mcol[2];
    bool mcol_ok[2];
#pragma unroll
    for (int t = 0; t < 2; ++t) {
        const int col = (threadIdx.x & 31) + 32 * t;
        mcol[t] = x0 - (kWin - 1) + col;
        mcol_ok[t] = col < kHalo && mcol[t] >= 0 && mcol[t] < ow;
    }
    auto issue_maps = [&](const int ch, float *dst) {
        for (int qr = threadIdx.x >> 5; qr < 3 * kHalo; qr += kLossThreads / 32) {
            const int q = qr / kHalo, r = qr - q * kHalo;
            const int my = y0 - (kWin - 1) + r;
            const bool row_ok = my >= 0 && my < oh;
            const float *row = maps + (q * 3 + ch) * plane + (row_ok ? (size_t)my * ow : 0);
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                if ((int)(threadIdx.x & 31) + 32 * t < kHalo) {
                    const bool ok = row_ok && mcol_ok[t];
                    cp_async4(dst + qr * kInStride + (threadIdx.x & 31) + 32 * t, ok ? row + mcol[t] : maps, ok);
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int ch = 0; ch < 3; ++ch) {
        // (a second buffer to prefetch the next channel costs a resident block per SM: measured slower)
        const float *cur = sin;
        issue_maps(ch, sin);
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        // horizontal: out(x) = sum_k w[k] * m(x - k); halo column of m(x - k) is (x - x0) + 10 - k
        for (int task = threadIdx.x; task < kHalo * (kLossTile / 4); task += kLossThreads) {
            const int r = task >> 3, g = task & 7;
            float a[3][4];
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const float *pm = cur + (q * kHalo + r) * kInStride + 4 * g;
#pragma unroll
                for (int o = 0; o < 4; ++o) a[q][o] = 0.0f;
#pragma unroll
                for (int j = 0; j < 4 + kWin - 1; ++j) {
                    const float v = pm[j];
#pragma unroll
                    for (int o = 0; o < 4; ++o) {
                        const int k = o + (kWin - 1) - j;
                        if (k >= 0 && k < kWin) a[q][o] = fmaf(w[k], v, a[q][o]);
                    }
                }
#pragma unroll
                for (int o = 0; o < 4; ++o) sm[(q * kHalo + r) * kMidStride + 4 * g + o] = a[q][o];
            }
        }
        __syncthreads();
        float f[3][kRowsPerThread];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
#pragma unroll
            for (int o = 0; o < kRowsPerThread; ++o) f[q][o] = 0.0f;
#pragma unroll
            for (int j = 0; j < kRowsPerThread + kWin - 1; ++j) {
                const float v = sm[(q * kHalo + vg * kRowsPerThread + j) * kMidStride + vc];
#pragma unroll
                for (int o = 0; o < kRowsPerThread; ++o) {
                    const int k = o + (kWin - 1) - j;
                    if (k >= 0 && k < kWin) f[q][o] = fmaf(w[k], v, f[q][o]);
                }
            }
        }
        const int x = x0 + vc;
#pragma unroll
        for (int o = 0; o < kRowsPerThread; ++o) {
            const int ly = vg * kRowsPerThread + o, y = y0 + ly;
            float g = 0.0f;
            if (y < height && x < width) {
                const float raw = image[y * ist.h + x * ist.w + ch * ist.c];
                const float t = gt[y * gst.h + x * gst.w + ch * gst.c];
                const float xc = clamp01(raw);
                const float d = xc - t;
                const float sgn = d > 0.0f ? 1.0f : (d < 0.0f ? -1.0f : 0.0f);
                g = g_ssim * (f[0][o] + 2.0f * xc * f[1][o] + t * f[2][o]) + g_l1 * sgn;
                g = (raw >= 0.0f && raw <= 1.0f) ? g * up : 0.0f;
            }
            sout[ly * (3 * kLossTile + 1) + 3 * vc + ch] = g;
        }
        __syncthreads();
    }
    // coalesced store of the tile: rows of up to 96 consecutive floats of the HWC gradient
    const int tw = min(kLossTile, width - x0), th = min(kLossTile, height - y0);
    for (int r = threadIdx.x >> 5; r < th; r += kLossThreads / 32) {
        float *row = grad_image + ((size_t)(y0 + r) * width + x0) * 3;
        for (int j = threadIdx.x & 31; j < 3 * tw; j += 32) row[j] = sout[r * (3 * kLossTile + 1) + j];
    }
}

// Scale regulariser, backward: d/ds_j = weight * upstream * exp(s_j)^2 / ||exp(s)|| / n_valid for
// valid points, written (or added) to columns 4:7 of grad_features [n,56].
__global__ void __launch_bounds__(kRegThreads)
scale_reg_bwd_kernel(const float *__restrict__ features, const int8_t *__restrict__ invalid_mask,
                     const long long n, const float reg_weight, const float *__restrict__ scalars,
                     const float *__restrict__ upstream, float *__restrict__ grad_features,
                     const int accumulate)
{
    const float scale = reg_weight * (upstream ? upstream[0] : 1.0f) / scalars[4];
    for (long long i = (long long)blockIdx.x * kRegThreads + threadIdx.x; i < n;
         i += (long long)gridDim.x * kRegThreads) {
        if (invalid_mask[i] != 0) continue;
        const float *row = features + i * 56 + 4;
        const float e0 = expf(row[0]), e1 = expf(row[1]), e2 = expf(row[2]);
        const float inv = scale / sqrtf(e0 * e0 + e1 * e1 + e2 * e2);
        float *out = grad_features + i * 56 + 4;
        if (accumulate) {
            out[0] += e0 * e0 * inv;
            out[1] += e1 * e1 * inv;
            out[2] += e2 * e2 * inv;
        } else {
            out[0] = e0 * e0 * inv;
            out[1] = e1 * e1 * inv;
            out[2] = e2 * e2 * inv;
        }
    }
}

constexpr size_t kFwdSmem = (size_t)(2 * 3 * kHalo * kInStride + 5 * kHalo * kMidStride) * sizeof(float);
constexpr size_t kBwdSmem =
    (size_t)(3 * kHalo * kInStride + 3 * kHalo * kMidStride + kLossTile * (3 * kLossTile + 1)) * sizeof(float);

inline int tiles_of(int v) { return (v + kLossTile - 1) / kLossTile; }

}  // namespace

extern "C" int64_t gs_loss_workspace_bytes(int height, int width)
{
    if (height < kWin || width < kWin) return GS_E_INVALID;
    const int64_t blocks = (int64_t)tiles_of(height - (kWin - 1)) * tiles_of(width - (kWin - 1));
    return (blocks * 2 + kRegBlocks * 2) * (int64_t)sizeof(float);
}

extern "C" int64_t gs_loss_maps_bytes(int height, int width)
{
    if (height < kWin || width < kWin) return GS_E_INVALID;
    return (int64_t)9 * (height - (kWin - 1)) * (width - (kWin - 1)) * (int64_t)sizeof(float);
}

extern "C" int gs_loss_fwd(const float *image, int64_t image_stride_h, int64_t image_stride_w,
                           int64_t image_stride_c, const float *gt, int64_t gt_stride_h,
                           int64_t gt_stride_w, int64_t gt_stride_c, int height, int width,
                           float lambda_value, const float *features, const int8_t *invalid_mask,
                           int64_t n, float reg_weight, float *maps, void *workspace, float *scalars,
                           void *stream)
{
    if (!image || !gt || !maps || !workspace || !scalars) return GS_E_INVALID;
    if (height < kWin || width < kWin) return GS_E_INVALID;
    if (features && (!invalid_mask || n < 0)) return GS_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    static GsPerDeviceOnce attr_set;
    int dev = 0;
    if (attr_set.need(dev)) {
        const cudaError_t e =
            cudaFuncSetAttribute(loss_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem);
        if (e != cudaSuccess) return (int)e;
        attr_set.mark(dev);
    }
    const dim3 grid(tiles_of(width - (kWin - 1)), tiles_of(height - (kWin - 1)));
    const int blocks = (int)(grid.x * grid.y);
    float *partials = (float *)workspace;
    float *reg_partials = partials + 2 * (size_t)blocks;
    const Strides ist{image_stride_h, image_stride_w, image_stride_c};
    const Strides gst{gt_stride_h, gt_stride_w, gt_stride_c};
    loss_fwd_kernel<<<grid, kLossThreads, kFwdSmem, st>>>(image, ist, gt, gst, height, width, maps, partials);
    int reg_blocks = 0;
    if (features) {
        reg_blocks = (int)((n + kRegThreads - 1) / kRegThreads);
        if (reg_blocks > kRegBlocks) reg_blocks = kRegBlocks;
        if (reg_blocks < 1) reg_blocks = 1;
        scale_reg_fwd_kernel<<<reg_blocks, kRegThreads, 0, st>>>(features, invalid_mask, n, reg_partials);
    }
    loss_finalize_kernel<<<1, 256, 0, st>>>(partials, blocks, reg_partials, reg_blocks, height, width,
                                            lambda_value, reg_weight, scalars);
    return (int)cudaGetLastError();
}

extern "C" int gs_loss_bwd(const float *image, int64_t image_stride_h, int64_t image_stride_w,
                           int64_t image_stride_c, const float *gt, int64_t gt_stride_h,
                           int64_t gt_stride_w, int64_t gt_stride_c, int height, int width,
                           float lambda_value, const float *maps, const float *upstream,
                           float *grad_image, void *stream)
{
    if (!image || !gt || !maps || !grad_image) return GS_E_INVALID;
    if (height < kWin || width < kWin) return GS_E_INVALID;
    static GsPerDeviceOnce attr_set;
    int dev = 0;
    if (attr_set.need(dev)) {
        const cudaError_t e =
            cudaFuncSetAttribute(loss_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBwdSmem);
        if (e != cudaSuccess) return (int)e;
        attr_set.mark(dev);
    }
    const double oh = height - (kWin - 1), ow = width - (kWin - 1);
    const float g_ssim = (float)(-(double)lambda_value / (3.0 * oh * ow));
    const float g_l1 = (float)((1.0 - (double)lambda_value) / (3.0 * (double)height * (double)width));
    const dim3 grid(tiles_of(width), tiles_of(height));
    const Strides ist{image_stride_h, image_stride_w, image_stride_c};
    const Strides gst{gt_stride_h, gt_stride_w, gt_stride_c};
    loss_bwd_kernel<<<grid, kLossThreads, kBwdSmem, (cudaStream_t)stream>>>(
        image, ist, gt, gst, height, width, g_ssim, g_l1, maps, upstream, grad_image);
    return (int)cudaGetLastError();
}

extern "C" int gs_scale_reg_bwd(const float *features, const int8_t *invalid_mask, int64_t n,
                                float reg_weight, const float *scalars, const float *upstream,
                                float *grad_features, int accumulate, void *stream)
{
    if (!features || !invalid_mask || !scalars || !grad_features || n < 0) return GS_E_INVALID;
    if (n == 0) return GS_OK;
    int blocks = (int)((n + kRegThreads - 1) / kRegThreads);
    if (blocks > kRegBlocks) blocks = kRegBlocks;
    scale_reg_bwd_kernel<<<blocks, kRegThreads, 0, (cudaStream_t)stream>>>(
        features, invalid_mask, n, reg_weight, scalars, upstream, grad_features, accumulate);
    return (int)cudaGetLastError();
}
