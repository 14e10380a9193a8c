/*
 * gs_oracle.c -- CPU restatement of the reference rasteriser hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import, link or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and there only as the checker
 * or as the timed CPU baseline.
 *
 * What it restates (paths relative to the reference checkout, R.py =
 * taichi_3d_gaussian_splatting/GaussianPointCloudRasterisation.py, G3D.py =
 * GaussianPoint3D.py, SH.py = SphericalHarmonics.py, U.py = utils.py):
 *
 *   gso_inverse_pose      U.py:396-432  (inverse_SE3_qt_torch), used R.py:845
 *   gso_cull              R.py:31-78    (filter_point_in_camera), G3D.py:14-62
 *   gso_compact           R.py:861-865  (ascending ids of the mask)
 *   gso_project           R.py:239-315  (generate_point_attributes_in_camera_plane),
 *                         R.py:196-236, G3D.py:30-87,111-116,162-191,333-349,
 *                         SH.py:10-45, U.py:257-272,351-353,495-510
 *   gso_tile_counts       R.py:81-128   (bbox + generate_num_overlap_tiles)
 *   gso_exclusive_scan    R.py:913-922  (cumsum + shift)
 *   gso_emit_keys         R.py:131-172
 *   gso_sort_pairs        R.py:947-950  (ascending int64 keys, stable in slot order)
 *   gso_tile_ranges       R.py:175-193
 *   gso_blend_fwd         R.py:318-485, U.py:275-284
 *   gso_blend_bwd         R.py:531-705, U.py:331-348
 *   gso_point_bwd         R.py:707-772, G3D.py:132-159,237-331,351-373, U.py:356-359
 *   gso_grad_epilogue     R.py:1102-1125,1167-1182
 *
 * Arithmetic profile ("literal fp32"): every binary operation of the Python
 * source is one IEEE-754 binary32 operation, evaluated in Python's parse order
 * (left to right), matrix products accumulate k = 0,1,2,.. (Taichi's matmul
 * helper), no FMA contraction, correctly rounded / and sqrt.  The one function
 * the reference takes from a third party (Taichi -> CUDA libdevice expf, not in
 * /root/reference; requirements.txt:1 pins only taichi>=1.6.0) is replaced by
 * gso_expf below: Cody-Waite reduction + the Cephes degree-5 polynomial, using
 * explicit fmaf at fixed places, so that the CUDA kernels can reproduce it bit
 * for bit.  Build with -ffp-contract=off (see oracle/Makefile).
 *
 * Parity pin: tests/test_oracle_golden.py checks this file against the
 * reference's own golden vectors (tile ranges T_R.py:18-51, Sigma'
 * tests/GaussianPoint3D_test.py:12-54, single point alpha/grads T_R.py:353-548,
 * pose inverse tests/utils_test.py:138-157) and against fixtures produced by
 * executing the reference's Python source under a numpy-float32 Taichi shim
 * (tests/golden/gen_golden.py, tests/test_reference_golden.py).  Against Taichi's real CUDA code generation the
 * parity is unpinned (Taichi cannot be installed in the build container).
 *
 * A second build with -DGSO_DOUBLE makes every `real` a double and uses libm
 * exp(): the arbiter for gradient tolerances.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef GSO_DOUBLE
typedef double real;
#define FN(name) gso64_##name
#define R_SQRT sqrt
#define R_FLOOR floor
#define R_FABS fabs
#define R_FMAX fmax
#define R_FMIN fmin
#define R_EXP exp
#else
typedef float real;
#define FN(name) gso_##name
#define R_SQRT sqrtf
#define R_FLOOR floorf
#define R_FABS fabsf
#define R_FMAX fmaxf
#define R_FMIN fminf
#define R_EXP gso_expf
#endif

#define TILE 16
#define BOUNDARY_TILES 3

/* ------------------------------------------------------------------------- */
/* exp: the replacement for the third-party expf (see header).               */
float gso_expf(float x)
{
    if (!(x >= -87.0f)) return (x != x) ? x : 0.0f;
    if (x > 88.0f) return INFINITY;
    float t = x * 1.44269504088896341f;
    float j = (t + 12582912.0f) - 12582912.0f; /* round to nearest even integer */
    float f = fmaf(j, -0.693145751953125f, x);
    f = fmaf(j, -1.42860682030941723212e-6f, f);
    float p = 1.9875691500e-4f;
    p = fmaf(p, f, 1.3981999507e-3f);
    p = fmaf(p, f, 8.3334519073e-3f);
    p = fmaf(p, f, 4.1665795894e-2f);
    p = fmaf(p, f, 1.6666665459e-1f);
    p = fmaf(p, f, 5.0000001201e-1f);
    float f2 = f * f;
    float r = fmaf(p, f2, f) + 1.0f;
    int32_t ji = (int32_t)j;
    uint32_t bits;
    memcpy(&bits, &r, 4);
    bits += ((uint32_t)ji) << 23;
    memcpy(&r, &bits, 4);
    return r;
}

void gso_expf_array(const float *x, float *y, int64_t n)
{
    for (int64_t i = 0; i < n; ++i) y[i] = gso_expf(x[i]);
}

/* Optional tile-row band for the two blend loops: lets bench.py time a bounded sample of a
 * large frame.  begin < 0 (default) = whole image.  Pixels outside the band are not written. */
static int g_band_begin = -1, g_band_end = -1;
void FN(set_tile_row_band)(int begin, int end) { g_band_begin = begin; g_band_end = end; }
static int tile_in_band(int tile_v) { return g_band_begin < 0 || (tile_v >= g_band_begin && tile_v < g_band_end); }

/* torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU arm sets its thread count itself */
void FN(set_num_threads)(int threads)
{
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#else
    (void)threads;
#endif
}

int FN(num_threads)(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------- */
/* small dense helpers, Taichi matmul order: acc = a[i,0]*b[0,j]; acc += ...  */
static void matmul(const real *a, const real *b, real *c, int n, int p, int m)
{
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) {
            real acc = a[i * p + 0] * b[0 * m + j];
            for (int k = 1; k < p; ++k) acc = acc + a[i * p + k] * b[k * m + j];
            c[i * m + j] = acc;
        }
}

static void transpose(const real *a, real *t, int n, int m)
{
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) t[j * n + i] = a[i * m + j];
}

/* G3D.py:30-48 */
static void rotation_from_quaternion(const real q[4], real R[9])
{
    real x = q[0], y = q[1], z = q[2], w = q[3];
    real xx = x * x, yy = y * y, zz = z * z;
    real xy = x * y, xz = x * z, yz = y * z;
    real wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1 - 2 * (yy + zz); R[1] = 2 * (xy - wz);     R[2] = 2 * (xz + wy);
    R[3] = 2 * (xy + wz);     R[4] = 1 - 2 * (xx + zz); R[5] = 2 * (yz - wx);
    R[6] = 2 * (xz - wy);     R[7] = 2 * (yz + wx);     R[8] = 1 - 2 * (xx + yy);
}

/* G3D.py:51-62 */
static void transform_from_qt(const real q[4], const real t[3], real T[16])
{
    real R[9];
    rotation_from_quaternion(q, R);
    T[0] = R[0]; T[1] = R[1]; T[2] = R[2];  T[3] = t[0];
    T[4] = R[3]; T[5] = R[4]; T[6] = R[5];  T[7] = t[1];
    T[8] = R[6]; T[9] = R[7]; T[10] = R[8]; T[11] = t[2];
    T[12] = 0;   T[13] = 0;   T[14] = 0;    T[15] = 1;
}

/* G3D.py:14-27 */
static void project_point(const real xyz[3], const real T[16], const real K[9],
                          real uv[2], real pcam[3])
{
    real h[4] = {xyz[0], xyz[1], xyz[2], 1};
    real hc[4];
    matmul(T, h, hc, 4, 4, 1);
    pcam[0] = hc[0]; pcam[1] = hc[1]; pcam[2] = hc[2];
    real kp[3];
    matmul(K, pcam, kp, 3, 3, 1);
    uv[0] = kp[0] / pcam[2];
    uv[1] = kp[1] / pcam[2];
}

/* taichi Matrix.normalized(): invlen = 1/sqrt(sum x*x); invlen * v */
static void normalize(real *v, int n)
{
    real s = v[0] * v[0];
    for (int i = 1; i < n; ++i) s = s + v[i] * v[i];
    real invlen = 1 / R_SQRT(s);
    for (int i = 0; i < n; ++i) v[i] = invlen * v[i];
}

/* SH.py:10-32 */
static void sh_basis(const real dir[3], real b[16])
{
    real d[3] = {dir[0], dir[1], dir[2]};
    normalize(d, 3);
    real x = d[0], y = d[1], z = d[2];
    b[0] = (real)0.28209479177387814;
    b[1] = (real)-0.48860251190291987 * y;
    b[2] = (real)0.48860251190291987 * z;
    b[3] = (real)-0.48860251190291987 * x;
    b[4] = (real)1.0925484305920792 * x * y;
    b[5] = (real)-1.0925484305920792 * y * z;
    b[6] = (real)0.94617469575755997 * z * z - (real)0.31539156525251999;
    b[7] = (real)-1.0925484305920792 * x * z;
    b[8] = (real)0.54627421529603959 * x * x - (real)0.54627421529603959 * y * y;
    b[9] = (real)0.59004358992664352 * y * ((real)-3.0 * x * x + y * y);
    b[10] = (real)2.8906114426405538 * x * y * z;
    b[11] = (real)0.45704579946446572 * y * ((real)1.0 - (real)5.0 * z * z);
    b[12] = (real)0.3731763325901154 * z * ((real)5.0 * z * z - (real)3.0);
    b[13] = (real)0.45704579946446572 * x * ((real)1.0 - (real)5.0 * z * z);
    b[14] = (real)1.4453057213202769 * z * (x * x - y * y);
    b[15] = (real)0.59004358992664352 * x * (-x * x + (real)3.0 * y * y);
}

static real dot16(const real *a, const real *b)
{
    real s = a[0] * b[0];
    for (int i = 1; i < 16; ++i) s = s + a[i] * b[i];
    return s;
}

static real sigmoid(real x) { return 1 / (1 + R_EXP(-x)); } /* U.py:351-353 */

/* G3D.py:65-87 */
static void projective_jacobian(const real K[9], const real p[3], real J[6])
{
    real fx = K[0], fy = K[4];
    real x = p[0], y = p[1], z = p[2];
    J[0] = fx / z; J[1] = 0;      J[2] = -(fx * x) / (z * z);
    J[3] = 0;      J[4] = fy / z; J[5] = -(fy * y) / (z * z);
}

/* ------------------------------------------------------------------------- */
/* U.py:396-432: q_inv = conj(q); t_inv = -rotate(q_inv/|q_inv|, t)           */
static void quat_mul(const real a[4], const real b[4], real o[4])
{
    real x0 = a[0], y0 = a[1], z0 = a[2], w0 = a[3];
    real x1 = b[0], y1 = b[1], z1 = b[2], w1 = b[3];
    o[0] = w0 * x1 + x0 * w1 + y0 * z1 - z0 * y1;
    o[1] = w0 * y1 - x0 * z1 + y0 * w1 + z0 * x1;
    o[2] = w0 * z1 + x0 * y1 - y0 * x1 + z0 * w1;
    o[3] = w0 * w1 - x0 * x1 - y0 * y1 - z0 * z1;
}

void FN(inverse_pose)(const real *q_pc_cam, const real *t_pc_cam, int n_obj,
                      real *q_cam_pc, real *t_cam_pc)
{
    for (int k = 0; k < n_obj; ++k) {
        const real *q = q_pc_cam + 4 * k, *t = t_pc_cam + 3 * k;
        real qi[4] = {-q[0], -q[1], -q[2], q[3]};
        /* torch.norm then divide (U.py:415) */
        real n = R_SQRT(((qi[0] * qi[0] + qi[1] * qi[1]) + qi[2] * qi[2]) + qi[3] * qi[3]);
        real qn[4] = {qi[0] / n, qi[1] / n, qi[2] / n, qi[3] / n};
        real v[4] = {t[0], t[1], t[2], 0};
        real qc[4] = {-qn[0], -qn[1], -qn[2], qn[3]};
        real tmp[4], r[4];
        quat_mul(qn, v, tmp);
        quat_mul(tmp, qc, r);
        q_cam_pc[4 * k + 0] = qi[0]; q_cam_pc[4 * k + 1] = qi[1];
        q_cam_pc[4 * k + 2] = qi[2]; q_cam_pc[4 * k + 3] = qi[3];
        t_cam_pc[3 * k + 0] = -r[0]; t_cam_pc[3 * k + 1] = -r[1]; t_cam_pc[3 * k + 2] = -r[2];
    }
}

/* R.py:31-78 */
void FN(cull)(const real *xyz, const int8_t *invalid_mask, const int32_t *object_id,
              int64_t n, const real *K, const real *q_cam_pc, const real *t_cam_pc,
              real near_plane, real far_plane, int width, int height, int8_t *mask)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        if (invalid_mask[i] == 1) { mask[i] = 0; continue; }
        int o = object_id[i];
        real T[16];
        transform_from_qt(q_cam_pc + 4 * o, t_cam_pc + 3 * o, T);
        real uv[2], pc[3];
        project_point(xyz + 3 * i, T, K, uv, pc);
        real u = uv[0], v = uv[1], z = pc[2];
        int vis = z > near_plane && z < far_plane &&
                  u >= (real)(-TILE * BOUNDARY_TILES) && u < (real)(width + TILE * BOUNDARY_TILES) &&
                  v >= (real)(-TILE * BOUNDARY_TILES) && v < (real)(height + TILE * BOUNDARY_TILES);
        mask[i] = vis ? 1 : 0;
    }
}

/* R.py:861-865 */
int64_t FN(compact)(const int8_t *mask, int64_t n, int32_t *ids)
{
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i)
        if (mask[i]) ids[m++] = (int32_t)i;
    return m;
}

/* G3D.py:162-191 -- Sigma' = J W Sigma W^T J^T, left to right */
static void project_covariance(const real q[4], const real s[3], const real T[16],
                               const real K[9], const real pcam[3], real cov[4])
{
    real J[6];
    projective_jacobian(K, pcam, J);
    real R[9];
    rotation_from_quaternion(q, R);
    real S[9] = {R_EXP(s[0]), 0, 0, 0, R_EXP(s[1]), 0, 0, 0, R_EXP(s[2])};
    real St[9], Rt[9], a[9], b[9], Sigma[9];
    transpose(S, St, 3, 3);
    transpose(R, Rt, 3, 3);
    matmul(R, S, a, 3, 3, 3);
    matmul(a, St, b, 3, 3, 3);
    matmul(b, Rt, Sigma, 3, 3, 3);
    real W[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    real Wt[9], Jt[6], u[6], v[6], w[6];
    transpose(W, Wt, 3, 3);
    transpose(J, Jt, 2, 3);
    matmul(J, W, u, 2, 3, 3);
    matmul(u, Sigma, v, 2, 3, 3);
    matmul(v, Wt, w, 2, 3, 3);
    matmul(w, Jt, cov, 2, 3, 2);
}

void FN(project_covariance)(const real *q, const real *s, const real *T, const real *K,
                            const real *pcam, real *cov)
{
    project_covariance(q, s, T, K, pcam, cov);
}

void FN(rotation_from_quaternion)(const real *q, real *R) { rotation_from_quaternion(q, R); }

/* U.py:257-272 (argument passed by value: caller's covariance stays unfiltered) */
static void conic_and_rescale(const real cov_in[4], real out[4])
{
    real c[4] = {cov_in[0], cov_in[1], cov_in[2], cov_in[3]};
    real det0 = c[0] * c[3] - c[1] * c[2];
    c[0] += (real)0.3;
    c[3] += (real)0.3;
    real det = c[0] * c[3] - c[1] * c[2];
    real rescale = R_SQRT(R_FMAX((real)0.0, det0 / det));
    real id = (real)1.0 / det;
    out[0] = id * c[3];
    out[1] = id * (-c[1]);
    out[2] = id * c[0];
    out[3] = rescale;
}

/* R.py:239-315.  features is updated in place (normalised quaternion). */
void FN(project)(const real *xyz, real *features, const int32_t *object_id,
                 const real *K, const real *q_cam_pc, const real *t_cam_pc,
                 const int32_t *ids, int64_t m,
                 real *point_uv, real *point_in_camera, real *conic, real *alpha,
                 real *color, real *radii)
{
#pragma omp parallel for schedule(static)
    for (int64_t idx = 0; idx < m; ++idx) {
        int64_t pid = ids[idx];
        real *f = features + 56 * pid;
        normalize(f, 4); /* R.py:196-205 */
        const real *p = xyz + 3 * pid;
        int o = object_id[pid];
        real T[16];
        transform_from_qt(q_cam_pc + 4 * o, t_cam_pc + 3 * o, T);
        /* U.py:495-510: ray origin = -(R^T) t */
        real nRt[9] = {-T[0], -T[4], -T[8], -T[1], -T[5], -T[9], -T[2], -T[6], -T[10]};
        real tt[3] = {T[3], T[7], T[11]};
        real origin[3];
        matmul(nRt, tt, origin, 3, 3, 1);

        real uv[2], pc[3], cov[4], cr[4];
        project_point(p, T, K, uv, pc);
        project_covariance(f, f + 4, T, K, pc, cov);
        conic_and_rescale(cov, cr);
        point_uv[2 * idx] = uv[0]; point_uv[2 * idx + 1] = uv[1];
        point_in_camera[3 * idx] = pc[0]; point_in_camera[3 * idx + 1] = pc[1];
        point_in_camera[3 * idx + 2] = pc[2];
        for (int k = 0; k < 4; ++k) conic[4 * idx + k] = cr[k];
        alpha[idx] = (real)1. / ((real)1. + R_EXP(-f[7]));

        real dir[3] = {p[0] - origin[0], p[1] - origin[1], p[2] - origin[2]};
        real basis[16];
        sh_basis(dir, basis);
        color[3 * idx + 0] = sigmoid(dot16(f + 8, basis));
        color[3 * idx + 1] = sigmoid(dot16(f + 24, basis));
        color[3 * idx + 2] = sigmoid(dot16(f + 40, basis));

        real a = cov[0], b = cov[1], c = cov[2], d = cov[3];
        real lam = (a + d + R_SQRT((a - d) * (a - d) + (real)4.0 * b * c)) / (real)2.0;
        radii[idx] = R_SQRT(lam) * (real)3.0;
    }
}

/* R.py:81-103.  Non-finite inputs give an empty box (documented deviation:
 * the reference's float->int cast of NaN is undefined). */
static void bbox(real u, real v, real radius, int width, int height, int box[4])
{
    if (!(isfinite((double)u) && isfinite((double)v) && isfinite((double)radius))) {
        box[0] = box[1] = box[2] = box[3] = 0;
        return;
    }
    real r = R_FMAX(radius, (real)1.0);
    real min_u = R_FMAX((real)0.0, u - r), max_u = u + r;
    real min_v = R_FMAX((real)0.0, v - r), max_v = v + r;
    int tw = width / TILE, th = height / TILE;
    /* clamp before the int cast so huge radii stay defined */
    real cap = (real)1.0e9;
    int tu0 = (int)R_FMIN(R_FLOOR(min_u / TILE), cap);
    tu0 = tu0 < tw ? tu0 : tw;
    int tu1 = (int)R_FMIN(R_FLOOR(max_u / TILE), cap) + 1;
    tu1 = (tu1 > tu0 + 1 ? tu1 : tu0 + 1);
    tu1 = tu1 < tw ? tu1 : tw;
    int tv0 = (int)R_FMIN(R_FLOOR(min_v / TILE), cap);
    tv0 = tv0 < th ? tv0 : th;
    int tv1 = (int)R_FMIN(R_FLOOR(max_v / TILE), cap) + 1;
    tv1 = (tv1 > tv0 + 1 ? tv1 : tv0 + 1);
    tv1 = tv1 < th ? tv1 : th;
    box[0] = tu0; box[1] = tu1; box[2] = tv0; box[3] = tv1;
}

/* R.py:106-128 */
void FN(tile_counts)(const real *point_uv, const real *radii, int64_t m, int width,
                     int height, int32_t *counts)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < m; ++i) {
        int b[4];
        bbox(point_uv[2 * i], point_uv[2 * i + 1], radii[i], width, height, b);
        counts[i] = (b[1] - b[0]) * (b[3] - b[2]);
    }
}

/* R.py:913-922 */
int64_t FN(exclusive_scan)(const int32_t *counts, int64_t m, int64_t *offsets)
{
    int64_t run = 0;
    for (int64_t i = 0; i < m; ++i) { offsets[i] = run; run += counts[i]; }
    return run;
}

/* R.py:131-172 */
void FN(emit_keys)(const real *point_uv, const real *point_in_camera, const real *radii,
                   const int64_t *offsets, int64_t m, int width, int height,
                   real depth_to_sort_key_scale, int64_t *keys, int32_t *payload)
{
    int tw = width / TILE;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < m; ++i) {
        int b[4];
        bbox(point_uv[2 * i], point_uv[2 * i + 1], radii[i], width, height, b);
        real z = point_in_camera[3 * i + 2];
        int32_t depth = (int32_t)(z * depth_to_sort_key_scale);
        for (int tu = b[0]; tu < b[1]; ++tu)
            for (int tv = b[2]; tv < b[3]; ++tv) {
                int64_t slot = offsets[i] + (int64_t)(b[3] - b[2]) * (tu - b[0]) + (tv - b[2]);
                int32_t tile = tu + tv * tw;
                keys[slot] = (int64_t)depth + ((int64_t)tile << 32);
                payload[slot] = (int32_t)i;
            }
    }
}

/* R.py:947-950: ascending signed int64, ties in slot order (stable LSD radix). */
void FN(sort_pairs)(int64_t *keys, int32_t *payload, int64_t k)
{
    if (k <= 1) return;
    int64_t *k2 = (int64_t *)malloc(sizeof(int64_t) * k);
    int32_t *p2 = (int32_t *)malloc(sizeof(int32_t) * k);
    int64_t *ka = keys, *kb = k2;
    int32_t *pa = payload, *pb = p2;
    for (int pass = 0; pass < 4; ++pass) {
        int shift = 16 * pass;
        size_t hist[65537];
        memset(hist, 0, sizeof(hist));
        for (int64_t i = 0; i < k; ++i) {
            uint64_t u = (uint64_t)ka[i] ^ 0x8000000000000000ull;
            hist[((u >> shift) & 0xffff) + 1]++;
        }
        for (int d = 0; d < 65536; ++d) hist[d + 1] += hist[d];
        for (int64_t i = 0; i < k; ++i) {
            uint64_t u = (uint64_t)ka[i] ^ 0x8000000000000000ull;
            size_t pos = hist[(u >> shift) & 0xffff]++;
            kb[pos] = ka[i];
            pb[pos] = pa[i];
        }
        int64_t *tk = ka; ka = kb; kb = tk;
        int32_t *tp = pa; pa = pb; pb = tp;
    }
    /* 4 passes: result is back in the caller's arrays */
    free(k2);
    free(p2);
}

/* R.py:175-193 (start/end must be zero-initialised by the caller, R.py:954-957) */
void FN(tile_ranges)(const int64_t *keys, int64_t k, int32_t *start, int32_t *end)
{
    if (k <= 0) return;
    for (int64_t i = 0; i + 1 < k; ++i) {
        int32_t t = (int32_t)(keys[i] >> 32), tn = (int32_t)(keys[i + 1] >> 32);
        if (t != tn) { start[tn] = (int32_t)(i + 1); end[t] = (int32_t)(i + 1); }
    }
    end[(int32_t)(keys[k - 1] >> 32)] = (int32_t)k;
}

/* U.py:275-284 */
static real density_conic(real px, real py, real mu, real mv, const real c[4])
{
    real dx = px - mu, dy = py - mv;
    real e = (real)-0.5 * (dx * dx * c[0] + dy * dy * c[2]) - dx * dy * c[1];
    return R_EXP(e) * c[3];
}

/* R.py:318-485.  Outputs are written for every pixel (tiles with an empty list
 * give zeros); *pairs_visited counts list entries visited before saturation. */
void FN(blend_fwd)(int height, int width, const int32_t *tile_start, const int32_t *tile_end,
                   const int32_t *payload, const real *point_uv, const real *point_in_camera,
                   const real *conic, const real *alpha_act, const real *color,
                   real *image, real *depth, real *acc_alpha, int32_t *last_effective,
                   int32_t *valid_count, int rgb_only, int64_t *pairs_visited)
{
    int tw = width / TILE, th = height / TILE;
    int64_t visited = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : visited)
    for (int tile = 0; tile < tw * th; ++tile) {
        int tile_u = tile % tw, tile_v = tile / tw;
        if (!tile_in_band(tile_v)) continue;
        int s = tile_start[tile], e = tile_end[tile];
        for (int o = 0; o < TILE * TILE; ++o) {
            int pu = tile_u * TILE + o % TILE, pv = tile_v * TILE + o / TILE;
            real T = (real)1.0, C[3] = {0, 0, 0}, D = 0, Wd = 0;
            int last = s, cnt = 0, saturated = 0;
            for (int i = s; i < e && !saturated; ++i) {
                int g = payload[i];
                visited++;
                real ga = density_conic(pu + (real)0.5, pv + (real)0.5, point_uv[2 * g],
                                        point_uv[2 * g + 1], conic + 4 * g);
                real a = ga * alpha_act[g];
                if (a < (real)1. / (real)255.) continue;
                a = R_FMIN(a, (real)0.99);
                real nT = T * (1 - a);
                if (nT < (real)0.0001) { saturated = 1; continue; }
                last = i + 1;
                for (int ch = 0; ch < 3; ++ch) C[ch] += color[3 * g + ch] * a * T;
                if (!rgb_only) {
                    real z = point_in_camera[3 * g + 2];
                    D += z * a * T;
                    Wd += a * T;
                    cnt += 1;
                }
                T = nT;
            }
            int64_t px = (int64_t)pv * width + pu;
            image[3 * px] = C[0]; image[3 * px + 1] = C[1]; image[3 * px + 2] = C[2];
            if (!rgb_only) {
                depth[px] = D / R_FMAX(Wd, (real)1e-6);
                acc_alpha[px] = (real)1. - T;
                last_effective[px] = last;
                valid_count[px] = cnt;
            }
        }
    }
    if (pairs_visited) *pairs_visited = visited;
}

/* R.py:531-705, U.py:331-348.  Accumulators are double so that the threaded sum
 * order does not show at float precision; they are indexed like the reference's
 * buffers: grad_uv [N,2], grad_logit [N] (feature column 7), magnitude [N] by
 * global point id; cov_buf [M,3], color_buf [M,3], affected [M] by list offset. */
void FN(blend_bwd)(int height, int width, const int32_t *tile_start, const int32_t *tile_end,
                   const int32_t *payload, const int32_t *ids, const real *image_grad,
                   const real *acc_alpha, const int32_t *last_effective,
                   const real *point_uv, const real *conic, const real *alpha_act,
                   const real *color,
                   double *grad_uv, double *grad_logit, double *magnitude,
                   double *cov_buf, double *color_buf, int32_t *affected,
                   real *map_on_image, int64_t *pairs_visited)
{
    int tw = width / TILE, th = height / TILE;
    int64_t visited = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : visited)
    for (int tile = 0; tile < tw * th; ++tile) {
        int tile_u = tile % tw, tile_v = tile / tw;
        if (!tile_in_band(tile_v)) continue;
        int s = tile_start[tile], e = tile_end[tile];
        for (int o = 0; o < TILE * TILE; ++o) {
            int pu = tile_u * TILE + o % TILE, pv = tile_v * TILE + o / TILE;
            int64_t px = (int64_t)pv * width + pu;
            int last = last_effective[px];
            real T = (real)1.0 - acc_alpha[px];
            real w[3] = {0, 0, 0};
            real g[3] = {image_grad[3 * px], image_grad[3 * px + 1], image_grad[3 * px + 2]};
            real tot[2] = {0, 0};
            for (int i = e - 1; i >= s; --i) {
                if (i >= last) continue;
                visited++;
                int gidx = payload[i];
                const real *c = conic + 4 * gidx;
                real dx = (pu + (real)0.5) - point_uv[2 * gidx];
                real dy = (pv + (real)0.5) - point_uv[2 * gidx + 1];
                real v0 = c[0] * dx + c[1] * dy;
                real v1 = c[1] * dx + c[2] * dy;
                real quad = dx * v0 + dy * v1;
                real p = R_EXP((real)-0.5 * quad) * c[3];
                real dpm[2] = {p * v0, p * v1};
                /* 0.5 * p * (inv_cov @ outer @ inv_cov) */
                real outer[4] = {dx * dx, dx * dy, dy * dx, dy * dy};
                real ic[4] = {c[0], c[1], c[1], c[2]};
                real t1[4], t2[4];
                matmul(ic, outer, t1, 2, 2, 2);
                matmul(t1, ic, t2, 2, 2, 2);
                real hp = (real)0.5 * p;
                real dpc[4] = {hp * t2[0], hp * t2[1], hp * t2[2], hp * t2[3]};
                real aa = alpha_act[gidx];
                real prod = p * aa;
#ifdef GSO_DOUBLE
                /* The fp64 build is the arbiter of gradient VALUES; which pairs take part is the
                 * reference's fp32 decision (R.py:634 on fp32 values), so it is taken here with the
                 * same float arithmetic as the fp32 build (the inputs are fp32 values widened). */
                float fdx = (pu + 0.5f) - (float)point_uv[2 * gidx];
                float fdy = (pv + 0.5f) - (float)point_uv[2 * gidx + 1];
                float fc0 = (float)c[0], fc1 = (float)c[1], fc2 = (float)c[2];
                float fv0 = fc0 * fdx + fc1 * fdy;
                float fv1 = fc1 * fdx + fc2 * fdy;
                float fquad = fdx * fv0 + fdy * fv1;
                float fprod = (gso_expf(-0.5f * fquad) * (float)c[3]) * (float)aa;
                int take = fprod >= 1.f / 255.f;
#else
                int take = prod >= (real)1. / (real)255.;
#endif
                if (take) {
                    real a = R_FMIN(prod, (real)0.99);
                    const real *col = color + 3 * gidx;
                    T = T / ((real)1. - a);
                    real dcol = a * T;
                    real agr = 0;
                    real e3[3];
                    for (int ch = 0; ch < 3; ++ch)
                        e3[ch] = (col[ch] * T - w[ch] / ((real)1. - a)) * g[ch];
                    for (int ch = 0; ch < 3; ++ch) w[ch] += col[ch] * a * T;
                    agr = e3[0] + e3[1] + e3[2];
                    real g_act = agr * p;
                    real g_logit = g_act * ((real)1. - aa) * aa;
                    real g_p = agr * aa;
                    real gm[2] = {g_p * dpm[0], g_p * dpm[1]};
                    tot[0] += R_FABS(gm[0]);
                    tot[1] += R_FABS(gm[1]);
                    int64_t pid = ids[gidx];
                    double vals[10] = {gm[0], gm[1], g_p * dpc[0], g_p * dpc[1], g_p * dpc[3],
                                       dcol * g[0], dcol * g[1], dcol * g[2], g_logit,
                                       R_SQRT(gm[0] * gm[0] + gm[1] * gm[1])};
#pragma omp atomic
                    grad_uv[2 * pid] += vals[0];
#pragma omp atomic
                    grad_uv[2 * pid + 1] += vals[1];
#pragma omp atomic
                    cov_buf[3 * gidx] += vals[2];
#pragma omp atomic
                    cov_buf[3 * gidx + 1] += vals[3];
#pragma omp atomic
                    cov_buf[3 * gidx + 2] += vals[4];
#pragma omp atomic
                    color_buf[3 * gidx] += vals[5];
#pragma omp atomic
                    color_buf[3 * gidx + 1] += vals[6];
#pragma omp atomic
                    color_buf[3 * gidx + 2] += vals[7];
#pragma omp atomic
                    grad_logit[pid] += vals[8];
#pragma omp atomic
                    magnitude[pid] += vals[9];
#pragma omp atomic
                    affected[gidx] += 1;
                }
            }
            map_on_image[2 * px] = tot[0];
            map_on_image[2 * px + 1] = tot[1];
        }
    }
    if (pairs_visited) *pairs_visited = visited;
}

/* G3D.py:237-331 -- the dense 4x9 . 9x9 . 9x{4,3} chain exactly as written */
static void covariance_jacobian(const real q[4], const real s[3], const real T[16],
                                const real K[9], const real pcam[3],
                                real dq[16] /*4x4*/, real ds[12] /*4x3*/)
{
    real J[6];
    projective_jacobian(K, pcam, J);
    real R[9];
    rotation_from_quaternion(q, R);
    real es[3] = {R_EXP(s[0]), R_EXP(s[1]), R_EXP(s[2])};
    real S[9] = {es[0], 0, 0, 0, es[1], 0, 0, 0, es[2]};
    real M[9];
    matmul(R, S, M, 3, 3, 3);
    real W[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    real U[6];
    matmul(J, W, U, 2, 3, 3);
#define U_(i, j) U[(i) * 3 + (j)]
#define M_(i, j) M[(i) * 3 + (j)]
#define R_(i, j) R[(i) * 3 + (j)]
    real dSp_dS[36] = {
        U_(0,0)*U_(0,0), U_(0,0)*U_(0,1), U_(0,0)*U_(0,2), U_(0,0)*U_(0,1), U_(0,1)*U_(0,1), U_(0,1)*U_(0,2), U_(0,0)*U_(0,2), U_(0,1)*U_(0,2), U_(0,2)*U_(0,2),
        U_(0,0)*U_(1,0), U_(0,0)*U_(1,1), U_(0,0)*U_(1,2), U_(0,1)*U_(1,0), U_(0,1)*U_(1,1), U_(0,1)*U_(1,2), U_(0,2)*U_(1,0), U_(0,2)*U_(1,1), U_(0,2)*U_(1,2),
        U_(0,0)*U_(1,0), U_(0,1)*U_(1,0), U_(0,2)*U_(1,0), U_(0,0)*U_(1,1), U_(0,1)*U_(1,1), U_(0,2)*U_(1,1), U_(0,0)*U_(1,2), U_(0,1)*U_(1,2), U_(0,2)*U_(1,2),
        U_(1,0)*U_(1,0), U_(1,0)*U_(1,1), U_(1,0)*U_(1,2), U_(1,0)*U_(1,1), U_(1,1)*U_(1,1), U_(1,1)*U_(1,2), U_(1,0)*U_(1,2), U_(1,1)*U_(1,2), U_(1,2)*U_(1,2)};
    real dS_dM[81] = {
        2*M_(0,0), 2*M_(0,1), 2*M_(0,2), 0, 0, 0, 0, 0, 0,
        M_(1,0), M_(1,1), M_(1,2), M_(0,0), M_(0,1), M_(0,2), 0, 0, 0,
        M_(2,0), M_(2,1), M_(2,2), 0, 0, 0, M_(0,0), M_(0,1), M_(0,2),
        M_(1,0), M_(1,1), M_(1,2), M_(0,0), M_(0,1), M_(0,2), 0, 0, 0,
        0, 0, 0, 2*M_(1,0), 2*M_(1,1), 2*M_(1,2), 0, 0, 0,
        0, 0, 0, M_(2,0), M_(2,1), M_(2,2), M_(1,0), M_(1,1), M_(1,2),
        M_(2,0), M_(2,1), M_(2,2), 0, 0, 0, M_(0,0), M_(0,1), M_(0,2),
        0, 0, 0, M_(2,0), M_(2,1), M_(2,2), M_(1,0), M_(1,1), M_(1,2),
        0, 0, 0, 0, 0, 0, 2*M_(2,0), 2*M_(2,1), 2*M_(2,2)};
    real dSp_dM[36];
    matmul(dSp_dS, dS_dM, dSp_dM, 4, 9, 9);
    real dM_dS[27] = {
        R_(0,0), 0, 0,  0, R_(0,1), 0,  0, 0, R_(0,2),
        R_(1,0), 0, 0,  0, R_(1,1), 0,  0, 0, R_(1,2),
        R_(2,0), 0, 0,  0, R_(2,1), 0,  0, 0, R_(2,2)};
    real dS_ds[9] = {es[0], 0, 0, 0, es[1], 0, 0, 0, es[2]};
    real tmp[12];
    matmul(dSp_dM, dM_dS, tmp, 4, 9, 3);
    matmul(tmp, dS_ds, ds, 4, 3, 3);
    real x = q[0], y = q[1], z = q[2], w = q[3];
    real sx = es[0], sy = es[1], sz = es[2];
    real dM_dq[36] = {
        0, -4*sx*y, -4*sx*z, 0,
        2*sy*y, 2*sy*x, -2*sy*w, -2*sy*z,
        2*sz*z, 2*sz*w, 2*sz*x, 2*sz*y,
        2*sx*y, 2*sx*x, 2*sx*w, 2*sx*z,
        -4*sy*x, 0, -4*sy*z, 0,
        -2*sz*w, 2*sz*z, 2*sz*y, -2*sz*x,
        2*sx*z, -2*sx*w, 2*sx*x, -2*sx*y,
        2*sy*w, 2*sy*z, 2*sy*y, 2*sy*x,
        -4*sz*x, -4*sz*y, 0, 0};
    matmul(dSp_dM, dM_dq, dq, 4, 9, 4);
#undef U_
#undef M_
#undef R_
}

void FN(covariance_jacobian)(const real *q, const real *s, const real *T, const real *K,
                             const real *pcam, real *dq, real *ds)
{
    covariance_jacobian(q, s, T, K, pcam, dq, ds);
}

/* G3D.py:132-159 */
static void position_jacobian(const real xyz[3], const real T[16], const real K[9], real out[6])
{
    real h[4] = {xyz[0], xyz[1], xyz[2], 1}, t[4];
    matmul(T, h, t, 4, 4, 1);
    real d[6] = {K[0] / t[2], K[1] / t[2], (-K[0] * t[0] - K[1] * t[1]) / (t[2] * t[2]),
                 K[3] / t[2], K[4] / t[2], (-K[3] * t[0] - K[4] * t[1]) / (t[2] * t[2])};
    real W[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    matmul(d, W, out, 2, 3, 3);
}

void FN(position_jacobian)(const real *xyz, const real *T, const real *K, real *out)
{
    position_jacobian(xyz, T, K, out);
}

/* R.py:707-772.  grad_xyz [N,3], grad_features [N,56] must hold the state left
 * by blend_bwd (zeros except column 7); rows of visible points are stored. */
void FN(point_bwd)(const real *xyz, const real *features, const int32_t *object_id,
                   const real *K, const real *q_cam_pc, const real *t_cam_pc,
                   const real *t_pc_cam, const int32_t *ids, int64_t m,
                   const real *point_in_camera, const real *grad_uv /*[N,2]*/,
                   const real *cov_buf /*[M,3]*/, const real *color_buf /*[M,3]*/,
                   real *grad_xyz, real *grad_features)
{
#pragma omp parallel for schedule(static)
    for (int64_t idx = 0; idx < m; ++idx) {
        int64_t pid = ids[idx];
        const real *f = features + 56 * pid;
        const real *p = xyz + 3 * pid;
        int o = object_id[pid];
        real guv[2] = {grad_uv[2 * pid], grad_uv[2 * pid + 1]};
        real g4[4] = {cov_buf[3 * idx], cov_buf[3 * idx + 1], cov_buf[3 * idx + 1],
                      cov_buf[3 * idx + 2]};
        real gc[3] = {color_buf[3 * idx], color_buf[3 * idx + 1], color_buf[3 * idx + 2]};
        real T[16];
        transform_from_qt(q_cam_pc + 4 * o, t_cam_pc + 3 * o, T);
        const real *origin = t_pc_cam + 3 * o;
        const real *pc = point_in_camera + 3 * idx;
        real duv[6], dq[16], ds[12];
        position_jacobian(p, T, K, duv);
        covariance_jacobian(f, f + 4, T, K, pc, dq, ds);
        real dir[3] = {p[0] - origin[0], p[1] - origin[1], p[2] - origin[2]};
        real basis[16];
        sh_basis(dir, basis);
        real gx[3], gq[4], gs[3];
        matmul(guv, duv, gx, 1, 2, 3);
        matmul(g4, dq, gq, 1, 4, 4);
        matmul(g4, ds, gs, 1, 4, 3);
        real *gf = grad_features + 56 * pid;
        for (int i = 0; i < 3; ++i) grad_xyz[3 * pid + i] = gx[i];
        for (int i = 0; i < 4; ++i) gf[i] = gq[i];
        for (int i = 0; i < 3; ++i) gf[4 + i] = gs[i];
        for (int ch = 0; ch < 3; ++ch) {
            real sg = sigmoid(dot16(f + 8 + 16 * ch, basis));
            real jac = sg * (1 - sg);
            for (int k = 0; k < 16; ++k) gf[8 + 16 * ch + k] = gc[ch] * (jac * basis[k]);
        }
    }
}

/* R.py:1102-1125, 1167-1182 */
void FN(grad_epilogue)(real *grad_features, int64_t n, int color_max_sh_band,
                       real q_factor, real s_factor, real alpha_factor,
                       real color_factor, real high_order_color_factor)
{
    int keep = color_max_sh_band <= 0 ? 1 : color_max_sh_band == 1 ? 4
               : color_max_sh_band == 2 ? 9 : 16;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        real *g = grad_features + 56 * i;
        for (int ch = 0; ch < 3; ++ch)
            for (int k = keep; k < 16; ++k) g[8 + 16 * ch + k] = 0;
        for (int k = 0; k < 4; ++k) g[k] *= q_factor;
        for (int k = 4; k < 7; ++k) g[k] *= s_factor;
        g[7] *= alpha_factor;
        for (int ch = 0; ch < 3; ++ch) {
            g[8 + 16 * ch] *= color_factor;
            for (int k = 1; k < 16; ++k) g[8 + 16 * ch + k] *= high_order_color_factor;
        }
    }
}
