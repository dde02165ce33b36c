/* Plain-C restatement of SCREAM's scene renderer -- TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library, and only as the checker / the timed CPU baseline.  Nothing under
 * scream_b200/ links or calls it.
 *
 * It restates /root/reference/models/render.py:27-73 one scalar operation at a time, each
 * rounded to fp32 exactly where the reference's separate ATen kernels round:
 *   rotation      render.py:68   (R @ P.T).T is an MKL sgemm; for N >= 15 columns each output is
 *                                fmaf(r2, z, fmaf(r1, y, r0 * x))   [measured, SURVEY 7.3.3]
 *   pose          train_3d_match.py:198-199   same sgemm chain, then a separately rounded + t
 *   z range       render.py:30-31   min / max over ALL n+m points; zmax - zmin in Python double,
 *                                then rounded to fp32 when it meets the fp32 tensor
 *   pv            render.py:31   1 - (z - zmin) / range   (true fp32 division)
 *   pixel centre  render.py:14-16   ((j - w//2) + 0.5) / (w//2)
 *   weight        render.py:42-43   dx = x - cx; d2 = dx*dx + dy*dy; arg = (-d2 / 2) * rho^2
 *   value         render.py:46,49   pv * exp(arg); max over points, first index on ties
 *   output        render.py:69   (img - 0.5) / 0.5
 * exp() is glibc expf here, an opaque <=1 ulp routine in the reference (ATen/MKL): values are
 * compared with a 1e-6 relative tolerance, everything before exp() bit for bit.
 * This file is also the scalar model of the arithmetic the CUDA kernels perform.
 *
 * Build: see oracle/Makefile  (-ffp-contract=off is REQUIRED: no implicit FMA contraction).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

static inline float rot_row(const float *r, float x, float y, float z)
{
    return fmaf(r[2], z, fmaf(r[1], y, r[0] * x));
}

int scream_oracle_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Transform stage only: rotated [V][N][3], pv [V][N], zrange [V][2].  Any output may be NULL. */
int scream_oracle_transform(const float *src, int64_t n, const float *tgt, int64_t m,
                            const float *pose, const float *views, int V,
                            float *rotated, float *pv, float *zrange)
{
    const int64_t N = n + m;
    if (n <= 0 || m <= 0 || V <= 0) return -1;
    float *P = (float *)malloc((size_t)N * 3 * sizeof(float));
    if (!P) return -2;
    for (int64_t i = 0; i < N; ++i) {
        const float *p = i < n ? src + 3 * i : tgt + 3 * (i - n);
        float x = p[0], y = p[1], z = p[2];
        if (pose && i < n) {                       /* train_3d_match.py:198 */
            float a = rot_row(pose + 0, x, y, z) + pose[3];
            float b = rot_row(pose + 4, x, y, z) + pose[7];
            float c = rot_row(pose + 8, x, y, z) + pose[11];
            x = a; y = b; z = c;
        }
        P[3 * i] = x; P[3 * i + 1] = y; P[3 * i + 2] = z;
    }
    for (int v = 0; v < V; ++v) {
        const float *R = views + 9 * v;
        float zmin = INFINITY, zmax = -INFINITY;
        for (int64_t i = 0; i < N; ++i) {
            float zr = rot_row(R + 6, P[3 * i], P[3 * i + 1], P[3 * i + 2]);
            if (zr < zmin) zmin = zr;
            if (zr > zmax) zmax = zr;
        }
        float range = (float)((double)zmax - (double)zmin);   /* render.py:31 */
        if (zrange) { zrange[2 * v] = zmin; zrange[2 * v + 1] = zmax; }
        for (int64_t i = 0; i < N; ++i) {
            float x = P[3 * i], y = P[3 * i + 1], z = P[3 * i + 2];
            float xr = rot_row(R + 0, x, y, z), yr = rot_row(R + 3, x, y, z), zr = rot_row(R + 6, x, y, z);
            if (rotated) {
                float *o = rotated + ((size_t)v * N + i) * 3;
                o[0] = xr; o[1] = yr; o[2] = zr;
            }
            if (pv) pv[(size_t)v * N + i] = 1.0f - (zr - zmin) / range;
        }
    }
    free(P);
    return 0;
}

/* Dense forward.  val/idx/second: [V][2][w][w]; imgs (normalised) may be NULL; second may be NULL. */
int scream_oracle_forward(const float *src, int64_t n, const float *tgt, int64_t m,
                          const float *pose, const float *views, int V, int w, double rho,
                          float *imgs, float *val, int32_t *idx, float *second, float *zrange)
{
    const int64_t N = n + m;
    if (n <= 0 || m <= 0 || V <= 0 || w <= 0) return -1;
    const int half = w / 2;
    const float rho2 = (float)(rho * rho);
    const size_t npix = (size_t)w * w;
    float *rot = (float *)malloc((size_t)V * N * 3 * sizeof(float));
    float *pv = (float *)malloc((size_t)V * N * sizeof(float));
    float *cen = (float *)malloc((size_t)w * sizeof(float));
    if (!rot || !pv || !cen) { free(rot); free(pv); free(cen); return -2; }
    int rc = scream_oracle_transform(src, n, tgt, m, pose, views, V, rot, pv, zrange);
    if (rc) { free(rot); free(pv); free(cen); return rc; }
    for (int j = 0; j < w; ++j) cen[j] = (((float)j - (float)half) + 0.5f) / (float)half;

    for (int v = 0; v < V; ++v) {
        const float *X = rot + (size_t)v * N * 3;
        const float *W = pv + (size_t)v * N;
        for (int c = 0; c < 2; ++c) {
            const int64_t lo = c ? n : 0, hi = c ? N : n;
#pragma omp parallel for schedule(static)
            for (int64_t p = 0; p < (int64_t)npix; ++p) {
                const float cx = cen[p % w], cy = cen[p / w];
                float best = -INFINITY, runner = -INFINITY;
                int32_t arg = 0;
                for (int64_t i = lo; i < hi; ++i) {
                    float dx = X[3 * i] - cx, dy = X[3 * i + 1] - cy;
                    float d2 = dx * dx + dy * dy;
                    float a = (-d2 / 2.0f) * rho2;
                    float t = W[i] * expf(a);
                    if (t > best) { runner = best; best = t; arg = (int32_t)(i - lo); }
                    else if (t > runner) runner = t;
                }
                const size_t o = ((size_t)v * 2 + c) * npix + (size_t)p;
                if (val) val[o] = best;
                if (idx) idx[o] = arg;
                if (second) second[o] = runner;
                if (imgs) imgs[o] = (best - 0.5f) / 0.5f;
            }
        }
    }
    free(rot); free(pv); free(cen);
    return 0;
}

/* Closed-form backward through a winner map, fp64 accumulation, single thread.
 * grad_src [n][3], grad_tgt [m][3] (doubles, overwritten). */
int scream_oracle_backward(const float *src, int64_t n, const float *tgt, int64_t m,
                           const float *pose, const float *views, int V, int w, double rho,
                           const float *grad_imgs, const int32_t *winner, const float *zrange,
                           double *grad_src, double *grad_tgt)
{
    const int64_t N = n + m;
    if (n <= 0 || m <= 0 || V <= 0 || w <= 0) return -1;
    const int half = w / 2;
    const double k = rho * rho / 2.0;
    const size_t npix = (size_t)w * w;
    float *rot = (float *)malloc((size_t)V * N * 3 * sizeof(float));
    if (!rot) return -2;
    int rc = scream_oracle_transform(src, n, tgt, m, pose, views, V, rot, NULL, NULL);
    if (rc) { free(rot); return rc; }
    memset(grad_src, 0, (size_t)n * 3 * sizeof(double));
    memset(grad_tgt, 0, (size_t)m * 3 * sizeof(double));
    for (int v = 0; v < V; ++v) {
        const float *R = views + 9 * v;
        const float *X = rot + (size_t)v * N * 3;
        const double zmin = zrange[2 * v], range = (double)zrange[2 * v + 1] - zmin;
        for (int c = 0; c < 2; ++c) {
            for (size_t p = 0; p < npix; ++p) {
                const size_t o = ((size_t)v * 2 + c) * npix + p;
                const int64_t i = winner[o] + (c ? n : 0);
                const double g = 2.0 * grad_imgs[o];
                const double cx = ((double)(p % w) - half + 0.5) / half, cy = ((double)(p / w) - half + 0.5) / half;
                const double dx = X[3 * i] - cx, dy = X[3 * i + 1] - cy;
                const double wg = exp(-(dx * dx + dy * dy) * k);
                const double pvi = 1.0 - (X[3 * i + 2] - zmin) / range;
                const double dd2 = g * pvi * wg * (-k);
                const double gx = 2 * dx * dd2, gy = 2 * dy * dd2, gz = -(g * wg) / range;
                double gp[3];
                for (int j = 0; j < 3; ++j) gp[j] = gx * R[j] + gy * R[3 + j] + gz * R[6 + j];
                if (c == 0 && pose) {              /* through the rigid transform of the source */
                    double q[3];
                    for (int j = 0; j < 3; ++j) q[j] = gp[0] * pose[j] + gp[1] * pose[4 + j] + gp[2] * pose[8 + j];
                    gp[0] = q[0]; gp[1] = q[1]; gp[2] = q[2];
                }
                double *dst = c ? grad_tgt + 3 * (i - n) : grad_src + 3 * i;
                dst[0] += gp[0]; dst[1] += gp[1]; dst[2] += gp[2];
            }
        }
    }
    free(rot);
    return 0;
}
