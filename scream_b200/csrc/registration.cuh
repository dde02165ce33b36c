// registration.cuh -- the step that follows the generator: nearest target of every predicted source point
// and the weighted Kabsch fit, on the device and without the reference's two host round trips.
//   pointnet.py:71    square_distance(src_ / s, tgt / s)[0].min(dim=1)        (utils.py:72-78)
//   pointnet.py:72-74 valid = dis < 0.075;  rigid_transform_3d(src[valid], tgt[ind[valid]])   (utils.py:138-178; the boolean
//                     mask is a device->host sync, utils.py:169 runs the 3x3 SVD on the CPU: a second one)
// Arithmetic of the distance (bit-exact, so that the arg-min and the `< threshold` mask are the reference's):
//   a = fl(p / s);  dot = fmaf(a2, b2, fmaf(a1, b1, a0 * b0))   (sgemm, K = 3)      utils.py:75
//   |a|^2 = fl(fl(a0^2 + a1^2) + a2^2)                                                utils.py:76-77
//   dist = fl(fl(fl(-2 * dot) + |a|^2) + |b|^2);  min keeps the first index on ties   utils.py:75-77, torch.min
// The fit accumulates its 16 sums in fp64 (the reference: fp32 sums, CPU LAPACK SVD) -- a value-level match (1e-5), like
// every floating-point reduction whose order the reference does not define.
#pragma once
#include "common.cuh"

namespace scream {

constexpr int NN_THREADS = 256;
constexpr int NN_TILE = 1024;          // targets per CTA (16 KB of float4 in shared memory)
constexpr int NN_PPT = 2;              // source points per thread

__global__ void nn_init_kernel(unsigned long long* __restrict__ best, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) best[i] = ~0ull;
}

// grid (ceil(n / (256 * NN_PPT)), ceil(m / NN_TILE)); NN_PPT source points per thread (one shared-memory read serves NN_PPT
// pairs), the CTA's tile of targets broadcast from shared memory.  Per pair: 1/NN_PPT LDS.128 + FMUL + 3 FFMA + FADD + FSETP +
// 2 SEL.  fl(fl(-2 * dot) + |a|^2) == fmaf(-2, dot, |a|^2): the product by -2 is exact, so the sum is rounded once either way.
__global__ void __launch_bounds__(NN_THREADS)
nn_search_kernel(const float* __restrict__ src, int64_t n, const float* __restrict__ tgt, int64_t m, float s,
                 unsigned long long* __restrict__ best)
{
    __shared__ float4 s_b[NN_TILE];
    const int64_t j0 = (int64_t)blockIdx.y * NN_TILE;
    const int cnt = (int)min((int64_t)NN_TILE, m - j0);
    for (int j = threadIdx.x; j < cnt; j += NN_THREADS) {
        const float* q = tgt + (size_t)(j0 + j) * 3;
        const float x = __fdiv_rn(__ldg(q), s), y = __fdiv_rn(__ldg(q + 1), s), z = __fdiv_rn(__ldg(q + 2), s);
        s_b[j] = make_float4(x, y, z, __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
    }
    __syncthreads();
    float ax[NN_PPT], ay[NN_PPT], az[NN_PPT], na[NN_PPT], bd[NN_PPT];
    int bj[NN_PPT];
    int64_t ip[NN_PPT];
#pragma unroll
    for (int u = 0; u < NN_PPT; ++u) {
        ip[u] = ((int64_t)blockIdx.x * NN_PPT + u) * NN_THREADS + threadIdx.x;
        const float* p = src + (size_t)min(ip[u], n - 1) * 3;             // past the end: the last point again (not stored)
        ax[u] = __fdiv_rn(__ldg(p), s); ay[u] = __fdiv_rn(__ldg(p + 1), s); az[u] = __fdiv_rn(__ldg(p + 2), s);
        na[u] = __fadd_rn(__fadd_rn(__fmul_rn(ax[u], ax[u]), __fmul_rn(ay[u], ay[u])), __fmul_rn(az[u], az[u]));
        bd[u] = INFINITY; bj[u] = -1;
    }
#pragma unroll 4
    for (int j = 0; j < cnt; ++j) {
        const float4 b = s_b[j];
#pragma unroll
        for (int u = 0; u < NN_PPT; ++u) {
            const float dot = __fmaf_rn(az[u], b.z, __fmaf_rn(ay[u], b.y, __fmul_rn(ax[u], b.x)));
            const float d = __fadd_rn(__fmaf_rn(-2.0f, dot, na[u]), b.w);
            if (d < bd[u]) { bd[u] = d; bj[u] = j; }                      // strict: the first index keeps ties; NaN never wins
        }
    }
#pragma unroll
    for (int u = 0; u < NN_PPT; ++u) {
        if (ip[u] >= n) continue;
        // nothing below +inf in this tile (overflowing or NaN coordinates): +inf at the tile's first index, as torch.min would
        // (-0) + (+0) = +0: torch.min sees the two zeros as equal (first index wins), the integer order below does not
        const unsigned long long key = ((unsigned long long)enc_f32(__fadd_rn(bd[u], 0.0f)) << 32) | (unsigned)(j0 + max(bj[u], 0));
        atomicMin(best + ip[u], key);
    }
}

__global__ void nn_finish_kernel(const unsigned long long* __restrict__ best, int64_t n,
                                 float* __restrict__ dist, int64_t* __restrict__ ind)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long k = best[i];
    if (dist) dist[i] = dec_f32((unsigned)(k >> 32));
    if (ind) ind[i] = (int64_t)(k & 0xFFFFFFFFull);
}

// ---- weighted Kabsch, utils.py:138-178 ---------------------------------------------------------------------------------
// sums per batch item: [0] W, [1..3] sum w a, [4..6] sum w b, [7..15] sum w a b^T (row-major)
constexpr int KB_THREADS = 256;

__global__ void __launch_bounds__(KB_THREADS)
kabsch_accum_kernel(const float* __restrict__ A, const float* __restrict__ Bp, const int64_t* __restrict__ b_index,
                    const float* __restrict__ weights, const float* __restrict__ valid_dist, float valid_thr,
                    float weight_threshold, int64_t k, int64_t b_rows, double* __restrict__ acc)
{
    const int bt = blockIdx.y;
    double v[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) v[t] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * KB_THREADS + threadIdx.x; i < k; i += (int64_t)gridDim.x * KB_THREADS) {
        const int64_t r = (int64_t)bt * k + i;
        float w = weights ? weights[r] : 1.0f;
        if (w < weight_threshold) w = 0.0f;                                  // utils.py:151
        if (valid_dist && !(valid_dist[r] < valid_thr)) w = 0.0f;            // pointnet.py:72 as a 0/1 weight
        if (w == 0.0f) continue;
        const int64_t rb = b_index ? b_index[r] : i;
        const float* a = A + (size_t)r * 3;
        const float* b = Bp + ((size_t)bt * b_rows + rb) * 3;
        const double a0 = a[0], a1 = a[1], a2 = a[2], b0 = b[0], b1 = b[1], b2 = b[2], wd = w;
        v[0] += wd;
        v[1] += wd * a0; v[2] += wd * a1; v[3] += wd * a2;
        v[4] += wd * b0; v[5] += wd * b1; v[6] += wd * b2;
        v[7] += wd * a0 * b0; v[8] += wd * a0 * b1; v[9] += wd * a0 * b2;
        v[10] += wd * a1 * b0; v[11] += wd * a1 * b1; v[12] += wd * a1 * b2;
        v[13] += wd * a2 * b0; v[14] += wd * a2 * b1; v[15] += wd * a2 * b2;
    }
    __shared__ double s_acc[16];
    if (threadIdx.x < 16) s_acc[threadIdx.x] = 0.0;
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 16; ++t) {
        double x = v[t];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[t], x);
    }
    __syncthreads();
    if (threadIdx.x < 16) atomicAdd(acc + (size_t)bt * 16 + threadIdx.x, s_acc[threadIdx.x]);
}

// one thread per batch item: H, its SVD by one-sided Jacobi rotations (H V = U S), R = V diag(1,1,det(V U^T)) U^T with the
// sign on the SMALLEST singular value (torch.svd sorts them in descending order, utils.py:169-174), t = cb - R ca.
__global__ void kabsch_solve_kernel(const double* __restrict__ acc, int bs, float* __restrict__ T)
{
    const int bt = blockIdx.x * blockDim.x + threadIdx.x;
    if (bt >= bs) return;
    const double* s = acc + (size_t)bt * 16;
    const double W = s[0], den = W + 1e-6;                                   // utils.py:155-158
    double ca[3], cb[3], G[3][3], V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int i = 0; i < 3; ++i) { ca[i] = s[1 + i] / den; cb[i] = s[4 + i] / den; }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            G[i][j] = s[7 + 3 * i + j] - ca[i] * s[4 + j] - s[1 + i] * cb[j] + W * ca[i] * cb[j];
    for (int sweep = 0; sweep < 40; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double al = 0, be = 0, ga = 0;
                for (int i = 0; i < 3; ++i) { al += G[i][p] * G[i][p]; be += G[i][q] * G[i][q]; ga += G[i][p] * G[i][q]; }
                if (ga == 0.0 || fabs(ga) <= 1e-300 + 1e-17 * sqrt(al * be)) continue;
                off = fmax(off, fabs(ga) / sqrt(al * be));
                const double zeta = (be - al) / (2.0 * ga);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
                for (int i = 0; i < 3; ++i) {
                    const double gp = G[i][p], gq = G[i][q], vp = V[i][p], vq = V[i][q];
                    G[i][p] = c * gp - sn * gq; G[i][q] = sn * gp + c * gq;
                    V[i][p] = c * vp - sn * vq; V[i][q] = sn * vp + c * vq;
                }
            }
        if (off < 1e-15) break;
    }
    double S[3], U[3][3];
    int lo = 0, hi = 0;
    for (int j = 0; j < 3; ++j) {
        S[j] = sqrt(G[0][j] * G[0][j] + G[1][j] * G[1][j] + G[2][j] * G[2][j]);
        if (S[j] < S[lo]) lo = j;
        if (S[j] > S[hi]) hi = j;
    }
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) U[i][j] = S[j] > 0 ? G[i][j] / S[j] : 0.0;
    if (!(S[lo] > 1e-12 * S[hi])) {                 // rank-deficient H: complete U with the cross product of the other two columns
        const int p = (lo + 1) % 3, q = (lo + 2) % 3;
        U[0][lo] = U[1][p] * U[2][q] - U[2][p] * U[1][q];
        U[1][lo] = U[2][p] * U[0][q] - U[0][p] * U[2][q];
        U[2][lo] = U[0][p] * U[1][q] - U[1][p] * U[0][q];
    }
    auto det3 = [](const double M[3][3]) {
        return M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0])
             + M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
    };
    const double d = det3(V) * det3(U);             // det(V U^T), utils.py:171
    double R[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double r = 0;
            for (int c = 0; c < 3; ++c) r += (c == lo ? d : 1.0) * V[i][c] * U[j][c];
            R[i][j] = r;
        }
    float* o = T + (size_t)bt * 16;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) o[4 * i + j] = (float)R[i][j];
        o[4 * i + 3] = (float)(cb[i] - (R[i][0] * ca[0] + R[i][1] * ca[1] + R[i][2] * ca[2]));
    }
    o[12] = 0.f; o[13] = 0.f; o[14] = 0.f; o[15] = 1.f;
}

}  // namespace scream
