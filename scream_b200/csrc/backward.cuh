// backward.cuh -- gradient scatter through the winner map.
// The reference has no backward source: autograd differentiates render.py:31,42-49,68-69.  For pixel p
// of cloud c in view v with winner i and upstream gradient G:
//   g      = G / 0.5                          (render.py:69; skipped with SCREAM_F_RAW)
//   d pv   = g * w_ip          d z'  = -d pv / range        (zmin, range are detached host scalars)
//   d d2   = g * pv * w_ip * (-rho^2 / 2)
//   d x'   = 2 (x' - cx) d d2     d y' = 2 (y' - cy) d d2
//   d P   += R_v^T (d x', d y', d z')         (render.py:68)   [then R_pose^T for a posed source point]
// One thread per (item, view, cloud, pixel); fp32 RED.ADD to the winner's row of grad_xyz.
#pragma once
#include "common.cuh"

namespace scream {

struct BwdParams {
    const float* xyz; const int64_t* offs; const float* pose; const float* views;
    int B, V, w;
    float negk;
    unsigned flags;
    const float* grad_imgs; const int32_t* winner; const float* zrange;
    float* grad_xyz;
    // pair mode (offs == nullptr; see FwdParams): virtual bases of the target cloud and of its gradient, sizes by value
    const float* xyz_t; float* grad_t; int64_t pair_n, pair_m;
};

constexpr int BW_THREADS = 256;

// gradient of one pixel w.r.t. its winner point (in the frame of xyz: through R_v^T and, for a posed source point, R_pose^T)
struct PixelGrad { float g0, g1, g2; };
__device__ __forceinline__ PixelGrad pixel_grad(const BwdParams& P, const float* __restrict__ R, const float* __restrict__ pose, bool posed,
                                                float x, float y, float z, float G, float zmin, float range, int row, int col, int half)
{
    if (posed) apply_pose(pose, x, y, z);
    const float xr = rot_row(R, x, y, z), yr = rot_row(R + 3, x, y, z), zr = rot_row(R + 6, x, y, z);
    const float pv = depth_weight_m(zr, zmin, range, (P.flags & 64u) != 0);   // 64u: SCREAM_F_HALF
    const float dx = __fsub_rn(xr, pixel_centre(col, half)), dy = __fsub_rn(yr, pixel_centre(row, half));
    const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    const float wgt = expf(__fmul_rn(d2, P.negk));
    const float gz = -(G * wgt) / range;
    const float dd2 = G * pv * wgt * P.negk;
    const float gx = 2.0f * dx * dd2, gy = 2.0f * dy * dd2;
    PixelGrad g;                                          // R_v^T g
    g.g0 = R[0] * gx + R[3] * gy + R[6] * gz;
    g.g1 = R[1] * gx + R[4] * gy + R[7] * gz;
    g.g2 = R[2] * gx + R[5] * gy + R[8] * gz;
    if (posed) {
        const float h0 = pose[0] * g.g0 + pose[4] * g.g1 + pose[8] * g.g2;
        const float h1 = pose[1] * g.g0 + pose[5] * g.g1 + pose[9] * g.g2;
        const float h2 = pose[2] * g.g0 + pose[6] * g.g1 + pose[10] * g.g2;
        g.g0 = h0; g.g1 = h1; g.g2 = h2;
    }
    return g;
}

__global__ void __launch_bounds__(BW_THREADS)
backward_scatter_kernel(BwdParams P)
{
    const int w = P.w, npix = w * w, half = w / 2;
    const int per_img = npix / BW_THREADS;                 // w % 64 == 0  ->  npix % 256 == 0
    int bid = blockIdx.x;
    const int part = bid % per_img; bid /= per_img;
    const int c = bid % 2; bid /= 2;
    const int v = bid % P.V;
    const int b = bid / P.V;

    __shared__ float s_R[9], s_pose[12];
    if (threadIdx.x < 9) s_R[threadIdx.x] = P.views[(size_t)v * 9 + threadIdx.x];
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && threadIdx.x >= 32 && threadIdx.x < 44) s_pose[threadIdx.x - 32] = P.pose[(size_t)b * 12 + threadIdx.x - 32];
    __syncthreads();

    const int p = part * BW_THREADS + threadIdx.x;
    const size_t o = (((size_t)b * P.V + v) * 2 + c) * npix + p;
    const int wi = P.winner[o];
    float G = P.grad_imgs[o];
    if (wi < 0 || G == 0.0f) return;
    if (!(P.flags & 1u)) G *= 2.0f;

    const ItemSpan sp = item_span(P.offs, b, P.pair_n, P.pair_m);
    const int64_t gi = (c ? sp.tgt0 : sp.src0) + wi;
    const float* xyz = (c && P.xyz_t != nullptr) ? P.xyz_t : P.xyz;
    const float zmin = P.zrange[((size_t)b * P.V + v) * 2], zmax = P.zrange[((size_t)b * P.V + v) * 2 + 1];
    const PixelGrad g = pixel_grad(P, s_R, s_pose, posed, xyz[gi * 3], xyz[gi * 3 + 1], xyz[gi * 3 + 2], G, zmin, z_range(zmin, zmax), p / w, p % w, half);
    float* dst = ((c && P.grad_t != nullptr) ? P.grad_t : P.grad_xyz) + gi * 3;
    atomicAdd(dst, g.g0); atomicAdd(dst + 1, g.g1); atomicAdd(dst + 2, g.g2);
}

// The same sums with the atomics in SHARED memory: one CTA per (item, cloud) walks all V views' pixels of that cloud, adds
// every pixel's gradient into a shared-memory copy of the cloud's gradient (3 floats per point), and writes it out once,
// coalesced.  75 M global RED.ADD of a 512-pair step (the scatter kernel sits at 73 % of the L2's throughput) become plain
// stores.  For grids that fill the GPU (B * 2 >= the SM count); a cloud beyond BG_CAP points takes the scatter's atomics here.
constexpr int BG_THREADS = 1024;
constexpr int BG_CAP = 18432;                             // points whose gradient fits 216 KB of shared memory

constexpr int BG_MAXV = 64;                               // views whose constants are staged at once

__global__ void __launch_bounds__(BG_THREADS, 1)
backward_gather_kernel(BwdParams P, int cap)
{
    extern __shared__ __align__(16) float s_acc[];       // [3 * npts]
    __shared__ float s_Rv[BG_MAXV][12];                  // R_v, zmin, range, -
    __shared__ float s_pose[12];
    const int w = P.w, npix = w * w, half = w / 2;
    const int c = blockIdx.x & 1, b = blockIdx.x >> 1;
    const ItemSpan sp = item_span(P.offs, b, P.pair_n, P.pair_m);
    const int64_t lo = c ? sp.tgt0 : sp.src0, hi = c ? sp.end : sp.tgt0;
    const int64_t npts = hi - lo;
    const bool local = npts <= (int64_t)cap;              // uniform over the CTA
    const float* xyz = ((c && P.xyz_t != nullptr) ? P.xyz_t : P.xyz) + lo * 3;
    float* dst = ((c && P.grad_t != nullptr) ? P.grad_t : P.grad_xyz) + lo * 3;
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && threadIdx.x < 12) s_pose[threadIdx.x] = P.pose[(size_t)b * 12 + threadIdx.x];
    // (the caller does not clear grad_xyz for this kernel: a cloud too large for shared memory clears its own rows first)
    if (local) for (int t = threadIdx.x; t < 3 * (int)npts; t += BG_THREADS) s_acc[t] = 0.0f;
    else for (int64_t t = threadIdx.x; t < 3 * npts; t += BG_THREADS) dst[t] = 0.0f;
    for (int v0 = 0; v0 < P.V; v0 += BG_MAXV) {
        const int nv = min(BG_MAXV, P.V - v0);
        __syncthreads();
        for (int t = threadIdx.x; t < nv * 12; t += BG_THREADS) {
            const int vv = t / 12, k = t % 12, v = v0 + vv;
            const float zmin = P.zrange[((size_t)b * P.V + v) * 2], zmax = P.zrange[((size_t)b * P.V + v) * 2 + 1];
            s_Rv[vv][k] = k < 9 ? P.views[(size_t)v * 9 + k] : (k == 9 ? zmin : (k == 10 ? z_range(zmin, zmax) : 0.0f));
        }
        __syncthreads();
        const size_t o0 = (((size_t)b * P.V + v0) * 2 + c) * npix;       // view v0 + vv: + vv * 2 * npix
        // Three pixels per thread are in flight: the one whose gradient is computed, the next one's winner coordinates (a
        // dependent gather: a third of the stall samples when exposed), the one after's winner / upstream gradient (a fourth,
        // with two gathers ahead, measured slower).  Pixel
        // (view, row, column) follows t = thread + 1024 i incrementally (no division in the loop).
        const int total = nv * npix, lane = threadIdx.x & 31;
        const int dcol = BG_THREADS % w, drow = BG_THREADS / w;
        int vv = 0, row = (int)threadIdx.x / w, col = (int)threadIdx.x % w;      // of the pixel being computed (npix >= 4096 > thread)
        int wi1 = -1, wi2 = -1; float G1 = 0.0f, G2 = 0.0f, x1 = 0.0f, y1 = 0.0f, z1 = 0.0f;
        auto load_wg = [&](int t, int& wi, float& G) {
            wi = -1; G = 0.0f;
            if (t < total) { const size_t o = o0 + (size_t)(t / npix) * npix + t; wi = __ldg(P.winner + o); G = __ldg(P.grad_imgs + o); }
        };
        auto gather = [&](int wi, float G, float& x, float& y, float& z) {
            x = y = z = 0.0f;
            if (wi >= 0 && G != 0.0f) { const float* q = xyz + (size_t)wi * 3; x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2); }
        };
        load_wg(threadIdx.x, wi1, G1);
        load_wg(threadIdx.x + BG_THREADS, wi2, G2);
        gather(wi1, G1, x1, y1, z1);
        for (int t = threadIdx.x; t < total; t += BG_THREADS) {          // uniform over the warp (npix is a multiple of the CTA size)
            const int wi0 = wi1; const float G0 = G1, x0 = x1, y0 = y1, z0 = z1;
            wi1 = wi2; G1 = G2;
            gather(wi1, G1, x1, y1, z1);
            load_wg(t + 2 * BG_THREADS, wi2, G2);
            const bool valid = wi0 >= 0 && G0 != 0.0f;
            PixelGrad g = {0.0f, 0.0f, 0.0f};
            if (valid) {
                const float Gs = (P.flags & 1u) ? G0 : G0 * 2.0f;
                g = pixel_grad(P, s_Rv[vv], s_pose, posed, x0, y0, z0, Gs, s_Rv[vv][9], s_Rv[vv][10], row, col, half);
            }
            // Neighbouring pixels often share their winner, and lanes of a warp that add to the same word make its
            // compare-and-swap loop retry once per lane: runs of equal winners are summed over the warp first (segmented
            // scan: 15 shuffles) and the last lane of a run adds the total.
            const int key = valid ? wi0 : -1 - lane;                  // invalid lanes: runs of their own
            const int prev = __shfl_up_sync(0xFFFFFFFFu, key, 1), next = __shfl_down_sync(0xFFFFFFFFu, key, 1);
            const unsigned heads = __ballot_sync(0xFFFFFFFFu, lane == 0 || prev != key);
            const int start = 31 - __clz(heads & (0xFFFFFFFFu >> (31 - lane)));
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const float a0 = __shfl_up_sync(0xFFFFFFFFu, g.g0, d), a1 = __shfl_up_sync(0xFFFFFFFFu, g.g1, d), a2 = __shfl_up_sync(0xFFFFFFFFu, g.g2, d);
                if (lane - d >= start) { g.g0 += a0; g.g1 += a1; g.g2 += a2; }
            }
            if (valid && (lane == 31 || next != key)) {
                // (shared memory has no native fp32 add: each of these is a compare-and-swap loop; one 128-bit CAS per point was slower)
                if (local) { atomicAdd(s_acc + 3 * wi0, g.g0); atomicAdd(s_acc + 3 * wi0 + 1, g.g1); atomicAdd(s_acc + 3 * wi0 + 2, g.g2); }
                else { float* a = dst + (size_t)wi0 * 3; atomicAdd(a, g.g0); atomicAdd(a + 1, g.g1); atomicAdd(a + 2, g.g2); }
            }
            // the next pixel of this thread: 1024 further on, possibly in the next view
            col += dcol; row += drow;
            if (col >= w) { col -= w; ++row; }
            if (row >= w) { row -= w; ++vv; }
        }
    }
    if (local) {
        __syncthreads();
        for (int t = threadIdx.x; t < 3 * (int)npts; t += BG_THREADS) dst[t] = s_acc[t];
    }
}

}  // namespace scream
