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
                                                float x, float y, float z, float G, float zmin, float range, int p, int w, int half)
{
    if (posed) apply_pose(pose, x, y, z);
    const float xr = rot_row(R, x, y, z), yr = rot_row(R + 3, x, y, z), zr = rot_row(R + 6, x, y, z);
    const float pv = depth_weight_m(zr, zmin, range, (P.flags & 64u) != 0);   // 64u: SCREAM_F_HALF
    const float dx = __fsub_rn(xr, pixel_centre(p % w, half)), dy = __fsub_rn(yr, pixel_centre(p / w, half));
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
    const PixelGrad g = pixel_grad(P, s_R, s_pose, posed, xyz[gi * 3], xyz[gi * 3 + 1], xyz[gi * 3 + 2], G, zmin, z_range(zmin, zmax), p, w, half);
    float* dst = ((c && P.grad_t != nullptr) ? P.grad_t : P.grad_xyz) + gi * 3;
    atomicAdd(dst, g.g0); atomicAdd(dst + 1, g.g1); atomicAdd(dst + 2, g.g2);
}

// The same sums with the atomics in SHARED memory: one CTA per (item, cloud) walks all V views' pixels of that cloud, adds
// every pixel's gradient into a shared-memory copy of the cloud's gradient (3 floats per point), and writes it out once,
// coalesced.  75 M global RED.ADD of a 512-pair step (the scatter kernel sits at 73 % of the L2's throughput) become plain
// stores.  For grids that fill the GPU (B * 2 >= the SM count); a cloud beyond BG_CAP points takes the scatter's atomics here.
constexpr int BG_THREADS = 1024;
constexpr int BG_CAP = 18432;                             // points whose gradient fits 216 KB of shared memory

#ifndef BG_FLIGHT_N
#define BG_FLIGHT_N 1
#endif
constexpr int BG_FLIGHT = BG_FLIGHT_N;                              // pixels in flight per thread
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
        // BG_FLIGHT pixels per thread and trip: their winner / gradient loads, then the winners' coordinates (a dependent
        // gather), are in flight together -- one CTA per SM has 24 serial trips of two L2 round trips otherwise
        for (int t0 = threadIdx.x; t0 < nv * npix; t0 += BG_FLIGHT * BG_THREADS) {
            int wi[BG_FLIGHT]; float G[BG_FLIGHT], x[BG_FLIGHT], y[BG_FLIGHT], z[BG_FLIGHT];
#pragma unroll
            for (int u = 0; u < BG_FLIGHT; ++u) {
                const int t = t0 + u * BG_THREADS;
                wi[u] = -1; G[u] = 0.0f;
                if (t < nv * npix) {
                    const int vv = t / npix, p = t - vv * npix;
                    const size_t o = o0 + (size_t)vv * 2 * npix + p;
                    wi[u] = __ldg(P.winner + o); G[u] = __ldg(P.grad_imgs + o);
                }
            }
#pragma unroll
            for (int u = 0; u < BG_FLIGHT; ++u) {
                x[u] = y[u] = z[u] = 0.0f;
                if (wi[u] >= 0 && G[u] != 0.0f) { const float* q = xyz + (size_t)wi[u] * 3; x[u] = __ldg(q); y[u] = __ldg(q + 1); z[u] = __ldg(q + 2); }
            }
#pragma unroll
            for (int u = 0; u < BG_FLIGHT; ++u) {
                if (t0 + u * BG_THREADS >= nv * npix) break;          // uniform over the warp (npix is a multiple of the CTA size)
                const bool valid = wi[u] >= 0 && G[u] != 0.0f;
                PixelGrad g = {0.0f, 0.0f, 0.0f};
                if (valid) {
                    const int t = t0 + u * BG_THREADS, vv = t / npix, p = t - vv * npix;
                    const float Gs = (P.flags & 1u) ? G[u] : G[u] * 2.0f;
                    g = pixel_grad(P, s_Rv[vv], s_pose, posed, x[u], y[u], z[u], Gs, s_Rv[vv][9], s_Rv[vv][10], p, w, half);
                }
                // Neighbouring pixels often share their winner, and lanes of a warp that add to the same word make its
                // compare-and-swap loop retry once per lane: runs of equal winners are summed over the warp first (segmented
                // scan: 15 shuffles) and the last lane of a run adds the total.
                const int key = valid ? wi[u] : -1 - (int)(threadIdx.x & 31);           // invalid lanes: runs of their own
                const int prev = __shfl_up_sync(0xFFFFFFFFu, key, 1), next = __shfl_down_sync(0xFFFFFFFFu, key, 1);
                const int lane = threadIdx.x & 31;
                const unsigned heads = __ballot_sync(0xFFFFFFFFu, lane == 0 || prev != key);
                const int start = 31 - __clz(heads & (0xFFFFFFFFu >> (31 - lane)));
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const float o0 = __shfl_up_sync(0xFFFFFFFFu, g.g0, d), o1 = __shfl_up_sync(0xFFFFFFFFu, g.g1, d), o2 = __shfl_up_sync(0xFFFFFFFFu, g.g2, d);
                    if (lane - d >= start) { g.g0 += o0; g.g1 += o1; g.g2 += o2; }
                }
                if (valid && (lane == 31 || next != key)) {
                    // (shared memory has no native fp32 add: each of these is a compare-and-swap loop; one 128-bit CAS per point was slower)
                    if (local) { atomicAdd(s_acc + 3 * wi[u], g.g0); atomicAdd(s_acc + 3 * wi[u] + 1, g.g1); atomicAdd(s_acc + 3 * wi[u] + 2, g.g2); }
                    else { float* a = dst + (size_t)wi[u] * 3; atomicAdd(a, g.g0); atomicAdd(a + 1, g.g1); atomicAdd(a + 2, g.g2); }
                }
            }
        }
    }
    if (local) {
        __syncthreads();
        for (int t = threadIdx.x; t < 3 * (int)npts; t += BG_THREADS) dst[t] = s_acc[t];
    }
}

}  // namespace scream
