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
    float x = xyz[gi * 3], y = xyz[gi * 3 + 1], z = xyz[gi * 3 + 2];
    if (posed) apply_pose(s_pose, x, y, z);
    const float xr = rot_row(s_R, x, y, z), yr = rot_row(s_R + 3, x, y, z), zr = rot_row(s_R + 6, x, y, z);
    const float zmin = P.zrange[((size_t)b * P.V + v) * 2], zmax = P.zrange[((size_t)b * P.V + v) * 2 + 1];
    const float range = z_range(zmin, zmax);
    const float pv = depth_weight_m(zr, zmin, range, (P.flags & 64u) != 0);   // 64u: SCREAM_F_HALF
    const float dx = __fsub_rn(xr, pixel_centre(p % w, half)), dy = __fsub_rn(yr, pixel_centre(p / w, half));
    const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    const float wgt = expf(__fmul_rn(d2, P.negk));

    const float gz = -(G * wgt) / range;
    const float dd2 = G * pv * wgt * P.negk;
    const float gx = 2.0f * dx * dd2, gy = 2.0f * dy * dd2;
    // R_v^T g
    float g0 = s_R[0] * gx + s_R[3] * gy + s_R[6] * gz;
    float g1 = s_R[1] * gx + s_R[4] * gy + s_R[7] * gz;
    float g2 = s_R[2] * gx + s_R[5] * gy + s_R[8] * gz;
    if (posed) {
        float h0 = s_pose[0] * g0 + s_pose[4] * g1 + s_pose[8] * g2;
        float h1 = s_pose[1] * g0 + s_pose[5] * g1 + s_pose[9] * g2;
        float h2 = s_pose[2] * g0 + s_pose[6] * g1 + s_pose[10] * g2;
        g0 = h0; g1 = h1; g2 = h2;
    }
    float* dst = ((c && P.grad_t != nullptr) ? P.grad_t : P.grad_xyz) + gi * 3;
    atomicAdd(dst, g0); atomicAdd(dst + 1, g1); atomicAdd(dst + 2, g2);
}

}  // namespace scream
