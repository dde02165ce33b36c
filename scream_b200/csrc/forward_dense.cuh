// forward_dense.cuh -- dense exact forward (the parity anchor).
//   render.py:39-49  for every pixel, max over ALL points of the cloud of pv_i * exp(-d2/2*rho^2),
//                    first index on ties (torch.max(dim=0)).
// Gather formulation, no atomics: a thread owns 4 horizontally adjacent pixels and scans the cloud in
// index order from a shared-memory stage of transformed points (x', y', pv), so "strictly greater"
// reproduces the lowest-index tie-break.  A pair is skipped only when its squared distance proves its
// value is strictly below the pixel's current best (cull_d2, common.cuh) -- the result is identical to
// evaluating every pair.
#pragma once
#include "common.cuh"

namespace scream {

constexpr int FD_THREADS = 128;
constexpr int FD_PX = 4;                              // pixels per thread (same row)
constexpr int FD_TILE = FD_THREADS * FD_PX;           // 512 pixels per CTA
constexpr int FD_CHUNK = 1024;                        // points staged per round

struct FwdParams {
    const float* xyz; const int64_t* offs; const float* pose; const float* views;
    const unsigned* zenc;
    int B, V, w;
    float negk, inv_negk;
    unsigned flags;
    float* imgs; int32_t* winner; float* zrange;
    const int* surv; const unsigned* surv_cnt; int surv_cap;   // survivor lists of the giant-cloud path (last filter round), or nullptr
    const int* surv1; const unsigned* surv1_cnt; int surv1_cap; // first-round lists: used for an image whose last-round list overflowed
    // culled kernel only:
    const float* xyz_t;          // pair mode (offs == nullptr): VIRTUAL base of the target cloud, tgt - 3 * pair_n floats, so that
    int64_t pair_n, pair_m;      //   xyz_t + 3 * (pair_n + i) is target point i; the source cloud is xyz
    int nparts;                  // CTAs per image (>= 1): CTA `part` owns the 32-pixel chunks c with c % nparts == part
    int fuse_zrange;             // the CTA reduces the depth range of its (item, view) itself (small grids: two launches less)
    int tiny;                    // clouds of up to FC_LINEAR_MAX points are left to forward_tiny_kernel
    int prow0 = 0, pcol0 = 0;    // forward_culled_lattice_kernel: (b, a) = it renders the pixels (2 r + b, 2 c + a) of a w == 128 image, one of
                                 // its four 64 x 64 lattices (giant clouds: one launch per lattice pass, scream_render.cu)
};

__device__ __forceinline__ float zero_or_cull(float best, float inv_negk) {
    // best == +-0: contributions with exp argument below -106 are exactly 0 and can never win or tie.
    // The same radius is used while best is (near-)denormal: there fp32 has no relative precision, so a
    // log-domain margin cannot prove "strictly below"; only "exactly zero" can.
    if (best >= 0.0f && best < 1e-37f) return __fmul_rn(-106.0f, inv_negk);
    return cull_d2(best, inv_negk);
}

__global__ void __launch_bounds__(FD_THREADS)
forward_dense_kernel(FwdParams P)
{
    const int w = P.w, npix = w * w, half = w / 2;
    const int tiles = npix / FD_TILE;
    int bid = blockIdx.x;
    const int tile = bid % tiles; bid /= tiles;
    const int c = bid % 2; bid /= 2;
    const int v = bid % P.V;
    const int b = bid / P.V;

    __shared__ float4 s_pts[FD_CHUNK];
    __shared__ float s_R[9], s_pose[12];
    if (threadIdx.x < 9) s_R[threadIdx.x] = P.views[(size_t)v * 9 + threadIdx.x];
    if (P.pose && threadIdx.x >= 32 && threadIdx.x < 44) s_pose[threadIdx.x - 32] = P.pose[(size_t)b * 12 + threadIdx.x - 32];

    const ItemSpan sp = item_span(P.offs, b);
    const int64_t lo = c ? sp.tgt0 : sp.src0, hi = c ? sp.end : sp.tgt0;
    const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]);
    const float zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
    const float range = z_range(zmin, zmax);
    const bool posed = P.pose != nullptr && c == 0;

    const int p0 = tile * FD_TILE + threadIdx.x * FD_PX;     // first pixel of this thread
    const float cy = pixel_centre(p0 / w, half);
    float cx[FD_PX], best[FD_PX], t2[FD_PX];
    int arg[FD_PX];
#pragma unroll
    for (int u = 0; u < FD_PX; ++u) {
        cx[u] = pixel_centre(p0 % w + u, half);
        best[u] = -INFINITY; t2[u] = INFINITY; arg[u] = 0;
    }
    __syncthreads();

    for (int64_t c0 = lo; c0 < hi; c0 += FD_CHUNK) {
        const int cnt = (int)min((int64_t)FD_CHUNK, hi - c0);
        for (int j = threadIdx.x; j < cnt; j += FD_THREADS) {
            const float* q = P.xyz + (size_t)(c0 + j) * 3;
            float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
            if (posed) apply_pose(s_pose, x, y, z);
            float xr = rot_row(s_R, x, y, z), yr = rot_row(s_R + 3, x, y, z), zr = rot_row(s_R + 6, x, y, z);
            s_pts[j] = make_float4(xr, yr, depth_weight_m(zr, zmin, range, (P.flags & 64u) != 0), 0.f);   // 64u: SCREAM_F_HALF
        }
        __syncthreads();
        const int ibase = (int)(c0 - lo);
#pragma unroll 2
        for (int j = 0; j < cnt; ++j) {
            const float4 q = s_pts[j];
            const float dy = __fsub_rn(q.y, cy), dy2 = __fmul_rn(dy, dy);
#pragma unroll
            for (int u = 0; u < FD_PX; ++u) {
                const float dx = __fsub_rn(q.x, cx[u]);
                const float d2 = __fadd_rn(__fmul_rn(dx, dx), dy2);
                if (d2 <= t2[u]) {
                    const float val = pair_value(d2, q.z, P.negk);
                    if (val > best[u]) {
                        best[u] = val; arg[u] = ibase + j;
                        t2[u] = zero_or_cull(val, P.inv_negk);
                    }
                }
            }
        }
        __syncthreads();
    }

    const size_t o = (((size_t)b * P.V + v) * 2 + c) * npix + p0;
    const bool raw = P.flags & 1u;
    float out[FD_PX];
#pragma unroll
    for (int u = 0; u < FD_PX; ++u) {
        float r = (hi > lo) ? best[u] : 0.0f;
        if (zmin != zmin || zmax != zmax) { r = __int_as_float(0x7FC00000); arg[u] = 0; }   // NaN depth: see forward_culled.cuh
        out[u] = raw ? r : __fmul_rn(__fsub_rn(r, 0.5f), 2.0f);     // (img - 0.5) / 0.5, render.py:69
    }
    *reinterpret_cast<float4*>(P.imgs + o) = make_float4(out[0], out[1], out[2], out[3]);
    if (P.winner) {
        const int none = (hi > lo) ? 0 : -1;
        *reinterpret_cast<int4*>(P.winner + o) =
            make_int4(arg[0] + none, arg[1] + none, arg[2] + none, arg[3] + none);
    }
    if (P.zrange && tile == 0 && c == 0 && threadIdx.x == 0) {
        P.zrange[((size_t)b * P.V + v) * 2] = zmin;
        P.zrange[((size_t)b * P.V + v) * 2 + 1] = zmax;
    }
}

}  // namespace scream
