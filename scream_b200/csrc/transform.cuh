// transform.cuh -- stage 1 of the path: rigid pose + view rotation + depth range.
//   render.py:63      cat(src, tgt)            -> CSR offsets, no copy
//   train_3d_match.py:198-199  R_gt src + t    -> fused (source cloud only)
//   render.py:68      R_v @ P.T                -> fmaf chain per view
//   render.py:30-31   min / max of rotated z over all n+m points, as host scalars
//                     -> device reduction (redux.sync on order-preserving uints + atomics), no host sync
#pragma once
#include "common.cuh"

namespace scream {

constexpr int ZR_THREADS = 256;
constexpr int ZR_PTS_PER_THREAD = 4;
constexpr int ZR_PTS_PER_CTA = ZR_THREADS * ZR_PTS_PER_THREAD;
constexpr int MAX_VIEWS_SMEM = 64;   // views handled per pass of the z-range kernel

// zenc: [B][V][2] uint, pre-set to (0xFFFFFFFF, 0) by zrange_init_kernel.
__global__ void zrange_init_kernel(unsigned* __restrict__ zenc, int count_pairs) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count_pairs) { zenc[2 * i] = 0xFFFFFFFFu; zenc[2 * i + 1] = 0u; }
}

// grid: (chunks_max, B).  Each thread keeps up to 4 (posed) points in registers and walks the views.
__global__ void __launch_bounds__(ZR_THREADS)
zrange_kernel(const float* __restrict__ xyz, const int64_t* __restrict__ offs,
              const float* __restrict__ pose, const float* __restrict__ views,
              int V, unsigned* __restrict__ zenc)
{
    const int b = blockIdx.y;
    const ItemSpan sp = item_span(offs, b);
    const int64_t npts = sp.end - sp.src0;
    const int64_t base = (int64_t)blockIdx.x * ZR_PTS_PER_CTA;
    if (base >= npts) return;

    __shared__ float s_views[MAX_VIEWS_SMEM * 3];   // z rows only
    __shared__ unsigned s_min[MAX_VIEWS_SMEM], s_max[MAX_VIEWS_SMEM];
    __shared__ float s_pose[12];
    if (pose && threadIdx.x < 12) s_pose[threadIdx.x] = pose[(size_t)b * 12 + threadIdx.x];

    float px[ZR_PTS_PER_THREAD], py[ZR_PTS_PER_THREAD], pz[ZR_PTS_PER_THREAD];
    bool ok[ZR_PTS_PER_THREAD];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ZR_PTS_PER_THREAD; ++k) {
        int64_t i = base + (int64_t)k * ZR_THREADS + threadIdx.x;
        ok[k] = i < npts;
        px[k] = py[k] = pz[k] = 0.f;
        if (ok[k]) {
            const float* p = xyz + (size_t)(sp.src0 + i) * 3;
            px[k] = __ldg(p); py[k] = __ldg(p + 1); pz[k] = __ldg(p + 2);
            if (pose && sp.src0 + i < sp.tgt0) apply_pose(s_pose, px[k], py[k], pz[k]);
        }
    }
    for (int v0 = 0; v0 < V; v0 += MAX_VIEWS_SMEM) {
        const int nv = min(MAX_VIEWS_SMEM, V - v0);
        __syncthreads();
        for (int t = threadIdx.x; t < nv * 3; t += ZR_THREADS)
            s_views[t] = views[(size_t)(v0 + t / 3) * 9 + 6 + t % 3];
        if (threadIdx.x < nv) { s_min[threadIdx.x] = 0xFFFFFFFFu; s_max[threadIdx.x] = 0u; }
        __syncthreads();
        for (int v = 0; v < nv; ++v) {
            unsigned lo = 0xFFFFFFFFu, hi = 0u;
#pragma unroll
            for (int k = 0; k < ZR_PTS_PER_THREAD; ++k) {
                if (ok[k]) {
                    unsigned e = enc_f32(rot_row(s_views + 3 * v, px[k], py[k], pz[k]));
                    lo = min(lo, e); hi = max(hi, e);
                }
            }
            lo = __reduce_min_sync(0xFFFFFFFFu, lo);
            hi = __reduce_max_sync(0xFFFFFFFFu, hi);
            if ((threadIdx.x & 31) == 0) { atomicMin(&s_min[v], lo); atomicMax(&s_max[v], hi); }
        }
        __syncthreads();
        if (threadIdx.x < nv) {
            unsigned* z = zenc + ((size_t)b * V + v0 + threadIdx.x) * 2;
            atomicMin(z, s_min[threadIdx.x]);
            atomicMax(z + 1, s_max[threadIdx.x]);
        }
    }
}

// caller-supplied zrange floats [B][V][2] -> zenc (SCREAM_F_ZRANGE_IN)
__global__ void zrange_encode_kernel(const float* __restrict__ zrange, unsigned* __restrict__ zenc, int count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) zenc[i] = enc_f32(zrange[i]);
}

// decode zenc -> zrange floats [B][V][2]
__global__ void zrange_decode_kernel(const unsigned* __restrict__ zenc, float* __restrict__ zrange, int count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) zrange[i] = dec_f32(zenc[i]);
}

// Debug / parity kernel: writes rotated [V][total][3] and pv [V][total].  grid (chunks, B).
__global__ void transform_dump_kernel(const float* __restrict__ xyz, const int64_t* __restrict__ offs,
                                      int64_t total_pts, const float* __restrict__ pose,
                                      const float* __restrict__ views, int V,
                                      const unsigned* __restrict__ zenc,
                                      float* __restrict__ rotated, float* __restrict__ pv)
{
    const int b = blockIdx.y;
    const ItemSpan sp = item_span(offs, b);
    const int64_t i = sp.src0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sp.end) return;
    float x = xyz[i * 3], y = xyz[i * 3 + 1], z = xyz[i * 3 + 2];
    if (pose && i < sp.tgt0) apply_pose(pose + (size_t)b * 12, x, y, z);
    for (int v = 0; v < V; ++v) {
        const float* R = views + (size_t)v * 9;
        float xr = rot_row(R, x, y, z), yr = rot_row(R + 3, x, y, z), zr = rot_row(R + 6, x, y, z);
        if (rotated) {
            float* o = rotated + ((size_t)v * total_pts + i) * 3;
            o[0] = xr; o[1] = yr; o[2] = zr;
        }
        if (pv) {
            float zmin = dec_f32(zenc[((size_t)b * V + v) * 2]), zmax = dec_f32(zenc[((size_t)b * V + v) * 2 + 1]);
            pv[(size_t)v * total_pts + i] = depth_weight(zr, zmin, z_range(zmin, zmax));
        }
    }
}

}  // namespace scream
