// transform.cuh -- stage 1 of the path: rigid pose + view rotation + depth range.
//   render.py:63      cat(src, tgt)            -> CSR offsets, no copy
//   train_3d_match.py:198-199  R_gt src + t    -> fused (source cloud only)
//   render.py:68      R_v @ P.T                -> fmaf chain per view
//   render.py:30-31   min / max of rotated z over all n+m points, as host scalars
//                     -> device reduction (redux.sync on order-preserving uints + atomics), no host sync
#pragma once
#include "common.cuh"
#include <cuda_pipeline.h>

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

// ---- large clouds (>= ZW_MIN_PTS points in an item): views in registers, points streamed ---------------------
// grid (chunks, view groups, B).  A thread keeps the z rows and the running min / max of ZW_VPG views in
// registers and walks ZW_PTS_PER_THREAD points two at a time: per (point, view) 3 FMA-pipe instructions and one
// three-input FMNMX3 (sm_100, .NaN: a NaN depth poisons the range like torch.min / torch.max do), against ~10 in the kernel above.  The cloud is read V / ZW_VPG times (from L2
// when it fits).  Float min / max order -0 < +0 like the integer keys do, so both kernels give the same bits.
constexpr int ZW_THREADS = 256;
#ifndef ZW_VPG_N
#define ZW_VPG_N 16
#endif
constexpr int ZW_VPG = ZW_VPG_N;
constexpr int ZW_PTS_PER_THREAD = 64;
constexpr int ZW_STAGES = 3;                // cp.async stages per thread (4 points each)
constexpr int ZW_PTS_PER_CTA = ZW_THREADS * ZW_PTS_PER_THREAD;
constexpr int64_t ZW_MIN_PTS = 1 << 18;

__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float r; asm("min.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float r; asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;
}

// packed fp32 x 2 (FMUL2 / FFMA2, sm_100): both halves rounded like the scalar instructions
#ifndef ZW_PACKED
#define ZW_PACKED 1
#endif
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t zw_pack(float lo, float hi) { f32x2_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void zw_unpack(f32x2_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2_t zw_mul2(f32x2_t a, f32x2_t b) { f32x2_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2_t zw_fma2(f32x2_t a, f32x2_t b, f32x2_t c) { f32x2_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

__global__ void __launch_bounds__(ZW_THREADS)
zrange_wide_kernel(const float* __restrict__ xyz, const int64_t* __restrict__ offs,
                   const float* __restrict__ pose, const float* __restrict__ views,
                   int V, unsigned* __restrict__ zenc)
{
    const int b = blockIdx.z, v0 = blockIdx.y * ZW_VPG;
    const ItemSpan sp = item_span(offs, b);
    const int64_t npts = sp.end - sp.src0;
    const int64_t base = (int64_t)blockIdx.x * ZW_PTS_PER_CTA;
    if (base >= npts) return;

    __shared__ unsigned s_min[ZW_VPG], s_max[ZW_VPG];
    __shared__ float s_pose[12];
    if (pose && threadIdx.x < 12) s_pose[threadIdx.x] = pose[(size_t)b * 12 + threadIdx.x];
    if (threadIdx.x < ZW_VPG) { s_min[threadIdx.x] = 0xFFFFFFFFu; s_max[threadIdx.x] = 0u; }

    float r[ZW_VPG][3], lo[ZW_VPG], hi[ZW_VPG];
#pragma unroll
    for (int v = 0; v < ZW_VPG; ++v) {
        const float* R = views + (size_t)min(v0 + v, V - 1) * 9 + 6;     // past the end: the last view again
        r[v][0] = __ldg(R); r[v][1] = __ldg(R + 1); r[v][2] = __ldg(R + 2);
        lo[v] = INFINITY; hi[v] = -INFINITY;
    }
#if ZW_PACKED
    // the z rows of two views side by side: one FMUL2 + two FFMA2 give a point's depth in both (the point's coordinate
    // is the same in both halves), half the FMA-pipe instructions of the scalar form and the same bits
    static_assert(ZW_VPG % 2 == 0, "views are paired");
    f32x2_t rp[ZW_VPG / 2][3];
#pragma unroll
    for (int v = 0; v < ZW_VPG / 2; ++v)
#pragma unroll
        for (int k = 0; k < 3; ++k) rp[v][k] = zw_pack(r[2 * v][k], r[2 * v + 1][k]);
#endif
    __syncthreads();

    // The points come through shared memory with cp.async, ZW_STAGES - 1 iterations ahead (the loop is otherwise bound
    // by the load latency), four per thread and iteration.  A thread only reads back what it copied itself: no block
    // barrier, wait_prior is enough.  Indices are local to the CTA's chunk (32-bit).
    __shared__ float s_stage[ZW_STAGES][12][ZW_THREADS];
    const float* pbase = xyz + (size_t)(sp.src0 + base) * 3;
    const int nloc = (int)min((int64_t)ZW_PTS_PER_CTA, npts - base);                 // points of this chunk
    const int nsrc = (int)max((int64_t)0, min((int64_t)nloc, sp.tgt0 - sp.src0 - base)); // ... of which source points (posed)
    const int tid = threadIdx.x;
    auto issue = [&](int i0, int st) {
        if (i0 < nloc) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * ZW_THREADS < nloc ? i0 + u * ZW_THREADS : i0;     // tail: the first point again
                const float* p = pbase + i * 3;
#pragma unroll
                for (int k = 0; k < 3; ++k) __pipeline_memcpy_async(&s_stage[st][3 * u + k][tid], p + k, 4);
            }
        }
        __pipeline_commit();
    };
#pragma unroll
    for (int st = 0; st < ZW_STAGES - 1; ++st) issue(tid + st * 4 * ZW_THREADS, st);
    int st = 0;
#pragma unroll 1
    for (int i0 = tid; i0 < nloc; i0 += 4 * ZW_THREADS) {
        issue(i0 + (ZW_STAGES - 1) * 4 * ZW_THREADS, st == 0 ? ZW_STAGES - 1 : st - 1);
        __pipeline_wait_prior(ZW_STAGES - 1);
        float px[4], py[4], pz[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { px[u] = s_stage[st][3 * u][tid]; py[u] = s_stage[st][3 * u + 1][tid]; pz[u] = s_stage[st][3 * u + 2][tid]; }
        st = st == ZW_STAGES - 1 ? 0 : st + 1;
        if (pose && i0 < nsrc) {                          // warp-uniform up to the one warp that straddles the clouds
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * ZW_THREADS < nloc ? i0 + u * ZW_THREADS : i0;
                if (i < nsrc) apply_pose(s_pose, px[u], py[u], pz[u]);
            }
        }
#if ZW_PACKED
        f32x2_t X[4], Y[4], Z[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { X[u] = zw_pack(px[u], px[u]); Y[u] = zw_pack(py[u], py[u]); Z[u] = zw_pack(pz[u], pz[u]); }
#pragma unroll
        for (int v = 0; v < ZW_VPG / 2; ++v) {
            float z0[4], z1[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)                       // rot_row(): fmaf(r2, z, fmaf(r1, y, r0 * x)) in both halves
                zw_unpack(zw_fma2(rp[v][2], Z[u], zw_fma2(rp[v][1], Y[u], zw_mul2(rp[v][0], X[u]))), z0[u], z1[u]);
            lo[2 * v] = fmin3(fmin3(lo[2 * v], z0[0], z0[1]), z0[2], z0[3]);
            hi[2 * v] = fmax3(fmax3(hi[2 * v], z0[0], z0[1]), z0[2], z0[3]);
            lo[2 * v + 1] = fmin3(fmin3(lo[2 * v + 1], z1[0], z1[1]), z1[2], z1[3]);
            hi[2 * v + 1] = fmax3(fmax3(hi[2 * v + 1], z1[0], z1[1]), z1[2], z1[3]);
        }
#else
#pragma unroll
        for (int v = 0; v < ZW_VPG; ++v) {
            const float za = rot_row(r[v], px[0], py[0], pz[0]), zb = rot_row(r[v], px[1], py[1], pz[1]);
            const float zc = rot_row(r[v], px[2], py[2], pz[2]), zd = rot_row(r[v], px[3], py[3], pz[3]);
            lo[v] = fmin3(fmin3(lo[v], za, zb), zc, zd);
            hi[v] = fmax3(fmax3(hi[v], za, zb), zc, zd);
        }
#endif
    }
#pragma unroll
    for (int v = 0; v < ZW_VPG; ++v) {
        const unsigned l = __reduce_min_sync(0xFFFFFFFFu, enc_f32(lo[v]));
        const unsigned h = __reduce_max_sync(0xFFFFFFFFu, enc_f32(hi[v]));
        if ((threadIdx.x & 31) == 0) { atomicMin(&s_min[v], l); atomicMax(&s_max[v], h); }
    }
    __syncthreads();
    if (threadIdx.x < ZW_VPG && v0 + threadIdx.x < V) {
        unsigned* z = zenc + ((size_t)b * V + v0 + threadIdx.x) * 2;
        atomicMin(z, s_min[threadIdx.x]);
        atomicMax(z + 1, s_max[threadIdx.x]);
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
