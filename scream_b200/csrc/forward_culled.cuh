// forward_culled.cuh -- culled exact forward: the product kernel.
//
// Same result, bit for bit, as forward_dense.cuh (and so as render.py:39-49): for every pixel the max
// over the cloud of pv_i * exp(-d2/2*rho^2) with the lowest index on ties -- but a (point, pixel) pair
// is only evaluated when it can still win.  One CTA renders one (item, view, cloud) image:
//
//   P1  transform (pose, view rotation, pv) every point of the cloud, bin it into a G x G grid of
//       cells over [-1,1]^2 (clamped), shared-memory histogram; the atomic's return value is the
//       point's rank inside its cell.
//   P2  exclusive scan of the histogram -> cell start offsets (row-major: a run of cells in one grid
//       row is one contiguous run of records).
//   P3  counting-sort scatter of (x', y') / pv / index records into shared memory (SoA).
//   P4  gather: a thread owns 4 horizontally adjacent pixels and searches the cells around them in
//       growing square windows.  A window only has to reach as far as cull_d2(best) (common.cuh): a
//       point farther than that from the pixel has a value strictly below the pixel's current best.
//       Inside a window every record costs one 8-byte shared load and a squared-distance test per pixel;
//       exp() is only evaluated for the few survivors.
//   Clouds larger than the shared-memory capacity are rendered in batches; the per-pixel best is carried
//   through the output buffers (max is associative; earlier batches hold the lower indices).
#pragma once
#include "common.cuh"
#include "forward_dense.cuh"   // FwdParams, zero_or_cull

namespace scream {

constexpr int FC_THREADS = 512;
constexpr int FC_PPT = 24;                               // points per thread per batch
constexpr int FC_CAP = FC_THREADS * FC_PPT;              // 12288 points per batch
constexpr int FC_G = 64;                                 // cell grid side
constexpr int FC_CELLS = FC_G * FC_G;
constexpr size_t FC_SMEM = (size_t)(FC_CELLS + 1) * 4 + (size_t)FC_CAP * (8 + 4 + 4) + 64;

struct Rect { int j0, j1, i0, i1; };

__device__ __forceinline__ int cell_of(float x) {
    // fl(fl(x + 1) * G/2), clamped; NaN -> 0
    float u = __fmul_rn(__fadd_rn(x, 1.0f), (float)(FC_G / 2));
    u = fminf(fmaxf(u, 0.0f), (float)(FC_G - 1));
    return (int)u;                                       // u >= 0: truncation == floor
}

// cells that can hold a point within `rc` cells of the pixel-centre span [ul, ur] x {uv}
__device__ __forceinline__ Rect window(float ul, float ur, float uv, float rc) {
    Rect W;
    W.j0 = max(0, (int)floorf(ul - rc)); W.j1 = min(FC_G - 1, (int)floorf(ur + rc));
    W.i0 = max(0, (int)floorf(uv - rc)); W.i1 = min(FC_G - 1, (int)floorf(uv + rc));
    return W;
}

// radius in cells (inflated: covers the <= 4e-6-cell rounding of cell_of and of the sqrt)
__device__ __forceinline__ float radius_cells(float t2) {
    float r = __fmul_rn(sqrtf(t2), (float)(FC_G / 2));
    return fminf(__fmaf_rn(r, 1e-4f, r) + 1e-3f, (float)(2 * FC_G));
}

__global__ void __launch_bounds__(FC_THREADS, 1)
forward_culled_kernel(FwdParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned* s_start = reinterpret_cast<unsigned*>(smem_raw);                    // [CELLS + 1]
    float2* s_xy = reinterpret_cast<float2*>(smem_raw + ((FC_CELLS + 1) * 4 + 15) / 16 * 16);
    float* s_pv = reinterpret_cast<float*>(s_xy + FC_CAP);
    int* s_idx = reinterpret_cast<int*>(s_pv + FC_CAP);
    __shared__ float s_R[9], s_pose[12];
    __shared__ unsigned s_warp_tot[FC_THREADS / 32];

    const int tid = threadIdx.x;
    const int w = P.w, npix = w * w, half = w / 2;
    int bid = blockIdx.x;
    const int c = bid % 2; bid /= 2;
    const int v = bid % P.V;
    const int b = bid / P.V;

    if (tid < 9) s_R[tid] = P.views[(size_t)v * 9 + tid];
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];

    const ItemSpan sp = item_span(P.offs, b);
    const int64_t lo = c ? sp.tgt0 : sp.src0, hi = c ? sp.end : sp.tgt0;
    const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]);
    const float zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
    const float range = z_range(zmin, zmax);
    const size_t img_off = (((size_t)b * P.V + v) * 2 + c) * npix;
    const bool raw = P.flags & 1u;
    const float cells_per_ndc = (float)(FC_G / 2);

    if (P.zrange && c == 0 && tid == 0) {
        P.zrange[((size_t)b * P.V + v) * 2] = zmin;
        P.zrange[((size_t)b * P.V + v) * 2 + 1] = zmax;
    }
    __syncthreads();

    int batch = 0;
    int64_t c0 = lo;
    do {                                                 // at least one pass, so empty clouds still write
        const int cnt = (int)max((int64_t)0, min((int64_t)FC_CAP, hi - c0));
        const bool last = c0 + cnt >= hi;

        // ---------------- P1: transform + histogram -------------------------------------------
        for (int t = tid; t <= FC_CELLS; t += FC_THREADS) s_start[t] = 0u;
        __syncthreads();
        unsigned packed[FC_PPT];
#pragma unroll
        for (int k = 0; k < FC_PPT; ++k) {
            const int j = k * FC_THREADS + tid;
            packed[k] = 0xFFFFFFFFu;
            if (j < cnt) {
                const float* q = P.xyz + (size_t)(c0 + j) * 3;
                float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
                if (posed) apply_pose(s_pose, x, y, z);
                const float xr = rot_row(s_R, x, y, z), yr = rot_row(s_R + 3, x, y, z);
                const int cell = cell_of(yr) * FC_G + cell_of(xr);
                const unsigned rank = atomicAdd(&s_start[cell], 1u);
                packed[k] = ((unsigned)cell << 16) | (rank & 0xFFFFu);       // cnt <= 12288 < 65536
            }
        }
        __syncthreads();

        // ---------------- P2: exclusive scan over the cells -----------------------------------
        {
            constexpr int PER = FC_CELLS / FC_THREADS;                       // 8 consecutive cells per thread
            unsigned loc[PER], sum = 0;
#pragma unroll
            for (int k = 0; k < PER; ++k) { loc[k] = s_start[tid * PER + k]; sum += loc[k]; }
            unsigned inc = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
                if ((tid & 31) >= d) inc += o;
            }
            if ((tid & 31) == 31) s_warp_tot[tid >> 5] = inc;
            __syncthreads();
            if (tid < 32) {
                unsigned t = tid < FC_THREADS / 32 ? s_warp_tot[tid] : 0u, ti = t;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    unsigned o = __shfl_up_sync(0xFFFFFFFFu, ti, d);
                    if (tid >= d) ti += o;
                }
                if (tid < FC_THREADS / 32) s_warp_tot[tid] = ti - t;         // exclusive warp offsets
            }
            __syncthreads();
            unsigned run = s_warp_tot[tid >> 5] + inc - sum;
#pragma unroll
            for (int k = 0; k < PER; ++k) { s_start[tid * PER + k] = run; run += loc[k]; }
            if (tid == FC_THREADS - 1) s_start[FC_CELLS] = run;
        }
        __syncthreads();

        // ---------------- P3: scatter records into cell order ---------------------------------
#pragma unroll
        for (int k = 0; k < FC_PPT; ++k) {
            if (packed[k] != 0xFFFFFFFFu) {
                const int j = k * FC_THREADS + tid;
                const float* q = P.xyz + (size_t)(c0 + j) * 3;
                float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
                if (posed) apply_pose(s_pose, x, y, z);
                const float xr = rot_row(s_R, x, y, z), yr = rot_row(s_R + 3, x, y, z), zr = rot_row(s_R + 6, x, y, z);
                const unsigned pos = s_start[packed[k] >> 16] + (packed[k] & 0xFFFFu);
                s_xy[pos] = make_float2(xr, yr);
                s_pv[pos] = depth_weight(zr, zmin, range);
                s_idx[pos] = (int)(c0 - lo) + j;
            }
        }
        __syncthreads();

        // ---------------- P4: per-pixel window search -----------------------------------------
        const int groups_per_row = w / FD_PX;
        for (int g = tid; g < npix / FD_PX; g += FC_THREADS) {
            const int row = g / groups_per_row, col0 = (g % groups_per_row) * FD_PX;
            const float cy = pixel_centre(row, half);
            float cx[FD_PX], best[FD_PX], t2[FD_PX];
            int arg[FD_PX];
            const size_t o = img_off + (size_t)row * w + col0;
            if (batch == 0) {
#pragma unroll
                for (int u = 0; u < FD_PX; ++u) { best[u] = 0.0f; arg[u] = 0; }
            } else {
                const float4 pb = *reinterpret_cast<const float4*>(P.imgs + o);
                best[0] = pb.x; best[1] = pb.y; best[2] = pb.z; best[3] = pb.w;
                if (P.winner) {
                    const int4 pa = *reinterpret_cast<const int4*>(P.winner + o);
                    arg[0] = pa.x; arg[1] = pa.y; arg[2] = pa.z; arg[3] = pa.w;
                } else {
                    arg[0] = arg[1] = arg[2] = arg[3] = -1;
                }
            }
            float t2max = 0.0f;
#pragma unroll
            for (int u = 0; u < FD_PX; ++u) {
                cx[u] = pixel_centre(col0 + u, half);
                t2[u] = zero_or_cull(best[u], P.inv_negk);
                t2max = fmaxf(t2max, t2[u]);
            }
            const float ul = __fmul_rn(__fadd_rn(cx[0], 1.0f), cells_per_ndc);
            const float ur = __fmul_rn(__fadd_rn(cx[FD_PX - 1], 1.0f), cells_per_ndc);
            const float uv = __fmul_rn(__fadd_rn(cy, 1.0f), cells_per_ndc);

            Rect done; done.j0 = 1; done.j1 = 0; done.i0 = 1; done.i1 = 0;   // empty
            float r = 1.0f;
            while (cnt > 0) {
                const float need = radius_cells(t2max);
                const float rr = fminf(r, need);
                const Rect W = window(ul, ur, uv, rr);
                for (int i = W.i0; i <= W.i1; ++i) {
                    const bool inside = i >= done.i0 && i <= done.i1;
                    // up to two runs of cells in this grid row: left of / right of the finished rectangle
                    for (int part = 0; part < 2; ++part) {
                        int ja, jb;
                        if (!inside) { if (part) break; ja = W.j0; jb = W.j1; }
                        else if (part == 0) { ja = W.j0; jb = done.j0 - 1; }
                        else { ja = done.j1 + 1; jb = W.j1; }
                        if (ja > jb) continue;
                        const int s0 = (int)s_start[i * FC_G + ja], s1 = (int)s_start[i * FC_G + jb + 1];
                        for (int s = s0; s < s1; ++s) {
                            const float2 q = s_xy[s];
                            const float dy = __fsub_rn(q.y, cy), dy2 = __fmul_rn(dy, dy);
#pragma unroll
                            for (int u = 0; u < FD_PX; ++u) {
                                const float dx = __fsub_rn(q.x, cx[u]);
                                const float d2 = __fadd_rn(__fmul_rn(dx, dx), dy2);
                                if (d2 <= t2[u]) {
                                    const float val = pair_value(d2, s_pv[s], P.negk);
                                    const int id = s_idx[s];
                                    if (val > best[u] || (val == best[u] && id < arg[u])) {
                                        best[u] = val; arg[u] = id;
                                        t2[u] = zero_or_cull(val, P.inv_negk);
                                        t2max = fmaxf(fmaxf(t2[0], t2[1]), fmaxf(t2[2], t2[3]));
                                    }
                                }
                            }
                        }
                    }
                }
                done = W;
                if (rr >= radius_cells(t2max)) break;
                r = __fmul_rn(r, 2.0f);
            }

            float out[FD_PX];
#pragma unroll
            for (int u = 0; u < FD_PX; ++u)
                out[u] = (raw || !last) ? best[u] : __fmul_rn(__fsub_rn(best[u], 0.5f), 2.0f);
            *reinterpret_cast<float4*>(P.imgs + o) = make_float4(out[0], out[1], out[2], out[3]);
            if (P.winner) {
                const int none = (hi > lo) ? 0 : -1;
                *reinterpret_cast<int4*>(P.winner + o) = make_int4(arg[0] + none, arg[1] + none, arg[2] + none, arg[3] + none);
            }
        }
        __syncthreads();
        c0 += cnt; ++batch;
    } while (c0 < hi);
}

}  // namespace scream
