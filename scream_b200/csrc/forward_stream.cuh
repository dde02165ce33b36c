// forward_stream.cuh -- giant clouds (BASELINE config 5: 10^6 .. 10^8 points per pair): an exact
// pre-filter that streams the points and keeps only those that can win a pixel, so that the culled
// kernel (forward_culled.cuh) renders a few thousand survivors per image instead of the whole cloud.
// This is north_star's "projection + z-buffer" stage: a per-pixel atomicMax on float bits, staged
// through a shared-memory image tile.
//
//   S1 seed    every point -> its home cell (the cell of the 64 x 64 grid that holds it; for w == 64 a cell
//              is a pixel) -> value at that pixel, rounded DOWN -> atomicMax on the uint bits of the value
//              (monotone for values >= 0) in a shared-memory tile, flushed to the global seed image.
//              seed[q] is the value of a real point at pixel q, hence a lower bound of the pixel's max.
//   S2 bound   per image: eff[q]  = max(seed[q], max_q' seed[q'] * exp(-k dmax^2(q',q)))   (lower bound of
//              the max at q, also for pixels no point calls home), and the per-cell admission threshold
//              thr[a] = min_{q != a} eff[q] * exp(+k dmin^2(a,q)):  a point of cell a can only reach
//              eff[q] at another pixel q if its pv is at least thr[a]  (val <= pv * exp(-k dmin^2)).
//   S3 filter  every point again: it survives iff its value at its home pixel (rounded UP) reaches
//              eff[home], or its pv reaches thr[home cell].  Survivors (cloud-local indices) are appended
//              to a per-image list.  Every point that attains a pixel's maximum survives -- including all
//              points that tie -- so rendering the survivors is bit-identical to rendering the cloud.
//   S4 render  forward_culled_kernel on the survivor lists (FwdParams::surv*); an image whose list
//              overflowed is rendered from the full cloud in batches (correct, slow).
// dmax / dmin: largest / smallest distance between a point anywhere in cell q' (resp. a) and the pixel
// centre of cell q: per axis |delta| + 1/2 resp. max(0, |delta| - 1/2) cells.
#pragma once
#include "common.cuh"
#include "forward_dense.cuh"
#include "forward_culled.cuh"

namespace scream {

constexpr int FS_THREADS = 1024;                        // S1: one CTA per SM (128 KB of tiles), 32 warps
constexpr int FS_FILTER_THREADS = 1024;
constexpr int FS_FILTER_CHUNK = 65536;                   // points per CTA task of S3 (amortises staging the bounds)
constexpr int FS_CHUNK = 32768;                          // points per CTA task
constexpr int FS_SURV_MIN = 4 * FC_CAP;                  // survivors kept per image: at least 49152, else total_pts / 16
constexpr int64_t FS_MIN_AVG_PTS = 4 * FC_CAP;           // stream path when total_pts / B reaches this
constexpr float FS_DOWN = 1.0f - 1e-4f;                  // covers ex2.approx + argument rounding at |arg| <= 108
constexpr float FS_UP = 1.0f + 1e-4f;

struct StreamParams {
    const float* xyz; const int64_t* offs; const float* pose; const float* views; const unsigned* zenc;
    int B, V;
    float negk;                                          // -rho^2/2
    unsigned* seed;                                      // [B][V][2][4096] uint bits of float >= 0 (S1), then eff (S2; -1 = unknown)
    float2* bnd;                                         // [B][V][2][4096] (eff, thr) written by S2, read by S3
    int* surv; unsigned* surv_cnt; int surv_cap;         // [B][V][2][surv_cap], [B][V][2]
};

constexpr int FS_VPC = 4;                                // views per CTA task of S1 (a point is loaded once for all of them)
constexpr int FS_VPC_FILTER = 2;                         // ... of S3: 2 views x 2 clouds x 32 KB of bounds staged in shared memory
constexpr float FS_PV_ABS = 4e-7f;                       // |1 - (z-zmin)*rcp(range) - pv_exact| (two roundings + rcp) stays below this

// fast exp for bounds only: ex2.approx(a * log2e), error <= 2 + 1.16|a| ulp (|a| <= 108: < 1.6e-5 relative)
__device__ __forceinline__ float exp_fast(float a) { return __expf(a); }

// per-view constants of a CTA task
struct StreamView { float R[9]; float zmin, rcp_range; };

struct StreamPoint { float pv, d2c; int cell; bool inside; };

// transform for BOUNDS: the renderer's rotation, pv through a reciprocal multiply (|error| <= FS_PV_ABS, applied in
// the safe direction by the callers), home cell = floor of the cell coordinate (clamped like the renderer's grid) and
// d2c = squared distance, in cells^2, to the pixel centre of that cell (w == 64: pixel = cell).  The cell function
// need not be bit-identical to the renderer's: S1 and S3 share it, and the bounds of S2 carry 1e-4 of slack.
__device__ __forceinline__ StreamPoint stream_point(float x, float y, float z, const StreamView& W) {
    StreamPoint sp;
    const float ux = __fmaf_rn(rot_row(W.R, x, y, z), FC_CELLS_PER_NDC, FC_CELLS_PER_NDC);
    const float uy = __fmaf_rn(rot_row(W.R + 3, x, y, z), FC_CELLS_PER_NDC, FC_CELLS_PER_NDC);
    sp.pv = __fmaf_rn(-(rot_row(W.R + 6, x, y, z) - W.zmin), W.rcp_range, 1.0f);
    const float jf = fminf(fmaxf(floorf(ux), 0.0f), (float)(FC_G - 1)), jf2 = fminf(fmaxf(floorf(uy), 0.0f), (float)(FC_G - 1));
    sp.inside = ux >= 0.0f && ux < (float)FC_G && uy >= 0.0f && uy < (float)FC_G;     // really inside its cell (not clamped)
    const float fx = ux - jf - 0.5f, fy = uy - jf2 - 0.5f;
    sp.d2c = __fmaf_rn(fx, fx, fy * fy);
    sp.cell = (int)jf2 * FC_G + (int)jf;                 // NaN coordinates: cell 0, inside == false
    return sp;
}

// exp(-k d^2) at the home pixel, fast exp; negkc = negk / (cells per NDC)^2
__device__ __forceinline__ float home_weight(const StreamPoint& sp, float negkc) { return exp_fast(sp.d2c * negkc); }

__device__ __forceinline__ void stream_load_views(const StreamParams& P, int b, int v0, int nv, StreamView* s_view) {
    for (int t = threadIdx.x; t < nv * 11; t += blockDim.x) {
        const int vv = t / 11, k = t % 11, v = v0 + vv;
        if (k < 9) s_view[vv].R[k] = P.views[(size_t)v * 9 + k];
        else {
            const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
            if (k == 9) s_view[vv].zmin = zmin; else s_view[vv].rcp_range = 1.0f / z_range(zmin, zmax);
        }
    }
}

// ---- S1: seed ----------------------------------------------------------------------------------
// grid (chunks, ceil(V / FS_VPC), B).
constexpr size_t FS_SEED_SMEM = (size_t)FS_VPC * 2 * FC_CELLS * sizeof(unsigned);

__global__ void __launch_bounds__(FS_THREADS)
stream_seed_kernel(StreamParams P)
{
    extern __shared__ unsigned s_tile[];                 // [FS_VPC][2][4096]
    __shared__ StreamView s_view[FS_VPC];
    __shared__ float s_pose[12];
    const int v0 = blockIdx.y * FS_VPC, nv = min(FS_VPC, P.V - v0), b = blockIdx.z;
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t base = sp.src0 + (int64_t)blockIdx.x * FS_CHUNK;
    if (base >= sp.end) return;
    const int tid = threadIdx.x;
    stream_load_views(P, b, v0, nv, s_view);
    if (P.pose && tid < 12) s_pose[tid] = P.pose[(size_t)b * 12 + tid];
    for (int t = tid; t < nv * 2 * FC_CELLS; t += FS_THREADS) s_tile[t] = 0u;
    __syncthreads();
    const float negkc = P.negk / (FC_CELLS_PER_NDC * FC_CELLS_PER_NDC);
    const int64_t stop = min(sp.end, base + FS_CHUNK);
    for (int64_t gi = base + tid; gi < stop; gi += FS_THREADS) {
        const int c = gi >= sp.tgt0;
        const float* q = P.xyz + (size_t)gi * 3;
        float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
        if (P.pose != nullptr && c == 0) apply_pose(s_pose, x, y, z);
        for (int vv = 0; vv < nv; ++vv) {
            const StreamPoint pt = stream_point(x, y, z, s_view[vv]);
            const float pv_lo = pt.pv - FS_PV_ABS;
            if (!(pv_lo > 0.0f) || !pt.inside) continue;     // clamped points may be anywhere outside: they cannot seed
            const float val = pv_lo * home_weight(pt, negkc) * FS_DOWN;
            const unsigned bits = __float_as_uint(val);
            unsigned* slot = &s_tile[(vv * 2 + c) * FC_CELLS + pt.cell];
            if (val > 0.0f && bits > *slot) atomicMax(slot, bits);
        }
    }
    __syncthreads();
    unsigned* out = P.seed + ((size_t)b * P.V + v0) * 2 * FC_CELLS;
    for (int t = tid; t < nv * 2 * FC_CELLS; t += FS_THREADS) {
        const unsigned bits = s_tile[t];
        if (bits != 0u && bits > out[t]) atomicMax(&out[t], bits);
    }
}

// ---- S2: per-image bounds ----------------------------------------------------------------------
// grid (V * 2, B), 1024 threads, 4 cells per thread.  Both bounds are separable in the two grid axes
// (exp(-k(dx^2+dy^2)) = exp(-k dx^2) exp(-k dy^2), and max / min distribute over the product of non-negative
// factors), so each is a column pass followed by a row pass over a window of `radius` cells: O(4096 * radius)
// per image instead of O(4096 * radius^2).  The own cell is excluded from thr[a] exactly: the column pass
// leaves out the own row, and the row pass adds the own row without the own column.
constexpr size_t FS_BOUND_SMEM(int radius) { return (size_t)4 * FC_CELLS * sizeof(float) + (size_t)3 * (radius + 1) * sizeof(float); }

__global__ void __launch_bounds__(1024)
stream_bound_kernel(StreamParams P, int radius)
{
    extern __shared__ float s_dyn[];
    float* s_seed = s_dyn;                               // seed, later E (eff as used by the threshold pass)
    float* s_a = s_dyn + FC_CELLS;                       // column pass of eff, later column pass of thr
    float* s_u = s_dyn + 2 * FC_CELLS;                   // column pass of the distance to the nearest seeded cell
    float* s_eff = s_dyn + 3 * FC_CELLS;
    float* t_max = s_dyn + 4 * FC_CELLS;                 // exp(-k ((d + 1/2) cell)^2)
    float* t_min = t_max + radius + 1;                   // exp(+k (max(d - 1/2, 0) cell)^2), capped (finite)
    float* t_d2 = t_min + radius + 1;                    // (max(d - 1/2, 0) cell)^2
    const int vc = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const size_t img = ((size_t)b * P.V * 2 + vc) * FC_CELLS;
    const float cell = 1.0f / FC_CELLS_PER_NDC;
    for (int d = tid; d <= radius; d += 1024) {
        const float hi = ((float)d + 0.5f) * cell, lo = fmaxf((float)d - 0.5f, 0.0f) * cell;
        t_max[d] = expf(hi * hi * P.negk) * (1.0f - 1e-4f);                     // accurate exp: denormals kept
        t_min[d] = fminf(expf(-lo * lo * P.negk) * (1.0f - 1e-4f), 3.0e38f);
        t_d2[d] = lo * lo;
    }
    for (int t = tid; t < FC_CELLS; t += 1024) s_seed[t] = __uint_as_float(P.seed[img + t]);
    __syncthreads();
    const float zero_d2 = FC_EXP_ZERO / (-P.negk) * 1.001f;                      // NDC^2 beyond which a value is exactly 0

    // eff, column pass
    for (int q = tid; q < FC_CELLS; q += 1024) {
        const int i = q >> 6, j = q & 63;
        float a = 0.0f, u = INFINITY;
        for (int i2 = max(0, i - radius); i2 <= min(FC_G - 1, i + radius); ++i2) {
            const float sd = s_seed[i2 * FC_G + j];
            if (sd > 0.0f) { const int d = abs(i2 - i); a = fmaxf(a, sd * t_max[d]); u = fminf(u, t_d2[d]); }
        }
        s_a[q] = a; s_u[q] = u;
    }
    __syncthreads();
    // eff, row pass.  -1: unknown (a seeded cell within the zero radius but the bound underflowed), 0: exactly zero
    for (int q = tid; q < FC_CELLS; q += 1024) {
        const int i = q >> 6, j = q & 63;
        float e = s_seed[q], m = INFINITY;
        for (int j2 = max(0, j - radius); j2 <= min(FC_G - 1, j + radius); ++j2) {
            const int d = abs(j2 - j);
            e = fmaxf(e, s_a[i * FC_G + j2] * t_max[d]);
            m = fminf(m, s_u[i * FC_G + j2] + t_d2[d]);
        }
        s_eff[q] = e > 0.0f ? e : (m <= zero_d2 ? -1.0f : 0.0f);
    }
    __syncthreads();
    // E = eff as the threshold pass uses it: unknown -> 0 (admits everything in reach), exactly zero -> +inf (skipped:
    // the winner of an all-zero pixel is index 0 whatever the points)
    for (int q = tid; q < FC_CELLS; q += 1024) {
        const float e = s_eff[q];
        s_seed[q] = e > 0.0f ? e : (e < 0.0f ? 0.0f : INFINITY);
    }
    __syncthreads();
    // thr, column pass without the own row
    for (int q = tid; q < FC_CELLS; q += 1024) {
        const int i = q >> 6, j = q & 63;
        float h = INFINITY;
        for (int i2 = max(0, i - radius); i2 <= min(FC_G - 1, i + radius); ++i2)
            if (i2 != i) h = fminf(h, s_seed[i2 * FC_G + j] * t_min[abs(i2 - i)]);
        s_a[q] = h;
    }
    __syncthreads();
    // thr, row pass: other rows through s_a (all columns), own row through E (other columns only)
    for (int q = tid; q < FC_CELLS; q += 1024) {
        const int i = q >> 6, j = q & 63;
        float thr = INFINITY;
        for (int j2 = max(0, j - radius); j2 <= min(FC_G - 1, j + radius); ++j2) {
            const float f = t_min[abs(j2 - j)];
            thr = fminf(thr, s_a[i * FC_G + j2] * f);
            if (j2 != j) thr = fminf(thr, s_seed[i * FC_G + j2] * f);
        }
        P.bnd[img + q] = make_float2(s_eff[q], thr * (1.0f - 1e-4f));     // eff: -1 unknown, 0 exactly zero
    }
}

// ---- S3: filter --------------------------------------------------------------------------------
// grid (chunks of FS_FILTER_CHUNK, ceil(V / FS_VPC_FILTER), B).  The (eff, thr) bounds of the CTA's views are staged
// in shared memory: the per-point lookup is a random 8-byte read, which L1 serves at one sector per cycle.
constexpr int FS_LIST_CAP = 5120;                        // survivors a CTA collects per image before it appends them in one piece
constexpr size_t FS_FILTER_SMEM = (size_t)FS_VPC_FILTER * 2 * (FC_CELLS * sizeof(float2) + FS_LIST_CAP * sizeof(int));

__global__ void __launch_bounds__(FS_FILTER_THREADS)
stream_filter_kernel(StreamParams P)
{
    extern __shared__ float2 s_bnd[];                    // [FS_VPC_FILTER][2][4096], then int s_list[FS_VPC_FILTER * 2][FS_LIST_CAP]
    int* s_list = reinterpret_cast<int*>(s_bnd + FS_VPC_FILTER * 2 * FC_CELLS);
    __shared__ unsigned s_cnt[FS_VPC_FILTER * 2], s_base[FS_VPC_FILTER * 2];
    __shared__ StreamView s_view[FS_VPC_FILTER];
    __shared__ float s_pose[12];
    const int v0 = blockIdx.y * FS_VPC_FILTER, nv = min(FS_VPC_FILTER, P.V - v0), b = blockIdx.z;
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t base = sp.src0 + (int64_t)blockIdx.x * FS_FILTER_CHUNK;
    if (base >= sp.end) return;
    const int tid = threadIdx.x, lane = tid & 31;
    stream_load_views(P, b, v0, nv, s_view);
    if (P.pose && tid < 12) s_pose[tid] = P.pose[(size_t)b * 12 + tid];
    if (tid < FS_VPC_FILTER * 2) s_cnt[tid] = 0u;
    {
        const float4* src = reinterpret_cast<const float4*>(P.bnd + ((size_t)b * P.V + v0) * 2 * FC_CELLS);
        float4* dst = reinterpret_cast<float4*>(s_bnd);
        for (int t = tid; t < nv * FC_CELLS; t += FS_FILTER_THREADS) dst[t] = __ldg(src + t);      // 2 clouds x 4096 float2 = 4096 float4 per view
    }
    __syncthreads();
    const float negkc = P.negk / (FC_CELLS_PER_NDC * FC_CELLS_PER_NDC);
    const int64_t stop = min(sp.end, base + FS_FILTER_CHUNK);
    const int64_t iters = (stop - base + FS_FILTER_THREADS - 1) / FS_FILTER_THREADS;
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t gi = base + it * FS_FILTER_THREADS + tid;
        const bool live = gi < stop;
        const int c = live && gi >= sp.tgt0;
        float x = 0.f, y = 0.f, z = 0.f;
        if (live) {
            const float* q = P.xyz + (size_t)gi * 3;
            x = __ldcs(q); y = __ldcs(q + 1); z = __ldcs(q + 2);
            if (P.pose != nullptr && c == 0) apply_pose(s_pose, x, y, z);
        }
        for (int vv = 0; vv < nv; ++vv) {
            const size_t img0 = ((size_t)b * P.V + v0 + vv) * 2;
            bool keep = false;
            if (live) {
                const StreamPoint pt = stream_point(x, y, z, s_view[vv]);
                const float pv_hi = pt.pv + FS_PV_ABS;
                if (pv_hi > 0.0f) {
                    const float2 bd = s_bnd[(vv * 2 + c) * FC_CELLS + pt.cell];                // (eff, thr)
                    // points outside [-1,1]^2 (clamped into border cells) never seed and are never dropped: S2's
                    // "exactly zero" pixels are only zero for the points inside
                    keep = !pt.inside || (pv_hi * home_weight(pt, negkc) * FS_UP >= bd.x) || (pv_hi >= bd.y);
                }
            }
            const unsigned any = __ballot_sync(0xFFFFFFFFu, keep);
            if (any == 0u) continue;
            // a warp straddles at most the src/tgt boundary: one ballot per cloud.  Survivors go to the CTA's
            // shared-memory list of the image (one shared atomic per warp); what does not fit goes straight to global.
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const unsigned m = __ballot_sync(0xFFFFFFFFu, keep && c == cc);
                if (m) {
                    const int l = vv * 2 + cc;
                    unsigned pos = 0;
                    if (lane == 0) pos = atomicAdd(&s_cnt[l], (unsigned)__popc(m));
                    pos = __shfl_sync(0xFFFFFFFFu, pos, 0) + __popc(m & ((1u << lane) - 1u));
                    if (keep && c == cc) {
                        const int li = (int)(gi - (cc ? sp.tgt0 : sp.src0));
                        if (pos < (unsigned)FS_LIST_CAP) s_list[l * FS_LIST_CAP + pos] = li;
                        else {
                            const unsigned gp = atomicAdd(&P.surv_cnt[img0 + cc], 1u);
                            if (gp < (unsigned)P.surv_cap) P.surv[(img0 + cc) * (size_t)P.surv_cap + gp] = li;
                        }
                    }
                }
            }
        }
    }
    __syncthreads();
    if (tid < nv * 2) {
        const unsigned n = min(s_cnt[tid], (unsigned)FS_LIST_CAP);
        s_cnt[tid] = n;
        s_base[tid] = n ? atomicAdd(&P.surv_cnt[((size_t)b * P.V + v0) * 2 + tid], n) : 0u;
    }
    __syncthreads();
    for (int l = 0; l < nv * 2; ++l) {
        const unsigned n = s_cnt[l], gb = s_base[l];
        int* dst = P.surv + (((size_t)b * P.V + v0) * 2 + l) * (size_t)P.surv_cap;
        for (unsigned t = tid; t < n; t += FS_FILTER_THREADS)
            if (gb + t < (unsigned)P.surv_cap) dst[gb + t] = s_list[l * FS_LIST_CAP + t];
    }
}

}  // namespace scream
