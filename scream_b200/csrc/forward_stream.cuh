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
constexpr int FS_FILTER_CHUNK_MAX = 131072;              // points per CTA task of S3 (amortises staging the thresholds)
constexpr int FS_FILTER_CHUNK_MIN = 8192;
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
    int seed_stride;                                     // S1 seeds from every seed_stride-th chunk of the cloud (any subset is a valid lower bound)
    int filter_chunk;                                    // points per CTA task of S3
};

constexpr int FS_VPC = 4;                                // views per CTA task of S1 (a point is loaded once for all of them)
constexpr int FS_VPC_FILTER = 5;                         // ... of S3: 5 views x 2 clouds x 16 KB of pv thresholds staged in shared memory
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
    const int64_t base = sp.src0 + (int64_t)blockIdx.x * P.seed_stride * FS_CHUNK;
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
// grid (ceil(V / FS_VPC_FILTER), chunks of P.filter_chunk points, B): the view groups of one chunk run back to back, so
// the chunk is read from HBM once and from L2 afterwards.
// Two steps.  (1) A point of cell a can only survive if its pv reaches  L[a] = min(eff[a] / FS_UP, thr[a])  (its value is
// at most pv: the Gaussian weight is <= 1).  L (one float per cell and image) is staged in shared memory for 5 views x 2
// clouds; per point and view this costs three float4 shared loads, ~30 FP32 ops and one random shared load.  The points
// that pass (the front layer and the rim of the cloud, a few per cent) are pushed on a shared-memory queue, one ballot per
// warp (every warp owns a queue; no block barrier in the loop).  (2) Whenever a warp's queue fills, and at the end, the warp processes its entries with all lanes busy: the full test
// against (eff, thr) from global memory, then one global atomic per warp and image to reserve list space.
constexpr int FS_WQ = 384;                               // queue entries per warp: (image of the CTA) << 20 | point of the chunk
constexpr size_t FS_FILTER_SMEM = (size_t)FS_VPC_FILTER * 2 * FC_CELLS * sizeof(float) + (size_t)(FS_FILTER_THREADS / 32) * FS_WQ * sizeof(unsigned);
static_assert(FS_WQ > 32 * FS_VPC_FILTER && FS_FILTER_CHUNK_MAX <= (1 << 20), "queue layout");

struct StreamViewPacked { float4 rx, ry, rz; };          // (R0..2, zmin), (R3..5, rcp_range), (R6..8, -)

__device__ __forceinline__ float ex2_ftz(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// the transform S1 and S3 share (same arithmetic as stream_point(): they must agree on the home cell)
struct FilterPoint { float ux, uy, pv_hi, cj, cr; bool inside; int cell; };
__device__ __forceinline__ FilterPoint filter_point(float x, float y, float z, const StreamViewPacked& W) {
    FilterPoint f;
    f.ux = __fmaf_rn(__fmaf_rn(W.rx.z, z, __fmaf_rn(W.rx.y, y, __fmul_rn(W.rx.x, x))), FC_CELLS_PER_NDC, FC_CELLS_PER_NDC);
    f.uy = __fmaf_rn(__fmaf_rn(W.ry.z, z, __fmaf_rn(W.ry.y, y, __fmul_rn(W.ry.x, x))), FC_CELLS_PER_NDC, FC_CELLS_PER_NDC);
    const float zr = __fmaf_rn(W.rz.z, z, __fmaf_rn(W.rz.y, y, __fmul_rn(W.rz.x, x)));
    f.pv_hi = __fmaf_rn(-(zr - W.rx.w), W.ry.w, 1.0f) + FS_PV_ABS;
    const float fj = floorf(f.ux), fi = floorf(f.uy);
    f.cj = fminf(fmaxf(fj, 0.0f), (float)(FC_G - 1)); f.cr = fminf(fmaxf(fi, 0.0f), (float)(FC_G - 1));
    f.inside = (f.cj == fj) && (f.cr == fi);             // really inside its cell (not clamped, not NaN)
    f.cell = (int)f.cr * FC_G + (int)f.cj;
    return f;
}

__global__ void __launch_bounds__(FS_FILTER_THREADS, 1)
stream_filter_kernel(StreamParams P)
{
    extern __shared__ float s_L[];                       // [FS_VPC_FILTER][2][4096], then unsigned s_q[32 warps][FS_WQ]
    unsigned* s_q = reinterpret_cast<unsigned*>(s_L + FS_VPC_FILTER * 2 * FC_CELLS);
    __shared__ StreamViewPacked s_view[FS_VPC_FILTER];
    __shared__ float s_pose[12];
    const int v0 = blockIdx.x * FS_VPC_FILTER, nv = min(FS_VPC_FILTER, P.V - v0), b = blockIdx.z;
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t base = sp.src0 + (int64_t)blockIdx.y * P.filter_chunk;
    if (base >= sp.end) return;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < nv) {
        const int v = v0 + tid;
        const float* R = P.views + (size_t)v * 9;
        const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
        s_view[tid].rx = make_float4(R[0], R[1], R[2], zmin);
        s_view[tid].ry = make_float4(R[3], R[4], R[5], 1.0f / z_range(zmin, zmax));
        s_view[tid].rz = make_float4(R[6], R[7], R[8], 0.0f);
    }
    const bool has_pose = P.pose != nullptr;
    if (has_pose && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];
    const size_t img0 = ((size_t)b * P.V + v0) * 2;
    const float2* bnd0 = P.bnd + img0 * FC_CELLS;
    for (int t = tid; t < nv * 2 * FC_CELLS; t += FS_FILTER_THREADS) {
        const float2 bd = __ldg(bnd0 + t);               // (eff, thr); eff: -1 unknown, 0 exactly zero -> L <= 0: everything passes
        s_L[t] = fminf(bd.x * (1.0f / FS_UP) * (1.0f - 1e-5f), bd.y);   // weight <= 1 + 2^-22 (ex2.approx)
    }
    __syncthreads();
    const float negkc = P.negk / (FC_CELLS_PER_NDC * FC_CELLS_PER_NDC);
    const float negkc2 = negkc * 1.4426950408889634f;   // exp(a) = 2^(a log2 e)
    // flushing tiny weights to zero is only safe when the weight AT THE HOME PIXEL (d2c <= 1/2 cell^2) cannot be tiny
    const bool ftz_ok = negkc * 0.5f > -60.0f;
    const int npts = (int)min((int64_t)P.filter_chunk, sp.end - base);
    const int split = (int)min((int64_t)npts, max((int64_t)0, sp.tgt0 - base));     // points [0, split) are source points

    // step 2 (per warp): the queued (image, point) pairs of this warp, all lanes busy
    unsigned* wq = s_q + (tid >> 5) * FS_WQ;
    int wcount = 0;                                      // warp-uniform
    auto drain = [&]() {
        for (int t0 = 0; t0 < wcount; t0 += 32) {
            const int t = t0 + lane;
            bool keep = false; int l = 0, li = 0;
            if (t < wcount) {
                const unsigned e = wq[t];
                l = (int)(e >> 20);
                const int pi = (int)(e & 0xFFFFFu);
                const int c = l & 1;
                const float* q = P.xyz + (size_t)(base + pi) * 3;
                float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
                if (has_pose && c == 0) apply_pose(s_pose, x, y, z);
                const FilterPoint f = filter_point(x, y, z, s_view[l >> 1]);
                li = (int)(base + pi - (c ? sp.tgt0 : sp.src0));
                keep = true;
                if (f.inside) {
                    const float2 bd = __ldg(bnd0 + (size_t)l * FC_CELLS + f.cell);
                    const float fx = f.ux - f.cj - 0.5f, fy = f.uy - f.cr - 0.5f;
                    const float d2c = __fmaf_rn(fx, fx, fy * fy);
                    const float wgt = ftz_ok ? ex2_ftz(d2c * negkc2) : exp_fast(d2c * negkc);
                    keep = (f.pv_hi * wgt * FS_UP >= bd.x) || (f.pv_hi >= bd.y);
                }
            }
            const unsigned km = __ballot_sync(0xFFFFFFFFu, keep);
            if (keep) {                                  // one global atomic per warp and image
                const unsigned peers = __match_any_sync(km, l);
                const int leader = __ffs(peers) - 1;
                unsigned gp = 0;
                if (lane == leader) gp = atomicAdd(&P.surv_cnt[img0 + l], (unsigned)__popc(peers));
                gp = __shfl_sync(peers, gp, leader) + __popc(peers & ((1u << lane) - 1u));
                if (gp < (unsigned)P.surv_cap) P.surv[(img0 + l) * (size_t)P.surv_cap + gp] = li;
            }
            __syncwarp();
        }
        wcount = 0;
    };

    // step 1: every point of the chunk against the pv thresholds of the CTA's views.  No block barrier in the loop: every
    // warp owns a queue (its fill level lives in a register) and drains it on its own.
    const unsigned lt = (1u << lane) - 1u;
    for (int p0 = 0; p0 < npts; p0 += FS_FILTER_THREADS) {
        const int pi = p0 + tid;
        const bool live = pi < npts;
        const int c = pi >= split;
        float x = 0.f, y = 0.f, z = 0.f;
        if (live) {
            const float* q = P.xyz + (size_t)(base + pi) * 3;
            x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
            if (has_pose && c == 0) apply_pose(s_pose, x, y, z);
        }
        const float floor_pv = live ? 0.0f : INFINITY;  // dead lanes never pass
        const float* Lc = s_L + c * FC_CELLS;
        const unsigned ebase = ((unsigned)c << 20) | (unsigned)pi;
        for (int vv = 0; vv < nv; ++vv) {
            const FilterPoint f = filter_point(x, y, z, s_view[vv]);
            const float lim = f.inside ? Lc[vv * 2 * FC_CELLS + f.cell] : 0.0f;    // clamped points are never dropped here
            const bool pass = f.pv_hi > floor_pv && f.pv_hi >= lim;
            const unsigned pm = __ballot_sync(0xFFFFFFFFu, pass);
            if (pm) {
                if (pass) wq[wcount + __popc(pm & lt)] = ebase + ((unsigned)vv << 21);
                wcount += __popc(pm);
            }
        }
        if (wcount > FS_WQ - 32 * FS_VPC_FILTER) { __syncwarp(); drain(); }      // one more round always fits
    }
    __syncwarp();
    drain();
}

}  // namespace scream
