// forward_stream.cuh -- giant clouds (BASELINE config 5: 10^6 .. 10^8 points per pair): an exact
// pre-filter that streams the points and keeps only those that can win a pixel, so that the culled
// kernel (forward_culled.cuh) renders a few thousand survivors per image instead of the whole cloud.
// This is north_star's "projection + z-buffer" stage: a per-pixel atomicMax on float bits, staged
// through a shared-memory image tile.
//
//   S1 seed    a sample of the points -> home cell (the cell of the 64 x 64 grid that holds the point; for w == 64 a cell
//              is a pixel) -> value at that pixel -> 64-bit atomicMax (value bits, point index) in a shared-memory tile,
//              flushed to the global seed image: per cell, one real point that does well at the cell's own pixel.
//   S2 bound   per image, in the score domain (s = d^2 + c, c = -ln(pv)/k): U(q) = best seed score at every pixel, its
//              squared-distance transform T at the cell corners, and a per-cell pv threshold (details at the kernel).
//   S2b masks  a voxel grid over each cloud's bounding box; per voxel one bit per view: "a point in here could pass that
//              view's cell test".  Conservative; lets S3 skip ~90 % of the (point, view) pairs on one 8-byte load.
//   S3 filter  every point, read once for all views: it survives iff it may reach U at some pixel -- voxel bit, then a pv
//              threshold per third of a cell, then the per-point test with the window line of its cell.  Survivors
//              (cloud-local indices) are appended to a per-image list.  Every point that attains a pixel's maximum
//              survives, including all points that tie.
//   round 2    (large clouds) seeds again from the survivors, plus per rim cell the survivor that does best at a point
//              out in the empty side; bounds again; the survivor lists filtered again, far candidates settled exactly.
//   S4 render  forward_culled_kernel on the survivor lists (FwdParams::surv*); an image whose list
//              overflowed is rendered from the full cloud in batches (correct, slow).
#pragma once
#include "common.cuh"
#include "forward_dense.cuh"
#include "forward_culled.cuh"

namespace scream {

constexpr int FS_THREADS = 1024;                        // S1: one CTA per SM (128 KB of tiles), 32 warps
constexpr int FS_FILTER_THREADS = 1024;
constexpr int FS_FILTER_CHUNK_MAX = 131072;              // points per CTA task of S3 (amortises staging the thresholds)
constexpr int FS_FILTER_CHUNK_MIN = 2048;
constexpr int FS_CHUNK = 32768;                          // points per sampled chunk of S1
constexpr int FS_SEED_CPB = 4;                           // sampled chunks per CTA of S1
constexpr int FS_SURV_MIN = 49152;                       // survivors kept per image: at least 49152, else total_pts / 16
constexpr int64_t FS_MIN_AVG_PTS = 147456;          // stream path when total_pts / B reaches this (147,456): measured crossover on
                                                         // B200 between 120 k and 240 k points per item (profiles/r01q_sweep_stream.jsonl)
constexpr float FS_DOWN = 1.0f - 1e-4f;                  // covers ex2.approx + argument rounding at |arg| <= 108
constexpr float FS_UP = 1.0f + 1e-4f;

struct StreamParams {
    const float* xyz; const int64_t* offs; const float* pose; const float* views; const unsigned* zenc;
    int B, V;
    float negk;                                          // -rho^2/2
    unsigned long long* seed;                            // [B][V][2][4096]  (value bits << 32 | cloud-local index) of the cell's seed point (S1)
    float* lpv;                                          // [B][V][2][4096]  pv a point of the cell needs at least (S2 -> S2b voxel masks)
    float* uimg;                                         // [B][V][2][4096]  U(q) (S2 -> the exact far test of the second round)
    float* lpv9;                                         // [B][V][2][192][192]  the same threshold per third of a cell (S2 -> S3 step 1)
    unsigned long long* seed2;                           // [B][V][2][4096]  second round: per rim cell the survivor that does best at the cell's AIM (~score bits, index)
    float2* aim;                                         // [B][V][2][4096]  a point FS_AIM cells outwards of a cell that borders empty space (NaN: none)
    float4* win;                                         // [B][V][2][4096][8] per cell, one 128-byte line: U of the 5 x 5 pixels around it (row-major,
                                                         //   -inf outside the image), then T_far at its 4 corners, 3 pads (S2 -> S3 step 2)
    int* surv; unsigned* surv_cnt; int surv_cap;         // [B][V][2][surv_cap], [B][V][2]   survivors of the first round
    int* surv2; unsigned* surv2_cnt; int surv2_cap;      // ... of the second round (re-seeded from the first round's survivors)
    int seed_stride;                                     // S1 seeds from every seed_stride-th chunk of the cloud (any subset gives valid bounds)
    int filter_chunk;                                    // points per CTA task of S3
    unsigned long long* stats;                           // [3] diagnostics: (point, view) pairs that passed the cell test / (unused) / the voxel mask of S3
    unsigned* bbox;                                      // [B][2][6] per cloud (source posed): max keys of ~enc(min x,y,z), enc(max x,y,z); zero-initialised
    unsigned long long* vox;                             // [B][view groups][2][vox_g^3] bit vv: a point of the voxel may pass view v0 + vv's cell test
    int vox_g;                                           // voxels per axis of the cloud's bounding box
    // point-sharded clouds: the first round's seeds come from ALL ranks.  seedtab [B][V][2][4096] (x', y', c_up,
    // value bits) of each cell's seed, exactly placed; built per rank by stream_seedtab_kernel, masked to the rank whose seed
    // is the best of the cell (all_reduce(MAX) over seed, rewritten as value bits << 32 | ~rank) and summed over the ranks
    // (all_reduce(SUM): one rank contributes, the others add zeros).  nullptr: seeds are local (P.seed holds point indices).
    float4* seedtab;
    int rank;
    // w == 128 (lattice passes, scream_render.cu): the image's pixel centres are four 64 x 64 lattices, the w == 64 lattice
    // moved by -+1/128 along x and y.  A pass filters for ONE of them in the frame that puts it on the cell centres (and
    // forward_culled_lattice_kernel then renders that lattice's pixels from the pass's lists): the
    // cell coordinate of a point is fl(32 x' + off) with off = 32 +- 0.25 (the additive constant of the same FMA, no extra
    // rounding) and the NDC position of a seed fl(x' + sh), sh = (off - 32) / 32 = +-1/128.  w == 64: off = 32, sh = 0.
    float offx, offy, shx, shy;
};

constexpr int FS_VPC = 2;                                // views per CTA task of S1 (2 views x 2 clouds x 32 KB of 64-bit tiles)
constexpr int FS_VPC_FILTER = 64;                        // ... of S3: one 64-bit voxel mask word covers them
#ifndef FS_VOX_BIG_MIN_N
#define FS_VOX_BIG_MIN_N (8 << 20)
#endif
constexpr int64_t FS_VOX_BIG_MIN = FS_VOX_BIG_MIN_N;       // points per item from which the 64^3 voxel grid replaces the 32^3 one
#ifndef FV_G_N
#define FV_G_N 64
#endif
constexpr int FV_G = FV_G_N;                                 // voxels per axis of a cloud's bounding box (S2b): g^3 mask words per group of 64 views
constexpr int FS_TC = (FC_G + 1) * (FC_G + 1);           // cell corners per image
constexpr float FS_AIM = 6.0f;                            // how far (cells) outside a rim cell its aim lies
constexpr int FS_SUB = 3;                                // S3 step 1 thresholds per cell and axis (odd: the pixel centre in the middle sub-cell)
constexpr float FS_PV_ABS = 4e-7f;                       // |1 - (z-zmin)*rcp(range) - pv_exact| (two roundings + rcp) stays below this

// fast exp for bounds only: ex2.approx(a * log2e), error <= 2 + 1.16|a| ulp (|a| <= 108: < 1.6e-5 relative)
__device__ __forceinline__ float exp_fast(float a) { return __expf(a); }

// per-view constants of a CTA task
struct StreamView { float R[9]; float zmin, rcp_range; };

struct StreamPoint { float pv, d2c; int cell; bool inside; };

// transform for BOUNDS: the renderer's rotation, pv through a reciprocal multiply (|error| <= FS_PV_ABS, applied in
// the safe direction by the callers), home cell = floor of the cell coordinate (clamped like the renderer's grid) and
// d2c = squared distance, in cells^2, to the pixel centre of that cell (w == 64: pixel = cell).  The cell function
// need not be bit-identical to the renderer's: S1 and S3 share it, and the bounds of S2 carry slack.
__device__ __forceinline__ StreamPoint stream_point(float x, float y, float z, const StreamView& W, float offx, float offy) {
    StreamPoint sp;
    const float ux = __fmaf_rn(rot_row(W.R, x, y, z), FC_CELLS_PER_NDC, offx);
    const float uy = __fmaf_rn(rot_row(W.R + 3, x, y, z), FC_CELLS_PER_NDC, offy);
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
// grid (chunks, ceil(V / FS_VPC), B).  Per cell and image: the sampled point with the best value at the cell's own pixel,
// as a 64-bit key (value bits, cloud-local index): atomicMax keeps value and identity together.  The value only ranks
// the candidates of a cell; S2 re-evaluates the chosen point exactly.
constexpr size_t FS_SEED_SMEM = (size_t)FS_VPC * 2 * FC_CELLS * sizeof(unsigned long long);

__global__ void __launch_bounds__(FS_THREADS)
stream_seed_kernel(StreamParams P)
{
    extern __shared__ unsigned long long s_tile[];       // [FS_VPC][2][4096]
    __shared__ StreamView s_view[FS_VPC];
    __shared__ float s_pose[12];
    const int v0 = blockIdx.y * FS_VPC, nv = min(FS_VPC, P.V - v0), b = blockIdx.z;
    const ItemSpan sp = item_span(P.offs, b);
    // the CTA takes FS_SEED_CPB sampled chunks (one flush of its tiles for all of them)
    if (sp.src0 + (int64_t)blockIdx.x * FS_SEED_CPB * P.seed_stride * FS_CHUNK >= sp.end) return;
    const int tid = threadIdx.x;
    stream_load_views(P, b, v0, nv, s_view);
    if (P.pose && tid < 12) s_pose[tid] = P.pose[(size_t)b * 12 + tid];
    for (int t = tid; t < nv * 2 * FC_CELLS; t += FS_THREADS) s_tile[t] = 0ull;
    __syncthreads();
    const float negkc = P.negk / (FC_CELLS_PER_NDC * FC_CELLS_PER_NDC);
    for (int kk = 0; kk < FS_SEED_CPB; ++kk) {
        const int64_t base = sp.src0 + ((int64_t)blockIdx.x * FS_SEED_CPB + kk) * P.seed_stride * FS_CHUNK;
        if (base >= sp.end) break;
        const int64_t stop = min(sp.end, base + FS_CHUNK);
        for (int64_t g0 = base + tid; g0 < stop; g0 += 4 * FS_THREADS) {     // four points in flight per thread
            float x[4], y[4], z[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t gi = g0 + (int64_t)u * FS_THREADS;
                x[u] = y[u] = z[u] = 0.0f;
                if (gi < stop) { const float* q = P.xyz + (size_t)gi * 3; x[u] = __ldg(q); y[u] = __ldg(q + 1); z[u] = __ldg(q + 2); }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t gi = g0 + (int64_t)u * FS_THREADS;
                if (gi >= stop) break;
                const int c = gi >= sp.tgt0;
                const unsigned li = (unsigned)(gi - (c ? sp.tgt0 : sp.src0));
                if (P.pose != nullptr && c == 0) apply_pose(s_pose, x[u], y[u], z[u]);
                for (int vv = 0; vv < nv; ++vv) {
                    const StreamPoint pt = stream_point(x[u], y[u], z[u], s_view[vv], P.offx, P.offy);
                    const float pv_lo = pt.pv - FS_PV_ABS;
                    if (!(pv_lo > 0.0f) || !pt.inside) continue;     // clamped points may be anywhere outside: they cannot seed
                    const float val = pv_lo * home_weight(pt, negkc);
                    const unsigned bits = __float_as_uint(val);
                    unsigned long long* slot = &s_tile[(vv * 2 + c) * FC_CELLS + pt.cell];
                    if (val > 0.0f && bits > (unsigned)(*reinterpret_cast<volatile unsigned long long*>(slot) >> 32))
                        atomicMax(slot, ((unsigned long long)bits << 32) | li);
                }
            }
        }
    }
    __syncthreads();
    unsigned long long* out = P.seed + ((size_t)b * P.V + v0) * 2 * FC_CELLS;
    for (int t = tid; t < nv * 2 * FC_CELLS; t += FS_THREADS) {
        const unsigned long long key = s_tile[t];
        if (key != 0ull && key > out[t]) atomicMax(&out[t], key);
    }
}

// Slack of the score tests (NDC^2).  Two points whose VALUES the renderer may order either way have scores within
// 4.8e-7 s + 6e-7 / k of each other (rounding of d2, of the exp argument, expf <= 2 ulp, the final product).  Everything
// proportional to s -- that, and the rounding of U, of the two transform passes, of the interpolation and of the point's
// cell position (together < 1.3e-6 s) -- is covered by rounding U UP by 6e-6 (U(q) is the score at stake at the critical
// pixel); pv is rounded in the safe direction on both sides.  What is left is the additive part:
__device__ __forceinline__ float stream_slack(float inv_k) { return __fmaf_rn(2e-6f, inv_k, 1e-10f); }

// ---- S2: per-image bounds, in the score domain -----------------------------------------------------
// With k = rho^2/2 and c_i = -ln(pv_i)/k the value of point i at pixel q is exp(-k s_i(q)), s_i(q) = |x_i - q|^2 + c_i,
// so "point i reaches the best value seen at q" reads  s_i(q) <= U(q)  -- no exp, no underflow.
//   U(q)  = min over the seed points (exact positions, c rounded up) of s_seed(q), capped where the value is exactly 0:
//           an upper bound of the minimum score at pixel q.  Ring search outwards from the pixel's cell with early exit.
//   T(X)  = min_q |X - q|^2 - U(q)  at the 65 x 65 cell CORNERS X: a squared-distance transform, separable, exact.
//           Point i can reach U at some pixel iff T(x_i) <= -c_i.  m(x) = T(x) - |x|^2 is a minimum of affine functions,
//           hence concave, so inside a cell it lies above the bilinear interpolation of its corner values:
//               T(x) >= bilerp(T corners)(t) - h^2 (t_x (1 - t_x) + t_y (1 - t_y)),   t = position in the cell, h = cell size
//           (exact whenever one pixel q is the minimiser over the whole cell, which is the rule away from the cloud).
//           This is the per-POINT test of S3 step 2: four floats per cell, no cell-level slack on the point's position.
//   lpv[a] = exp(-k Cmax[a]),  Cmax[a] = h^2/2 - min(T at the 4 corners of a): no point of cell a with pv below it passes
//           the per-point test -- the cheap per-cell threshold of S3 step 1.
// grid (V * 2, B), 1024 threads.
constexpr size_t FS_BOUND_SMEM = (size_t)FC_CELLS * sizeof(float4) + (size_t)FC_CELLS * sizeof(float)
                               + (size_t)2 * (FC_G + 1) * FC_G * sizeof(float) + (size_t)FS_TC * sizeof(float);

// ---- point-sharded clouds: this rank's seed table, and its mask after the cross-rank MAX --------------------------------
// grid (V * 2, B) x 1024.  For every cell the seed's (x', y', c_up) exactly as stream_bound_kernel places it, + its value bits;
// P.seed is rewritten from (value bits << 32 | point index) to (value bits << 32 | ~rank): the collective's payload.
__global__ void __launch_bounds__(1024)
stream_seedtab_kernel(StreamParams P)
{
    __shared__ float s_R[9], s_pose[12];
    const int vc = blockIdx.x, v = vc >> 1, c = vc & 1, b = blockIdx.y, tid = threadIdx.x;
    const size_t img = ((size_t)b * P.V + v) * 2 + c;
    if (tid < 9) s_R[tid] = P.views[(size_t)v * 9 + tid];
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t lo = c ? sp.tgt0 : sp.src0;
    const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
    const float rcp_range = 1.0f / z_range(zmin, zmax);
    const float inv_k = -1.0f / P.negk;
    __syncthreads();
    for (int cell = tid; cell < FC_CELLS; cell += 1024) {
        const unsigned long long key = P.seed[img * FC_CELLS + cell];
        float4 sd = make_float4(0.f, 0.f, 0.f, 0.f);
        unsigned long long out = 0ull;
        if (key != 0ull) {
            const float* q = P.xyz + (size_t)(lo + (int64_t)(unsigned)(key & 0xFFFFFFFFull)) * 3;
            float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
            if (posed) apply_pose(s_pose, x, y, z);
            const float pv_lo = __fmaf_rn(-(rot_row(s_R + 6, x, y, z) - zmin), rcp_range, 1.0f) - FS_PV_ABS;
            if (pv_lo > 0.0f) {
                sd = make_float4(rot_row(s_R, x, y, z) + P.shx, rot_row(s_R + 3, x, y, z) + P.shy,
                                 fmaxf(-logf(pv_lo) * inv_k, 0.0f) * (1.0f + 1e-5f) + 1e-12f, __uint_as_float((unsigned)(key >> 32)));
                out = (key & 0xFFFFFFFF00000000ull) | (unsigned long long)(0xFFFFFFFFu - (unsigned)P.rank);
            }
        }
        P.seedtab[img * FC_CELLS + cell] = sd;
        P.seed[img * FC_CELLS + cell] = out;
    }
}
// after all_reduce(MAX) over P.seed: a rank whose seed is not the cell's best clears its table entry (the SUM over the ranks
// then holds exactly the best rank's entry)
__global__ void __launch_bounds__(256)
stream_seedtab_mask_kernel(StreamParams P, long long total)
{
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const unsigned long long g = P.seed[t];
    const float4 sd = P.seedtab[t];
    const bool mine = g != 0ull && (unsigned)(g >> 32) == __float_as_uint(sd.w) && (unsigned)(g & 0xFFFFFFFFull) == 0xFFFFFFFFu - (unsigned)P.rank
                      && sd.z > 0.0f;             // c_up >= 1e-12 for every placed seed
    if (!mine) P.seedtab[t] = make_float4(0.f, 0.f, 0.f, 0.f);
}

__global__ void __launch_bounds__(1024)
stream_bound_kernel(StreamParams P)
{
    extern __shared__ float4 s_seedpt[];                 // [4096] (x', y', c_up, -) of the cell's seed; c = +inf: none
    float* s_U = reinterpret_cast<float*>(s_seedpt + FC_CELLS);          // [4096]
    float* s_A = s_U + FC_CELLS;                                         // [64][65] row pass: [pixel row][corner column]
    float* s_E = s_A + (FC_G + 1) * FC_G;                                // [64][65] row pass without the 4 columns around X
    float* s_T = s_E + (FC_G + 1) * FC_G;                                // [65][65]
    __shared__ float s_R[9], s_pose[12];
    const int vc = blockIdx.x, v = vc >> 1, c = vc & 1, b = blockIdx.y, tid = threadIdx.x;
    const size_t img = ((size_t)b * P.V + v) * 2 + c;
    if (tid < 9) s_R[tid] = P.views[(size_t)v * 9 + tid];
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t lo = c ? sp.tgt0 : sp.src0;
    const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
    const float rcp_range = 1.0f / z_range(zmin, zmax);
    const float inv_k = -1.0f / P.negk;
    const float h = 1.0f / FC_CELLS_PER_NDC;
    const float s_zero_up = FC_EXP_ZERO * inv_k * (1.0f + 1e-4f);       // scores beyond: value exactly 0
    __syncthreads();

    // seed points, exactly placed
    for (int cell = tid; cell < FC_CELLS; cell += 1024) {
        const unsigned long long key = P.seed[img * FC_CELLS + cell];
        float4 sd = make_float4(0.f, 0.f, INFINITY, 0.f);
        if (P.seedtab != nullptr) {                       // the best seed of ANY rank (see StreamParams::seedtab); key 0: none anywhere
            if (key != 0ull) { sd = P.seedtab[img * FC_CELLS + cell]; sd.w = 0.f; if (!(sd.z > 0.0f)) sd.z = INFINITY; }
        } else if (key != 0ull) {
            const float* q = P.xyz + (size_t)(lo + (int64_t)(unsigned)(key & 0xFFFFFFFFull)) * 3;
            float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
            if (posed) apply_pose(s_pose, x, y, z);
            const float pv_lo = __fmaf_rn(-(rot_row(s_R + 6, x, y, z) - zmin), rcp_range, 1.0f) - FS_PV_ABS;
            if (pv_lo > 0.0f)
                sd = make_float4(rot_row(s_R, x, y, z) + P.shx, rot_row(s_R + 3, x, y, z) + P.shy,
                                 fmaxf(-logf(pv_lo) * inv_k, 0.0f) * (1.0f + 1e-5f) + 1e-12f, 0.f);
        }
        s_seedpt[cell] = sd;
    }
    __syncthreads();

    // Rim cells (occupied, empty cells within 3): their AIM, a point FS_AIM cells out into the empty side.  The second round
    // keeps, per rim cell, the survivor that does best THERE as an extra seed: for the far pixels of the empty side the
    // cell's own seed (best at the cell's pixel) is a poor bound -- the outermost point wins them, by ~D h in score.
    for (int cell = tid; cell < FC_CELLS; cell += 1024) {
        const int i = cell >> 6, j = cell & 63;
        float2 a = make_float2(NAN, NAN);
        if (s_seedpt[cell].z < INFINITY) {
            int sx = 0, sy = 0;
            for (int di = -3; di <= 3; ++di)
                for (int dj = -3; dj <= 3; ++dj) {
                    const int qi = i + di, qj = j + dj;
                    if (qi >= 0 && qi < FC_G && qj >= 0 && qj < FC_G && !(s_seedpt[qi * FC_G + qj].z < INFINITY)) { sx += dj; sy += di; }
                }
            if (sx != 0 || sy != 0) {
                const float n = FS_AIM * h * rsqrtf((float)(sx * sx + sy * sy));
                a = make_float2(((float)j + 0.5f) * h - 1.0f + (float)sx * n, ((float)i + 0.5f) * h - 1.0f + (float)sy * n);
            }
        }
        P.aim[img * FC_CELLS + cell] = a;
    }
    // extra seeds of the second round, exactly placed, as a compact list (in s_A, which is free until the row pass)
    __shared__ int s_nextra;
    float4* s_extra = reinterpret_cast<float4*>(s_A);
    constexpr int FS_EXTRA_MAX = (FC_G + 1) * FC_G / 4;
    if (tid == 0) s_nextra = 0;
    __syncthreads();
    for (int cell = tid; cell < FC_CELLS; cell += 1024) {
        const unsigned long long key = P.seed2[img * FC_CELLS + cell];
        if (key == 0ull) continue;
        const float* q = P.xyz + (size_t)(lo + (int64_t)(unsigned)(key & 0xFFFFFFFFull)) * 3;
        float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
        if (posed) apply_pose(s_pose, x, y, z);
        const float pv_lo = __fmaf_rn(-(rot_row(s_R + 6, x, y, z) - zmin), rcp_range, 1.0f) - FS_PV_ABS;
        if (!(pv_lo > 0.0f)) continue;
        const int slot = atomicAdd(&s_nextra, 1);
        if (slot < FS_EXTRA_MAX)
            s_extra[slot] = make_float4(rot_row(s_R, x, y, z) + P.shx, rot_row(s_R + 3, x, y, z) + P.shy,
                                        fmaxf(-logf(pv_lo) * inv_k, 0.0f) * (1.0f + 1e-5f) + 1e-12f, 0.f);
    }
    __syncthreads();
    const int nextra = min(s_nextra, FS_EXTRA_MAX);

    // U(q): rings of cells around the pixel's own cell, until no farther seed can beat the best so far
    for (int q = tid; q < FC_CELLS; q += 1024) {
        const int i = q >> 6, j = q & 63;
        const float qx = ((float)j + 0.5f) * h - 1.0f, qy = ((float)i + 0.5f) * h - 1.0f;      // pixel centre (w == 64), exact
        float best = s_zero_up;
        for (int d = 0; d < FC_G; ++d) {
            const float reach = ((float)d - 0.5f) * h;                       // seeds lie inside their cells
            if (d > 0 && reach * reach * (1.0f - 1e-5f) >= best) break;
            const int i0 = i - d, i1 = i + d, j0 = max(j - d, 0), j1 = min(j + d, FC_G - 1);
            if (i0 >= 0) for (int jj = j0; jj <= j1; ++jj) {
                const float4 sd = s_seedpt[i0 * FC_G + jj];
                const float dx = sd.x - qx, dy = sd.y - qy;
                best = fminf(best, __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, sd.z)));
            }
            if (d > 0 && i1 < FC_G) for (int jj = j0; jj <= j1; ++jj) {
                const float4 sd = s_seedpt[i1 * FC_G + jj];
                const float dx = sd.x - qx, dy = sd.y - qy;
                best = fminf(best, __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, sd.z)));
            }
            for (int ii = max(i0 + 1, 0); ii <= min(i1 - 1, FC_G - 1); ++ii) {
                if (j - d >= 0) {
                    const float4 sd = s_seedpt[ii * FC_G + j - d];
                    const float dx = sd.x - qx, dy = sd.y - qy;
                    best = fminf(best, __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, sd.z)));
                }
                if (j + d < FC_G) {
                    const float4 sd = s_seedpt[ii * FC_G + j + d];
                    const float dx = sd.x - qx, dy = sd.y - qy;
                    best = fminf(best, __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, sd.z)));
                }
            }
        }
        for (int e = 0; e < nextra; ++e) {                                   // any real point bounds the minimum
            const float4 sd = s_extra[e];
            const float dx = sd.x - qx, dy = sd.y - qy;
            best = fminf(best, __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, sd.z)));
        }
        // Denormal range (k U > 80): values are multiples of 2^-149, so scores up to -ln(1 - u)/k <= 1.5 u/k apart
        // (u = 8 * 2^-149 * e^(k U)) may tie or swap; from u >= 1/2 on every non-zero value may.  Same rule as the
        // renderer's exact pass (forward_culled.cuh).
        // (lattice passes: a seed's x' + sh is rounded once more, <= 3e-8 per coordinate: 2 d 3e-8 <= 1e-7 sqrt(score))
        float up = best * (1.0f + 6e-6f) + ((P.shx != 0.0f || P.shy != 0.0f) ? 1e-7f * sqrtf(best) : 0.0f);
        const float e1 = up * -P.negk;
        if (e1 >= FC_EXP_DENORM) {
            const float lu = e1 - 101.2f;
            up = lu >= -0.7f ? s_zero_up : up + 1.6f * expf(lu) * inv_k;
        }
        s_U[q] = fminf(up, s_zero_up);
    }
    __syncthreads();

    // T_far at the corners: pixels outside the 4 x 4 block around corner (Y, X) -- every pixel that is NOT in the 5 x 5
    // neighbourhood of a cell touching the corner.  Row pass over all / over the far columns, then the column pass.
    for (int t = tid; t < (FC_G + 1) * FC_G; t += 1024) {
        const int qy = t / (FC_G + 1), X = t % (FC_G + 1);                   // lanes along X: U is read as a broadcast
        float a = INFINITY, e = INFINITY;
        for (int qx = 0; qx < FC_G; ++qx) {
            const float dx = ((float)(X - qx) - 0.5f) * h;
            const float val = __fmaf_rn(dx, dx, -s_U[qy * FC_G + qx]);
            a = fminf(a, val);
            if (qx < X - 2 || qx > X + 1) e = fminf(e, val);
        }
        s_A[qy * (FC_G + 1) + X] = a; s_E[qy * (FC_G + 1) + X] = e;      // [qy][X]: both passes free of bank conflicts
    }
    __syncthreads();
    for (int t = tid; t < FS_TC; t += 1024) {
        const int Y = t / (FC_G + 1), X = t % (FC_G + 1);
        float m = INFINITY;
        for (int qy = 0; qy < FC_G; ++qy) {
            const float dy = ((float)(Y - qy) - 0.5f) * h;
            const bool far_row = qy < Y - 2 || qy > Y + 1;
            m = fminf(m, __fmaf_rn(dy, dy, far_row ? s_A[qy * (FC_G + 1) + X] : s_E[qy * (FC_G + 1) + X]));
        }
        s_T[t] = m;
    }
    __syncthreads();
    // per-cell pv threshold: the largest c a point of the cell may have and still pass the per-point test of S3:
    // near pixels exactly (smallest distance of the cell to the pixel centre), far pixels through the corner bound
    for (int cell = tid; cell < FC_CELLS; cell += 1024) {
        const int i = cell >> 6, j = cell & 63;
        const float tmin = fminf(fminf(s_T[i * (FC_G + 1) + j], s_T[i * (FC_G + 1) + j + 1]),
                                 fminf(s_T[(i + 1) * (FC_G + 1) + j], s_T[(i + 1) * (FC_G + 1) + j + 1]));
        float cmax = 0.5f * h * h - tmin;
        for (int di = -2; di <= 2; ++di) {
            const int qi = i + di;
            if (qi < 0 || qi >= FC_G) continue;
            const float ay = fmaxf(fabsf((float)di) - 0.5f, 0.0f) * h;
            for (int dj = -2; dj <= 2; ++dj) {
                const int qj = j + dj;
                if (qj < 0 || qj >= FC_G) continue;
                const float ax = fmaxf(fabsf((float)dj) - 0.5f, 0.0f) * h;
                cmax = fmaxf(cmax, s_U[qi * FC_G + qj] - (ax * ax + ay * ay) * (1.0f - 1e-6f));
            }
        }
        cmax += stream_slack(inv_k);
        // pv >= exp(-k cmax); cmax < 0: nothing can pass (pv <= 1)
        P.lpv[img * FC_CELLS + cell] = expf(fmaxf(P.negk * cmax, -120.0f)) * (1.0f - 1e-4f);
    }
    // largest U over the 5 x 5 window of every cell (in s_A, free again): where it is small, the outer rings of the window
    // cannot carry the maximum below -- the own pixel's term is >= -(0.236 h)^2, a ring-1 term <= Umax - (0.5 h)^2, a ring-2
    // term <= Umax - (1.5 h)^2 -- and are skipped
    __syncthreads();
    for (int cell = tid; cell < FC_CELLS; cell += 1024) {
        const int i = cell >> 6, j = cell & 63;
        float um = 0.0f;
        for (int qi = max(i - 2, 0); qi <= min(i + 2, FC_G - 1); ++qi)
            for (int qj = max(j - 2, 0); qj <= min(j + 2, FC_G - 1); ++qj) um = fmaxf(um, s_U[qi * FC_G + qj]);
        s_A[cell] = um;
    }
    __syncthreads();
    // the same threshold per THIRD of a cell (3 x 3 sub-cells; the pixel centre lies in the middle one): in a dense cloud
    // the winner of a pixel sits within a fraction of a pixel of its centre, and the other eight sub-cells refuse outright
    for (int t = tid; t < FS_SUB * FC_G * FS_SUB * FC_G; t += 1024) {
        const int i3 = t / (FS_SUB * FC_G), j3 = t % (FS_SUB * FC_G), i = i3 / FS_SUB, j = j3 / FS_SUB;
        const int sa = j3 - FS_SUB * j, sb = i3 - FS_SUB * i;
        const float tmin = fminf(fminf(s_T[i * (FC_G + 1) + j], s_T[i * (FC_G + 1) + j + 1]),
                                 fminf(s_T[(i + 1) * (FC_G + 1) + j], s_T[(i + 1) * (FC_G + 1) + j + 1]));
        float cmax = 0.5f * h * h - tmin;
        // the sub-cell's box relative to the cell's pixel centre, in cells, widened by the rounding of the classification
        const float x0 = (float)sa / FS_SUB - 0.5f - 1e-4f, x1 = (float)(sa + 1) / FS_SUB - 0.5f + 1e-4f;
        const float y0 = (float)sb / FS_SUB - 0.5f - 1e-4f, y1 = (float)(sb + 1) / FS_SUB - 0.5f + 1e-4f;
        const float um = s_A[i * FC_G + j];
        const int rad = um <= 0.18f * h * h ? 0 : (um <= 2.0f * h * h ? 1 : 2);
        for (int di = -rad; di <= rad; ++di) {
            const int qi = i + di;
            if (qi < 0 || qi >= FC_G) continue;
            const float ay = fmaxf(fmaxf((float)di - y1, y0 - (float)di), 0.0f) * h;
            for (int dj = -rad; dj <= rad; ++dj) {
                const int qj = j + dj;
                if (qj < 0 || qj >= FC_G) continue;
                const float ax = fmaxf(fmaxf((float)dj - x1, x0 - (float)dj), 0.0f) * h;
                cmax = fmaxf(cmax, s_U[qi * FC_G + qj] - (ax * ax + ay * ay) * (1.0f - 1e-6f));
            }
        }
        cmax += stream_slack(inv_k);
        P.lpv9[img * (size_t)(FS_SUB * FC_G * FS_SUB * FC_G) + t] = expf(fmaxf(P.negk * cmax, -120.0f)) * (1.0f - 1e-4f);
    }
    for (int q = tid; q < FC_CELLS; q += 1024) P.uimg[img * FC_CELLS + q] = s_U[q];
    // the per-point test reads one 128-byte line per cell: 25 window values of U, 4 corner values of T_far
    float* win = reinterpret_cast<float*>(P.win + img * (size_t)FC_CELLS * 8);
    for (int t = tid; t < FC_CELLS * 32; t += 1024) {
        const int cell = t >> 5, k = t & 31, i = cell >> 6, j = cell & 63;
        float val = 0.0f;
        if (k < 25) {
            const int qi = i + k / 5 - 2, qj = j + k % 5 - 2;
            val = (qi >= 0 && qi < FC_G && qj >= 0 && qj < FC_G) ? s_U[qi * FC_G + qj] : -INFINITY;
        } else if (k < 29) {
            val = s_T[(i + ((k - 25) >> 1)) * (FC_G + 1) + j + ((k - 25) & 1)];
        }
        win[t] = val;
    }
}

// ---- S2b: voxel masks ----------------------------------------------------------------------------
// The cell test of S3 costs ~50 instructions per (point, view) and only a few per cent pass.  Most pairs can be refused
// without looking at the view: a voxel grid over each cloud's bounding box, and per voxel and view one bit -- "some point
// inside this voxel could pass the cell test" (largest pv over the voxel against the smallest threshold over the cells
// its projection touches, with slack for every rounding involved; conservative, never exact).  S3 then reads one byte
// per point and view group and only queues the pairs whose bit is set.
constexpr int FV_BBOX_THREADS = 256, FV_BBOX_PTS = FV_BBOX_THREADS * 32;

// grid (chunks of FV_BBOX_PTS points, B): bounding box of the source cloud (posed, the arithmetic of every other stage)
// and of the target cloud.
__global__ void __launch_bounds__(FV_BBOX_THREADS)
stream_bbox_kernel(StreamParams P)
{
    __shared__ float s_pose[12];
    __shared__ unsigned s_key[12];
    const int b = blockIdx.y, tid = threadIdx.x;
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t base = sp.src0 + (int64_t)blockIdx.x * FV_BBOX_PTS;
    if (base >= sp.end) return;
    if (P.pose && tid < 12) s_pose[tid] = P.pose[(size_t)b * 12 + tid];
    if (tid >= 32 && tid < 44) s_key[tid - 32] = 0u;
    __syncthreads();
    float lo[2][3], hi[2][3];
#pragma unroll
    for (int k = 0; k < 6; ++k) { lo[k / 3][k % 3] = INFINITY; hi[k / 3][k % 3] = -INFINITY; }
    const int64_t stop = min(sp.end, base + FV_BBOX_PTS);
    for (int64_t g0 = base + tid; g0 < stop; g0 += 4 * FV_BBOX_THREADS) {      // four points in flight per thread
      float px[4], py[4], pz[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t gi = g0 + (int64_t)u * FV_BBOX_THREADS;
        px[u] = py[u] = pz[u] = 0.0f;
        if (gi < stop) { const float* q = P.xyz + (size_t)gi * 3; px[u] = __ldg(q); py[u] = __ldg(q + 1); pz[u] = __ldg(q + 2); }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t gi = g0 + (int64_t)u * FV_BBOX_THREADS;
        if (gi >= stop) break;
        float x = px[u], y = py[u], z = pz[u];
        if (gi < sp.tgt0) {
            if (P.pose) apply_pose(s_pose, x, y, z);
            lo[0][0] = fminf(lo[0][0], x); lo[0][1] = fminf(lo[0][1], y); lo[0][2] = fminf(lo[0][2], z);
            hi[0][0] = fmaxf(hi[0][0], x); hi[0][1] = fmaxf(hi[0][1], y); hi[0][2] = fmaxf(hi[0][2], z);
        } else {
            lo[1][0] = fminf(lo[1][0], x); lo[1][1] = fminf(lo[1][1], y); lo[1][2] = fminf(lo[1][2], z);
            hi[1][0] = fmaxf(hi[1][0], x); hi[1][1] = fmaxf(hi[1][1], y); hi[1][2] = fmaxf(hi[1][2], z);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < 2; ++c) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const unsigned kl = __reduce_max_sync(0xFFFFFFFFu, ~enc_f32(lo[c][a]));
            const unsigned kh = __reduce_max_sync(0xFFFFFFFFu, enc_f32(hi[c][a]));
            if ((tid & 31) == 0) { atomicMax(&s_key[c * 6 + a], kl); atomicMax(&s_key[c * 6 + 3 + a], kh); }
        }
    }
    __syncthreads();
    if (tid < 12) atomicMax(P.bbox + (size_t)b * 12 + tid, s_key[tid]);
}

struct VoxelGrid { float lo[3], inv[3]; bool ok; };       // voxel index along a = min(int((p[a] - lo[a]) * inv[a]), g - 1)
__device__ __forceinline__ VoxelGrid voxel_grid(const unsigned* __restrict__ bb, int g, float* ext) {
    VoxelGrid G; G.ok = true;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = dec_f32(~bb[a]), hi = dec_f32(bb[3 + a]);
        const float e = hi - lo;
        G.ok = G.ok && (e >= 0.0f) && (e < INFINITY);                       // an empty cloud or NaN / inf coordinates: no grid
        G.lo[a] = lo; G.inv[a] = e > 0.0f ? (float)g / e : 0.0f;
        if (ext) ext[a] = e;
    }
    return G;
}

// grid (ceil(g^3 / 256), view groups * 2 clouds, B): one thread per voxel, the FS_VPC_FILTER views of the group.
constexpr int FV_FOOT_MAX = 8;                            // projections wider than this many cells: keep
constexpr int FV_VIEWS_PER_THREAD = 16;                   // a quarter of a mask word per thread

__global__ void __launch_bounds__(256)
stream_voxmask_kernel(StreamParams P)
{
    __shared__ StreamView s_view[FV_VIEWS_PER_THREAD];
    const int g = P.vox_g, nvox = g * g * g;
    // blockIdx.y = ((view group * 2 + cloud) * 4 + quarter): a thread does 16 of the 64 views of its voxel's word
    const int quarter = blockIdx.y & 3, c = (blockIdx.y >> 2) & 1, grp = blockIdx.y >> 3, b = blockIdx.z;
    const int v0 = grp * FS_VPC_FILTER + quarter * FV_VIEWS_PER_THREAD, nv = max(0, min(FV_VIEWS_PER_THREAD, P.V - v0));
    stream_load_views(P, b, v0, nv, s_view);
    __syncthreads();
    const int vox = blockIdx.x * 256 + threadIdx.x;
    if (vox >= nvox) return;
    const int groups = (P.V + FS_VPC_FILTER - 1) / FS_VPC_FILTER;
    unsigned short* out = reinterpret_cast<unsigned short*>(P.vox + (((size_t)b * groups + grp) * 2 + c) * (size_t)nvox + vox) + quarter;
    float ext[3];
    const VoxelGrid G = voxel_grid(P.bbox + ((size_t)b * 2 + c) * 6, g, ext);
    if (!G.ok) { *out = 0xFFFFu; return; }
    const int iv[3] = { vox % g, (vox / g) % g, vox / (g * g) };
    float cp[3], hs[3], M = 0.0f;
#pragma unroll
    for (int a = 0; a < 3; ++a) M = fmaxf(M, fmaxf(fabsf(G.lo[a]), fabsf(G.lo[a] + ext[a])));
    const float rnd = 2e-6f * M + 1e-30f;                 // rounding of a rotated coordinate, here plus in the per-point transform
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float s = ext[a] / (float)g;
        cp[a] = G.lo[a] + ((float)iv[a] + 0.5f) * s;
        hs[a] = 0.5f * s * (1.0f + 1e-3f) + rnd;          // index rounding (<= 2e-5 voxels) and the centre's own rounding
    }
    unsigned mask = 0u;
    for (int vv = 0; vv < nv; ++vv) {
        const StreamView& W = s_view[vv];
        const float uxc = __fmaf_rn(rot_row(W.R, cp[0], cp[1], cp[2]), FC_CELLS_PER_NDC, P.offx);
        const float uyc = __fmaf_rn(rot_row(W.R + 3, cp[0], cp[1], cp[2]), FC_CELLS_PER_NDC, P.offy);
        const float zc = rot_row(W.R + 6, cp[0], cp[1], cp[2]);
        const float ex = (fabsf(W.R[0]) * hs[0] + fabsf(W.R[1]) * hs[1] + fabsf(W.R[2]) * hs[2] + rnd) * FC_CELLS_PER_NDC + 1e-4f;
        const float ey = (fabsf(W.R[3]) * hs[0] + fabsf(W.R[4]) * hs[1] + fabsf(W.R[5]) * hs[2] + rnd) * FC_CELLS_PER_NDC + 1e-4f;
        const float ez = fabsf(W.R[6]) * hs[0] + fabsf(W.R[7]) * hs[1] + fabsf(W.R[8]) * hs[2] + rnd;
        const float fj0 = floorf(uxc - ex), fj1 = floorf(uxc + ex), fi0 = floorf(uyc - ey), fi1 = floorf(uyc + ey);
        bool keep = !(fj0 >= 0.0f && fi0 >= 0.0f && fj1 < (float)FC_G && fi1 < (float)FC_G);   // touches the outside (or NaN)
        keep = keep || fj1 - fj0 >= (float)FV_FOOT_MAX || fi1 - fi0 >= (float)FV_FOOT_MAX;
        if (!keep) {
            const float pv_max = __fmaf_rn(-((zc - ez) - W.zmin), W.rcp_range, 1.0f) + (FS_PV_ABS + 1e-6f);
            const float* L = P.lpv + (((size_t)b * P.V + v0 + vv) * 2 + c) * FC_CELLS;
            float lim = INFINITY;
            for (int i = (int)fi0; i <= (int)fi1; ++i)
                for (int j = (int)fj0; j <= (int)fj1; ++j) lim = fminf(lim, __ldg(L + i * FC_G + j));
            keep = pv_max >= lim;
        }
        if (keep) mask |= 1u << vv;
    }
    *out = (unsigned short)mask;
}

// ---- S3: filter --------------------------------------------------------------------------------
// grid (ceil(V / 64), 2 clouds x chunks of P.filter_chunk points, B).  A CTA works on one cloud's points and (up to) 64
// views; every point is read once.  Per round a warp takes 32 points:
//  (0) the 64-bit voxel mask of each point (S2b; one 8-byte load, the tables live in L2) -> the (view, point) pairs whose
//      bit is set are appended to the warp's pair queue, a ring in shared memory (prefix sum over the lanes' bit counts,
//      32 views at a time);
//  (1) the queue is worked off 32 pairs at a time, all lanes busy: coordinates from the warp's point ring, transform,
//      pv against the threshold of the point's cell (S2; one 4-byte load from L2); pairs that pass (the front layer and
//      the rim of the cloud, a few per cent) go on the warp's step-2 queue;
//  (2) whenever that queue holds 32: the per-point test T_lowerbound(x_i) <= -c_i (S2) with the 128-byte line of the
//      point's cell, then one global atomic per warp and image to reserve list space.
// Every point that attains a pixel's maximum passes -- all points that tie included -- so rendering the survivors is
// bit-identical to rendering the cloud.  No block barrier in the loop (queues are per warp, fill levels in registers).
#ifndef FS_HEAVY_N
#define FS_HEAVY_N 6
#endif
constexpr int FS_HEAVY = FS_HEAVY_N;                     // S3: from this many views on, a point's pairs are queued by the whole warp
constexpr int FS_WQ = 64;                                // step-2 queue entries per warp: view << 21 | cloud << 20 | point of the chunk
constexpr int FS_Q1 = 2048;                              // pair queue per warp, a ring (u16: slot << 11 | view << 5 | lane): < 32 left over + 32 * 32 per half round
constexpr size_t FS_FILTER_SMEM = (size_t)(FS_FILTER_THREADS / 32) * (FS_WQ * sizeof(unsigned) + FS_Q1 * sizeof(unsigned short) + 2 * 96 * sizeof(float))
                                + (size_t)3 * FS_VPC_FILTER * 8 * sizeof(float4);
static_assert(FS_WQ >= 64 && FS_FILTER_CHUNK_MAX <= (1 << 20) && FS_VPC_FILTER == 64 && 32 * 32 + 32 <= FS_Q1, "queue layout, one mask word per view group");

struct StreamViewPacked { float4 rx, ry, rz; };          // (R0..2, zmin), (R3..5, rcp_range), (R6..8, -)

__device__ __forceinline__ float ex2_ftz(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// the transform S1 and S3 share (same arithmetic as stream_point(): they must agree on the home cell)
// Points outside the cell grid (outside the image; in a lattice pass of a w == 128 render also the quarter-cell strip that the
// shifted frame leaves out): every pixel centre lies inside the grid, a convex box, so the point's projection onto the box is
// at least as close to EVERY pixel as the point itself -- a point whose projection (same depth weight) cannot reach U anywhere
// cannot either.  They are tested at the projected position (ux, uy clamped into the border cell), not kept wholesale: at
// w == 128 a third of a unit ball's final lists were strip points, hundreds of records in a few border cells that every far
// pixel near them walked.  NaN coordinates are kept (the renderer writes the NaN image by itself).
#ifndef FS_CLAMP_OUTSIDE
#define FS_CLAMP_OUTSIDE 1
#endif
struct FilterPoint { float ux, uy, pv_hi, cj, cr; bool inside, nan; int cell; };
__device__ __forceinline__ FilterPoint filter_point(float x, float y, float z, const StreamViewPacked& W, float offx, float offy) {
    FilterPoint f;
    f.ux = __fmaf_rn(__fmaf_rn(W.rx.z, z, __fmaf_rn(W.rx.y, y, __fmul_rn(W.rx.x, x))), FC_CELLS_PER_NDC, offx);
    f.uy = __fmaf_rn(__fmaf_rn(W.ry.z, z, __fmaf_rn(W.ry.y, y, __fmul_rn(W.ry.x, x))), FC_CELLS_PER_NDC, offy);
    const float zr = __fmaf_rn(W.rz.z, z, __fmaf_rn(W.rz.y, y, __fmul_rn(W.rz.x, x)));
    f.pv_hi = __fmaf_rn(-(zr - W.rx.w), W.ry.w, 1.0f) + FS_PV_ABS;
    const float fj = floorf(f.ux), fi = floorf(f.uy);
    f.cj = fminf(fmaxf(fj, 0.0f), (float)(FC_G - 1)); f.cr = fminf(fmaxf(fi, 0.0f), (float)(FC_G - 1));
    f.inside = (f.cj == fj) && (f.cr == fi);             // really inside its cell (not clamped, not NaN)
    f.nan = !(f.ux == f.ux) || !(f.uy == f.uy);
    f.cell = (int)f.cr * FC_G + (int)f.cj;
#if FS_CLAMP_OUTSIDE
    if (!f.inside && !f.nan) {                           // the projection onto the grid, inside the (closed) border cell
        f.ux = fminf(fmaxf(f.ux, f.cj), f.cj + 1.0f); f.uy = fminf(fmaxf(f.uy, f.cr), f.cr + 1.0f);
        f.inside = true;
    }
#endif
    return f;
}

// The per-point test of S3 (inside points only): may point f reach U at some pixel?  Pixels of the 5 x 5 neighbourhood of
// its cell exactly, all others through the interpolated corner values of T_far (lower bound, see S2).  One 128-byte line.
// Returns 2: it does at a pixel of the window; 1: only the far bound lets it through; 0: it cannot.  thr = the largest
// |x - q|^2 - U(q) that still passes.
__device__ __forceinline__ int point_may_win(const FilterPoint& f, const float4* __restrict__ win, float inv_k, float* thr_out = nullptr)
{
    const float h = 1.0f / FC_CELLS_PER_NDC;
    const float c_lo = -logf(f.pv_hi) * inv_k * (1.0f - 1e-5f);          // pv rounded up -> c rounded down
    const float thr = stream_slack(inv_k) - c_lo;
    if (thr_out) *thr_out = thr;
    const float4* line = win + (size_t)f.cell * 8;
    float w[32];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float4 v = __ldg(line + k);
        w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
    }
    const float tx = f.ux - f.cj, ty = f.uy - f.cr;                      // position in the cell, [0, 1)
    float best = INFINITY;                                               // min over the window of |x - q|^2 - U(q)
#pragma unroll
    for (int di = -2; di <= 2; ++di) {
        const float dy = (ty - (float)di - 0.5f) * h;
#pragma unroll
        for (int dj = -2; dj <= 2; ++dj) {
            const float dx = (tx - (float)dj - 0.5f) * h;
            best = fminf(best, __fmaf_rn(dx, dx, dy * dy) - w[(di + 2) * 5 + dj + 2]);
        }
    }
    const float top = __fmaf_rn(tx, w[26] - w[25], w[25]), bot = __fmaf_rn(tx, w[28] - w[27], w[27]);
    const float tlb = __fmaf_rn(ty, bot - top, top) - h * h * (tx * (1.0f - tx) + ty * (1.0f - ty));
    return best <= thr ? 2 : (tlb <= thr ? 1 : 0);
}

__global__ void __launch_bounds__(FS_FILTER_THREADS, 1)
stream_filter_kernel(StreamParams P)
{
    extern __shared__ __align__(16) unsigned s_q[];      // step-2 queues, pair queues, point rings
    unsigned short* s_q1 = reinterpret_cast<unsigned short*>(s_q + (FS_FILTER_THREADS / 32) * FS_WQ);
    float* s_ring = reinterpret_cast<float*>(s_q1 + (FS_FILTER_THREADS / 32) * FS_Q1);
    // the cell test reads a random view per lane: 8 copies of the constants, copy (lane & 7) at 16-byte unit 8 * view + copy,
    // so the 8 lanes of a quarter warp always hit 8 different bank groups
    float4* s_vrep = reinterpret_cast<float4*>(s_ring + (FS_FILTER_THREADS / 32) * 2 * 96);       // [3][64][8]
    __shared__ StreamViewPacked s_view[FS_VPC_FILTER];
    __shared__ float s_pose[12];
    __shared__ float s_grid[6];                          // voxel grid of this cloud: lo x,y,z, voxels per unit length x,y,z (S2b)
    __shared__ bool s_grid_ok;
    const int v0 = blockIdx.x * FS_VPC_FILTER, nv = min(FS_VPC_FILTER, P.V - v0), b = blockIdx.z;
    const int nchunk = gridDim.y >> 1, c = (int)blockIdx.y >= nchunk;
    const ItemSpan sp = item_span(P.offs, b);
    const int64_t cloud0 = c ? sp.tgt0 : sp.src0, cloud1 = c ? sp.end : sp.tgt0;
    const int64_t base = cloud0 + (int64_t)((int)blockIdx.y - c * nchunk) * P.filter_chunk;
    if (base >= cloud1) return;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < nv) {
        const int v = v0 + tid;
        const float* R = P.views + (size_t)v * 9;
        const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
        s_view[tid].rx = make_float4(R[0], R[1], R[2], zmin);
        s_view[tid].ry = make_float4(R[3], R[4], R[5], 1.0f / z_range(zmin, zmax));
        s_view[tid].rz = make_float4(R[6], R[7], R[8], 0.0f);
    }
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && tid >= 64 && tid < 76) s_pose[tid - 64] = P.pose[(size_t)b * 12 + tid - 64];
    if (tid == 96) {
        const VoxelGrid G = voxel_grid(P.bbox + ((size_t)b * 2 + c) * 6, P.vox_g, nullptr);
        for (int a = 0; a < 3; ++a) { s_grid[a] = G.lo[a]; s_grid[3 + a] = G.inv[a]; }
        s_grid_ok = G.ok;
    }
    __syncthreads();
    for (int t = tid; t < 3 * FS_VPC_FILTER * 8; t += FS_FILTER_THREADS) {
        const int k = t / (FS_VPC_FILTER * 8), vv = (t >> 3) % FS_VPC_FILTER;
        s_vrep[t] = vv < nv ? (k == 0 ? s_view[vv].rx : k == 1 ? s_view[vv].ry : s_view[vv].rz) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    const size_t img0 = ((size_t)b * P.V + v0) * 2;
    const float* lpv0 = P.lpv9 + (img0 + c) * (size_t)(FS_SUB * FC_G * FS_SUB * FC_G);       // view vv of this cloud: + vv * 2 * that
    const int vg = P.vox_g, groups = (P.V + FS_VPC_FILTER - 1) / FS_VPC_FILTER;
    const unsigned long long* vox = P.vox + (((size_t)b * groups + blockIdx.x) * 2 + c) * (size_t)(vg * vg * vg);
    const float inv_k = -1.0f / P.negk;
    const float offx = P.offx, offy = P.offy;
    const int npts = (int)min((int64_t)P.filter_chunk, cloud1 - base);
    const bool grid_ok = s_grid_ok;
    const float glx = s_grid[0], gly = s_grid[1], glz = s_grid[2], gix = s_grid[3], giy = s_grid[4], giz = s_grid[5];
    const unsigned long long vmask = nv == 64 ? ~0ull : (1ull << nv) - 1ull;       // views of this group

    // step 2 (per warp): the queued (view, point) pairs of this warp, all lanes busy
    unsigned* wq = s_q + (tid >> 5) * FS_WQ;
    int wcount = 0;                                      // warp-uniform
    unsigned long long n_cell = 0ull;                    // diagnostics
    auto drain = [&](bool flush) {                       // the newest 32 entries at a time; a partial batch only when flushing
        while (wcount >= 32 || (flush && wcount > 0)) {
            const int t0 = max(wcount - 32, 0), t = t0 + lane;
            n_cell += (unsigned)(wcount - t0);
            bool keep = false; int l = 0, li = 0;
            if (t < wcount) {
                const unsigned e = wq[t];
                l = (int)(e >> 20);
                const int pi = (int)(e & 0xFFFFFu);
                const float* q = P.xyz + (size_t)(base + pi) * 3;
                float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
                if (posed) apply_pose(s_pose, x, y, z);
                const FilterPoint f = filter_point(x, y, z, s_view[l >> 1], P.offx, P.offy);
                li = (int)(base + pi - cloud0);
                keep = !f.inside || point_may_win(f, P.win + (img0 + l) * (size_t)FC_CELLS * 8, inv_k) != 0;
            }
            wcount = t0;
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
    };

    const unsigned lt = (1u << lane) - 1u;
    unsigned short* q1 = s_q1 + (tid >> 5) * FS_Q1;      // ring of (slot << 11 | view << 5 | lane), FIFO
    float* ring = s_ring + (tid >> 5) * (2 * 96);        // the posed points of this round and the one before: [slot][x,y,z][lane]
    unsigned q1h = 0u, q1t = 0u;                         // warp-uniform head / tail (free running)
    int cur_p0 = 0, cur_slot = 0;                        // the round being appended: chunk index of lane 0's point, ring slot
    unsigned tail1 = 0u;                                 // queue tail when the last round started appending: older pairs live in the slot reused next
    // step 1 on the oldest 32 pairs of the queue (count < 32 only when flushing)
    auto cell_test = [&](int count) {
        const bool live = lane < count;
        const unsigned e = live ? (unsigned)q1[(q1h + lane) & (FS_Q1 - 1)] : 0u;
        q1h += (unsigned)count;
        const int slot = (int)(e >> 11), vv = (int)(e >> 5) & 63, sl = (int)(e & 31u);
        const float* pt = ring + slot * 96 + sl;
        const float px = pt[0], py = pt[32], pz = pt[64];
        const float4 rx = s_vrep[vv * 8 + (lane & 7)], ry = s_vrep[(FS_VPC_FILTER + vv) * 8 + (lane & 7)], rz = s_vrep[(2 * FS_VPC_FILTER + vv) * 8 + (lane & 7)];
        // the transform of filter_point() / stream_point(); sub-cell by integer floor of 3 x the cell coordinate
        const float ux = __fmaf_rn(__fmaf_rn(rx.z, pz, __fmaf_rn(rx.y, py, __fmul_rn(rx.x, px))), FC_CELLS_PER_NDC, offx);
        const float uy = __fmaf_rn(__fmaf_rn(ry.z, pz, __fmaf_rn(ry.y, py, __fmul_rn(ry.x, px))), FC_CELLS_PER_NDC, offy);
        const float zr = __fmaf_rn(rz.z, pz, __fmaf_rn(rz.y, py, __fmul_rn(rz.x, px)));
        const float pv_hi = __fmaf_rn(-(zr - rx.w), ry.w, 1.0f) + FS_PV_ABS;
        const int j3 = __float2int_rd(ux * (float)FS_SUB), i3 = __float2int_rd(uy * (float)FS_SUB);
#if FS_CLAMP_OUTSIDE
        // outside points: the sub-cell of their projection onto the grid (see FilterPoint); NaN coordinates pass
        const bool inside = (ux == ux) && (uy == uy);
        const int j3c = min(max(j3, 0), FS_SUB * FC_G - 1), i3c = min(max(i3, 0), FS_SUB * FC_G - 1);
        const float lim = __ldg(lpv0 + (size_t)vv * (2 * FS_SUB * FC_G * FS_SUB * FC_G) + (inside ? i3c * (FS_SUB * FC_G) + j3c : 0));
#else
        const bool inside = (unsigned)j3 < (unsigned)(FS_SUB * FC_G) && (unsigned)i3 < (unsigned)(FS_SUB * FC_G);
        const float lim = __ldg(lpv0 + (size_t)vv * (2 * FS_SUB * FC_G * FS_SUB * FC_G) + (inside ? i3 * (FS_SUB * FC_G) + j3 : 0));
#endif
        const bool pass = live && pv_hi > 0.0f && (pv_hi >= lim || !inside);         // NaN (and, without the projection, clamped) points are never dropped here
        const unsigned pm = __ballot_sync(0xFFFFFFFFu, pass);
        if (pm) {
            // the pair's point: entries are at most one round old, the other slot holds the round before
            if (pass) wq[wcount + __popc(pm & lt)] = ((unsigned)vv << 21) | ((unsigned)c << 20) | (unsigned)(cur_p0 - (slot != cur_slot ? FS_FILTER_THREADS : 0) + sl);
            wcount += __popc(pm);
        }
        if (wcount >= 32) { __syncwarp(); drain(false); }
    };
    float nx = 0.0f, ny = 0.0f, nz = 0.0f;               // the next round's point, loaded one round ahead
    if (tid < npts) { const float* q = P.xyz + (size_t)(base + tid) * 3; nx = __ldg(q); ny = __ldg(q + 1); nz = __ldg(q + 2); }
    for (int p0 = 0, slot = 0; p0 < npts; p0 += FS_FILTER_THREADS, slot ^= 1) {
        // the slot written below still holds the round before last: pairs of that round must be gone first (rare)
        if ((int)(tail1 - q1h) > 0) { cell_test((int)min(32u, q1t - q1h)); __syncwarp(); }
        tail1 = q1t;
        cur_p0 = p0 + (tid & ~31); cur_slot = slot;
        // step 0: this lane's point, its voxel mask, the pairs whose bit is set
        const int pi = p0 + tid;
        float x = nx, y = ny, z = nz;
        if (pi + FS_FILTER_THREADS < npts) {
            const float* q = P.xyz + (size_t)(base + pi + FS_FILTER_THREADS) * 3;
            nx = __ldg(q); ny = __ldg(q + 1); nz = __ldg(q + 2);
        }
        unsigned long long m = 0ull;
        if (pi < npts) {
            if (posed) apply_pose(s_pose, x, y, z);
            const int i0 = min((int)((x - glx) * gix), vg - 1), i1 = min((int)((y - gly) * giy), vg - 1),
                      i2 = min((int)((z - glz) * giz), vg - 1);
            m = (grid_ok ? __ldg(vox + (size_t)((i2 * vg + i1) * vg + i0)) : ~0ull) & vmask;
        }
        ring[slot * 96 + lane] = x; ring[slot * 96 + 32 + lane] = y; ring[slot * 96 + 64 + lane] = z;
        const unsigned etag = ((unsigned)slot << 11) | (unsigned)lane;
#pragma unroll
        for (int half = 0; half < 2; ++half) {           // 32 views at a time: at most 1024 new pairs, < 32 left from before
            unsigned mh = half ? (unsigned)(m >> 32) : (unsigned)m;
            if (__any_sync(0xFFFFFFFFu, mh != 0u)) {
                // every lane appends its own pairs: exclusive prefix of the bit counts, then one store per set bit
                const int cnt = __popc(mh);
                int inc = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane >= d) inc += o; }
                unsigned pos = q1t + (unsigned)(inc - cnt);
                q1t += (unsigned)__shfl_sync(0xFFFFFFFFu, inc, 31);
                // The bit counts are skewed (a rim point passes twenty views, most points none): a lane with many bits has the
                // whole warp write its pairs in one step (lane v writes view v); the others loop over their few bits.
                const unsigned heavy = __ballot_sync(0xFFFFFFFFu, cnt >= FS_HEAVY);
                for (unsigned hm = heavy; hm; hm &= hm - 1u) {
                    const int L = __ffs(hm) - 1;
                    const unsigned mL = __shfl_sync(0xFFFFFFFFu, mh, L), pL = __shfl_sync(0xFFFFFFFFu, pos, L);
                    if ((mL >> lane) & 1u)
                        q1[(pL + (unsigned)__popc(mL & lt)) & (FS_Q1 - 1)] = (unsigned short)(((unsigned)slot << 11) | ((unsigned)(lane + 32 * half) << 5) | (unsigned)L);
                }
                if (cnt >= FS_HEAVY) mh = 0u;
                while (mh) {
                    const int vv = __ffs(mh) - 1 + 32 * half;
                    mh &= mh - 1u;
                    q1[pos++ & (FS_Q1 - 1)] = (unsigned short)(etag | ((unsigned)vv << 5));
                }
                __syncwarp();
                while (q1t - q1h >= 32u) cell_test(32);
            }
        }
    }
    __syncwarp();
    while (q1t != q1h) cell_test((int)min(32u, q1t - q1h));
    __syncwarp();
    drain(true);
    if (lane == 0) {
        if (n_cell) atomicAdd(P.stats, n_cell);
        if (q1t) atomicAdd(P.stats + 2, (unsigned long long)q1t);
    }
}

// ---- second round: re-seed from the survivors, filter the survivor lists again ------------------------------------
// The first round's seeds are the best of a SAMPLE, so its survivors are the points that beat the best sample somewhere:
// about (points per cell) / (samples per cell) per cell and neighbour.  They contain every true winner, hence seeding
// again from them gives the true per-cell best, and the same exact test then leaves little more than the winners.
// Both kernels walk one image's list: grid (chunks of FS_R2_CHUNK list entries, V * 2, B).
constexpr int FS_R2_CHUNK = 32768;
constexpr size_t FS_RESEED_SMEM = (size_t)2 * FC_CELLS * sizeof(unsigned long long);
constexpr unsigned FS_R2_OVERFLOW = 0xFFFFFFFFu;         // surv2_cnt of an image whose first list overflowed

struct Round2Ctx { size_t img; int64_t lo; unsigned n1; bool posed; };
__device__ __forceinline__ Round2Ctx round2_ctx(const StreamParams& P) {
    Round2Ctx C;
    const int vc = blockIdx.y, v = vc >> 1, c = vc & 1, b = blockIdx.z;
    C.img = ((size_t)b * P.V + v) * 2 + c;
    const ItemSpan sp = item_span(P.offs, b);
    C.lo = c ? sp.tgt0 : sp.src0;
    C.n1 = P.surv_cnt[C.img];
    C.posed = P.pose != nullptr && c == 0;
    return C;
}

__global__ void __launch_bounds__(FS_THREADS)
stream_reseed_kernel(StreamParams P)
{
    extern __shared__ unsigned long long s_tile[];       // [2][4096]: best at the cell's own pixel / best at the cell's aim
    unsigned long long* s_tile2 = s_tile + FC_CELLS;
    __shared__ StreamView s_view;
    __shared__ float s_pose[12];
    const Round2Ctx C = round2_ctx(P);
    if (C.n1 > (unsigned)P.surv_cap || (unsigned)blockIdx.x * FS_R2_CHUNK >= C.n1) return;
    const int tid = threadIdx.x, b = blockIdx.z, v = blockIdx.y >> 1;
    stream_load_views(P, b, v, 1, &s_view);
    if (C.posed && tid < 12) s_pose[tid] = P.pose[(size_t)b * 12 + tid];
    for (int t = tid; t < 2 * FC_CELLS; t += FS_THREADS) s_tile[t] = 0ull;
    __syncthreads();
    const float inv_k = -1.0f / P.negk;
    const float2* aim = P.aim + C.img * FC_CELLS;
    const float negkc = P.negk / (FC_CELLS_PER_NDC * FC_CELLS_PER_NDC);
    const int* list = P.surv + C.img * (size_t)P.surv_cap;
    const unsigned stop = min(C.n1, (blockIdx.x + 1u) * FS_R2_CHUNK);
    for (unsigned t = blockIdx.x * FS_R2_CHUNK + tid; t < stop; t += FS_THREADS) {
        const unsigned li = (unsigned)__ldg(list + t);
        const float* q = P.xyz + (size_t)(C.lo + li) * 3;
        float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
        if (C.posed) apply_pose(s_pose, x, y, z);
        const StreamPoint pt = stream_point(x, y, z, s_view, P.offx, P.offy);
        const float pv_lo = pt.pv - FS_PV_ABS;
        if (!(pv_lo > 0.0f) || !pt.inside) continue;
        const float val = pv_lo * home_weight(pt, negkc);
        const unsigned bits = __float_as_uint(val);
        unsigned long long* slot = &s_tile[pt.cell];
        if (val > 0.0f && bits > (unsigned)(*reinterpret_cast<volatile unsigned long long*>(slot) >> 32))
            atomicMax(slot, ((unsigned long long)bits << 32) | li);
        // rim cells: the smallest score at the cell's aim (scores are positive floats: their bits order like uints)
        const float2 a = __ldg(aim + pt.cell);
        if (a.x == a.x) {
            const float dx = rot_row(s_view.R, x, y, z) + P.shx - a.x, dy = rot_row(s_view.R + 3, x, y, z) + P.shy - a.y;
            const float sc = __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, -logf(pv_lo) * inv_k));
            const unsigned nb = ~__float_as_uint(fmaxf(sc, 0.0f));
            unsigned long long* slot2 = &s_tile2[pt.cell];
            if (nb > (unsigned)(*reinterpret_cast<volatile unsigned long long*>(slot2) >> 32))
                atomicMax(slot2, ((unsigned long long)nb << 32) | li);
        }
    }
    __syncthreads();
    unsigned long long* out = P.seed + C.img * FC_CELLS;
    unsigned long long* out2 = P.seed2 + C.img * FC_CELLS;
    for (int t = tid; t < FC_CELLS; t += FS_THREADS) {
        const unsigned long long key = s_tile[t], key2 = s_tile2[t];
        if (key != 0ull && key > out[t]) atomicMax(&out[t], key);
        if (key2 != 0ull && key2 > out2[t]) atomicMax(&out2[t], key2);
    }
}

// The second filter also settles the points that only the far bound let through (the rim of a solid cloud: the
// interpolated bound is loose by up to ~h^2/2 there, and the cloud is dense): the whole warp tests one such point against
// every pixel -- 8 x 8 pixel blocks first (distance to the block against the block's largest U), then the pixels of the
// blocks that remain.  What is left is little more than the points that win a pixel.
__global__ void __launch_bounds__(FS_THREADS)
stream_refilter_kernel(StreamParams P)
{
    __shared__ StreamViewPacked s_view;
    __shared__ float s_pose[12];
    __shared__ float s_U[FC_CELLS], s_Ub[64];
    const Round2Ctx C = round2_ctx(P);
    if (C.n1 > (unsigned)P.surv_cap) {                    // first list incomplete: no second list either
        if (blockIdx.x == 0 && threadIdx.x == 0) P.surv2_cnt[C.img] = FS_R2_OVERFLOW;
        return;
    }
    if ((unsigned)blockIdx.x * FS_R2_CHUNK >= C.n1) return;
    const int tid = threadIdx.x, lane = tid & 31, b = blockIdx.z, v = blockIdx.y >> 1;
    if (tid == 0) {
        const float* R = P.views + (size_t)v * 9;
        const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
        s_view.rx = make_float4(R[0], R[1], R[2], zmin);
        s_view.ry = make_float4(R[3], R[4], R[5], 1.0f / z_range(zmin, zmax));
        s_view.rz = make_float4(R[6], R[7], R[8], 0.0f);
    }
    if (C.posed && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];
    for (int q = tid; q < FC_CELLS; q += FS_THREADS) s_U[q] = __ldg(P.uimg + C.img * FC_CELLS + q);
    __syncthreads();
    if (tid < 64) {
        const int bi = tid >> 3, bj = tid & 7;
        float m = -INFINITY;
        for (int k = 0; k < 64; ++k) m = fmaxf(m, s_U[(bi * 8 + (k >> 3)) * FC_G + bj * 8 + (k & 7)]);
        s_Ub[tid] = m;
    }
    __syncthreads();
    const float inv_k = -1.0f / P.negk;
    const float h = 1.0f / FC_CELLS_PER_NDC;
    const int* list = P.surv + C.img * (size_t)P.surv_cap;
    const unsigned first = blockIdx.x * FS_R2_CHUNK, stop = min(C.n1, first + FS_R2_CHUNK);
    for (unsigned t0 = first + (tid - lane); t0 < stop; t0 += FS_THREADS) {
        const unsigned t = t0 + lane;
        int code = 0, li = 0; float ux = 0.0f, uy = 0.0f, thr = 0.0f;
        if (t < stop) {
            li = __ldg(list + t);
            const float* q = P.xyz + (size_t)(C.lo + li) * 3;
            float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
            if (C.posed) apply_pose(s_pose, x, y, z);
            const FilterPoint f = filter_point(x, y, z, s_view, P.offx, P.offy);
            ux = f.ux; uy = f.uy;
            if (f.pv_hi > 0.0f) {
                if (!f.inside) code = 2;
                else {
                    // the sub-cell threshold of THIS round's bounds first (4 bytes): most of the list fails it
                    const int j3 = __float2int_rd(f.ux * (float)FS_SUB), i3 = __float2int_rd(f.uy * (float)FS_SUB);
#if FS_CLAMP_OUTSIDE
                    const bool in3 = true;               // a projected point may sit on the grid's far edge: the last sub-cell
                    const int j3c = min(max(j3, 0), FS_SUB * FC_G - 1), i3c = min(max(i3, 0), FS_SUB * FC_G - 1);
                    const float lim = __ldg(P.lpv9 + C.img * (size_t)(FS_SUB * FC_G * FS_SUB * FC_G) + i3c * (FS_SUB * FC_G) + j3c);
#else
                    const bool in3 = (unsigned)j3 < (unsigned)(FS_SUB * FC_G) && (unsigned)i3 < (unsigned)(FS_SUB * FC_G);
                    const float lim = in3 ? __ldg(P.lpv9 + C.img * (size_t)(FS_SUB * FC_G * FS_SUB * FC_G) + i3 * (FS_SUB * FC_G) + j3) : 0.0f;
#endif
                    (void)in3;
                    if (f.pv_hi >= lim) code = point_may_win(f, P.win + C.img * (size_t)FC_CELLS * 8, inv_k, &thr);
                }
            }
        }
        // the far candidates, one at a time, all lanes on it
        unsigned todo = __ballot_sync(0xFFFFFFFFu, code == 1);
        while (todo) {
            const int sl = __ffs(todo) - 1;
            todo &= todo - 1u;
            const float px = __shfl_sync(0xFFFFFFFFu, ux, sl), py = __shfl_sync(0xFFFFFFFFu, uy, sl), pt = __shfl_sync(0xFFFFFFFFu, thr, sl);
            unsigned blocks[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) {                // lane tests blocks lane and lane + 32: pixel centres [8 bj + 0.5, 8 bj + 7.5]
                const int blk = lane + 32 * r, bi = blk >> 3, bj = blk & 7;
                const float dx = fmaxf(fmaxf((float)(bj * 8) + 0.5f - px, px - ((float)(bj * 8) + 7.5f)), 0.0f) * h;
                const float dy = fmaxf(fmaxf((float)(bi * 8) + 0.5f - py, py - ((float)(bi * 8) + 7.5f)), 0.0f) * h;
                blocks[r] = __ballot_sync(0xFFFFFFFFu, __fmaf_rn(dx, dx, dy * dy) * (1.0f - 1e-5f) - s_Ub[blk] <= pt);
            }
            bool hit = false;
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                unsigned bm = blocks[r];
                while (bm && !hit) {
                    const int blk = __ffs(bm) - 1 + 32 * r, bi = blk >> 3, bj = blk & 7;
                    bm &= bm - 1u;
                    bool any = false;
#pragma unroll
                    for (int rr = 0; rr < 2; ++rr) {     // lane tests pixels lane and lane + 32 of the block
                        const int k = lane + 32 * rr, i = bi * 8 + (k >> 3), j = bj * 8 + (k & 7);
                        const float dx = (px - ((float)j + 0.5f)) * h, dy = (py - ((float)i + 0.5f)) * h;
                        // the extra rounding of px - (j + 0.5) (<= 4e-6 cells on >= 2.5 cells) goes to the safe side
                        any = any || (__fmaf_rn(dx, dx, dy * dy) * (1.0f - 8e-6f) - s_U[i * FC_G + j] <= pt);
                    }
                    hit = __any_sync(0xFFFFFFFFu, any);
                }
            }
            if (hit && lane == sl) code = 2;
        }
        const bool keep = code == 2;
        const unsigned km = __ballot_sync(0xFFFFFFFFu, keep);
        if (km) {
            unsigned gp = 0;
            if (lane == 0) gp = atomicAdd(&P.surv2_cnt[C.img], (unsigned)__popc(km));
            gp = __shfl_sync(0xFFFFFFFFu, gp, 0) + __popc(km & ((1u << lane) - 1u));
            if (keep && gp < (unsigned)P.surv2_cap) P.surv2[C.img * (size_t)P.surv2_cap + gp] = li;
        }
    }
}

}  // namespace scream
