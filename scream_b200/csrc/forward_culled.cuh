// forward_culled.cuh -- culled exact forward: the product kernel.
//
// Same result, bit for bit, as forward_dense.cuh (and so as render.py:39-49): for every pixel the max
// over the cloud of  val = pv_i * exp(-d2/2*rho^2)  with the lowest index on ties -- but almost no
// (point, pixel) pair is ever evaluated.  One CTA renders one (item, view, cloud) image.
//
// Search domain.  With k = rho^2/2 and c_i = -ln(pv_i)/k >= 0 the value is exp(-k * s), s = d2 + c_i,
// so the winner of a pixel is the point of minimum "score" s (a power diagram).  The search runs on the
// score (two FMAs per pair, no exp); exp() is evaluated once per pixel, for the point that won, in the
// reference's op order.  fp32 rounding can reorder two points only when their scores agree to ~1e-6
// relative: every pixel therefore tracks the runner-up score as well, and a pixel whose two best scores
// are within a (5x over-estimated) margin -- or whose value is in the denormal range, where fp32 has no
// relative precision -- is re-done by an exact pass that evaluates the reference expression for every
// point inside the margin and applies the (value, lowest index) order directly.
//
//   P1  transform (pose, view rotation) every point, bin it into a 64 x 64 grid of cells over
//       [-1,1]^2 (clamped), shared-memory histogram; the atomic's return value is the rank in the cell.
//   P2  exclusive scan of the histogram -> cell start offsets (row-major: a run of cells in one grid
//       row is one contiguous run of records).
//   P3  counting-sort scatter of (x', y', c, index) records into shared memory (SoA).
//   P4  per 64 x 64 pixel tile:
//       A  lane per pixel, 3 x 3 cells around the pixel: best / runner-up score.  Resolved when the best
//          score (+margin) is smaller than the squared distance to the edge of the window (a point
//          outside the window has s >= d2 > best).  ~90 % of the pixels of a dense cloud end here.
//       C  the rest (compacted into a queue): disc scan -- grid rows outwards from the pixel, per-row
//          column span from the circle equation, radius^2 = best score so far (+margin), capped where
//          exp() underflows to exactly zero -- then, if flagged, the exact pass over the same disc.
//   Clouds larger than the shared-memory capacity are rendered in batches; the per-pixel best is carried
//   through the output buffers (max is associative; earlier batches hold the lower indices).
#pragma once
#include "common.cuh"
#include "forward_dense.cuh"   // FwdParams

namespace scream {

constexpr int FC_THREADS = 1024;
constexpr int FC_PPT = 12;                               // points per thread per batch
constexpr int FC_CAP = FC_THREADS * FC_PPT;              // 12288 points per batch
#ifndef FC_NEAR_SPREAD_N
#define FC_NEAR_SPREAD_N 0
#endif
constexpr int FC_NEAR_SPREAD = FC_NEAR_SPREAD_N;                        // near queue spread over all warps also when the GPU is full (A/B)
#ifndef FC_FAR_TAKE_N
#define FC_FAR_TAKE_N 32
#endif
constexpr int FC_FAR_TAKE = FC_FAR_TAKE_N;                 // far pixels a warp takes at once when every image has a CTA of its own
#ifndef FC_COOP_FAR_N
#define FC_COOP_FAR_N 4
#endif
constexpr int FC_COOP_FAR = FC_COOP_FAR_N;                           // small grids: far pixels a warp searches together before their values are evaluated side by side
#ifndef FC_AHEAD_N
#define FC_AHEAD_N 2
#endif
#ifndef FC_ZFLIGHT_N
#define FC_ZFLIGHT_N 4
#endif
constexpr int FC_ZFLIGHT = FC_ZFLIGHT_N;                   // points in flight per thread in the fused depth-range pass
constexpr int FC_AHEAD = FC_AHEAD_N;                              // points whose global loads are in flight ahead of the one in work (P1, P3)
constexpr int FC_DIL_MAX = 31;                           // dilation steps whose masks fit the histogram's memory (31 x 64 x 8 bytes)
#ifndef FC_LINEAR_MAX_N
#define FC_LINEAR_MAX_N 128
#endif
constexpr int FC_LINEAR_MAX = FC_LINEAR_MAX_N;             // batches of up to this many points: every pixel scans all records
constexpr int FC_G = 64;                                 // cell grid side
constexpr int FC_CELLS = FC_G * FC_G;
constexpr int FC_TILE = 64;                              // pixel tile side of P4
constexpr int FC_TPIX = FC_TILE * FC_TILE;
constexpr float FC_CELLS_PER_NDC = (float)(FC_G / 2);
constexpr float FC_EXP_ZERO = 108.0f;                    // k*s beyond this: value is exactly +0
constexpr float FC_EXP_DENORM = 80.0f;                   // k*s beyond this: treat as denormal zone
constexpr float FC_EXP_MINUS1 = 18.1f;                   // k*s beyond this: (val-0.5)/0.5 == -1 exactly

// shared-memory carve-up (bytes)
constexpr size_t FC_OFF_REC = 0;                                             // record PAIRS (see FcShared): xy2 | c2 | idx
constexpr size_t FC_OFF_C2 = FC_OFF_REC + (size_t)FC_CAP * 8;
constexpr size_t FC_OFF_IDX = FC_OFF_C2 + (size_t)FC_CAP * 4;
constexpr size_t FC_OFF_START = FC_OFF_IDX + (size_t)FC_CAP * 4;             // u16 [CELLS + 8]
constexpr size_t FC_OFF_SCRATCH = FC_OFF_START + (size_t)(FC_CELLS + 8) * 2; // u32 hist[CELLS+4] | u16 cand[TPIX], queue[TPIX]
constexpr size_t FC_SMEM = FC_OFF_SCRATCH + (size_t)(FC_CELLS + 4) * 4;

// Records live in shared memory as PAIRS: xy2[p] = {x0, x1, y0, y1} (16 B), c2[p] = {c0, c1} (8 B), idx[2p], idx[2p+1],
// for records 2p and 2p+1: one LDS.128 + one LDS.64 feed the packed fp32x2 pipe (FADD2 / FFMA2, sm_100), two scores for
// four FMA-pipe instructions.  A run of records [rb, re) is walked as the pairs [rb >> 1, (re + 1) >> 1): the extra record
// at either end is a real point of a neighbouring cell (or the +inf pad behind the last record), and a real point is
// always a valid candidate.  The search loops address the arrays through 32-bit shared-window addresses (no generic ->
// shared conversion in the loops) and load the next pair before they work on the current one.
struct FcShared {
    unsigned xy2, c2;                                    // shared-window addresses of the pair arrays
    unsigned char* rec;                                  // generic pointer to the same block (scalar accessors)
    const unsigned short* start;
    const unsigned long long* rowmask; const unsigned long long* blockmask;   // cell occupancy: per grid row, per 4 rows
};
__device__ __forceinline__ float* rec_x(unsigned char* rec, int s) { return reinterpret_cast<float*>(rec) + ((s >> 1) << 2) + (s & 1); }
__device__ __forceinline__ float* rec_c(unsigned char* rec, int s) { return reinterpret_cast<float*>(rec + FC_OFF_C2 - FC_OFF_REC) + s; }
__device__ __forceinline__ int* rec_idx(unsigned char* rec, int s) { return reinterpret_cast<int*>(rec + FC_OFF_IDX - FC_OFF_REC) + s; }
__device__ __forceinline__ float4 lds128(unsigned a) {
    float4 v; asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a)); return v;
}
__device__ __forceinline__ float2 lds64(unsigned a) {
    float2 v; asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a)); return v;
}

__device__ __forceinline__ float2 sub2(float2 a, float b) {
    float2 d;
    asm("{ .reg .b64 ra, rb, rd;\n  mov.b64 ra, {%2, %3};\n  mov.b64 rb, {%4, %4};\n  sub.rn.f32x2 rd, ra, rb;\n  mov.b64 {%0, %1}, rd; }"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b));
    return d;
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{ .reg .b64 ra, rb, rc, rd;\n  mov.b64 ra, {%2, %3};\n  mov.b64 rb, {%4, %5};\n  mov.b64 rc, {%6, %7};\n"
        "  fma.rn.f32x2 rd, ra, rb, rc;\n  mov.b64 {%0, %1}, rd; }"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
__device__ __forceinline__ float fmin3f(float a, float b, float c) {     // NaN operands are ignored, like fminf
    float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;
}

__device__ __forceinline__ int cell_of(float x) {
    // fl(fl(x + 1) * G/2), clamped; NaN -> 0
    float u = __fmul_rn(__fadd_rn(x, 1.0f), FC_CELLS_PER_NDC);
    u = fminf(fmaxf(u, 0.0f), (float)(FC_G - 1));
    return (int)u;                                       // u >= 0: truncation == floor
}

// radius in cells of a squared NDC radius (inflated: covers the rounding of cell_of and of the sqrt)
__device__ __forceinline__ float radius_cells(float r2) {
    float r = __fmul_rn(sqrtf(r2), FC_CELLS_PER_NDC);
    return __fmaf_rn(r, 1e-4f, r) + 1e-3f;
}

// score margin inside which two points may swap order through fp32 rounding (5x the analysed bound)
__device__ __forceinline__ float score_margin(float s, float inv_k) {
    return __fmaf_rn(s, 4e-6f, __fmul_rn(3e-6f, inv_k));
}

struct PixelGeom { float cx, cy, ux, uy; };              // centre in NDC and in cell coordinates

// scores of records 2p and 2p+1 at pixel g: fl(dx*dx + fl(dy*dy + c)) each, the same bits wherever it is evaluated
__device__ __forceinline__ float2 pair_scores(float4 q, float2 c, const PixelGeom& g) {
    const float2 dx = sub2(make_float2(q.x, q.y), g.cx), dy = sub2(make_float2(q.z, q.w), g.cy);
    return fma2(dx, dx, fma2(dy, dy, c));
}
__device__ __forceinline__ float2 score_pair(const FcShared& S, int p, const PixelGeom& g) {
    return pair_scores(lds128(S.xy2 + 16u * (unsigned)p), lds64(S.c2 + 8u * (unsigned)p), g);
}
__device__ __forceinline__ float score_of(const FcShared& S, int s, const PixelGeom& g) {
    const float2 sc = score_pair(S, s >> 1, g);
    return (s & 1) ? sc.y : sc.x;
}
__device__ __forceinline__ float2 record_xy(const FcShared& S, int s) {
    const float* f = rec_x(S.rec, s);
    return make_float2(f[0], f[2]);
}
// one step of the (best, runner-up, best pair) search over a pair of scores
__device__ __forceinline__ void top2_step(float2 sc, int p, float& s1, float& s2, int& a1, bool& improved) {
    const float lo = fminf(sc.x, sc.y), hi = fmaxf(sc.x, sc.y);
    s2 = fmin3f(s2, fmaxf(lo, s1), hi);
    if (lo < s1) { s1 = lo; a1 = p; improved = true; }
}
// (best, runner-up, best pair) over the records [rb, re), software-pipelined: the loads of pair p + 1 are in flight while
// pair p is scored (the load past the end stays inside the CTA's shared memory and is never used)
__device__ __forceinline__ void scan_run(const FcShared& S, int rb, int re, const PixelGeom& g,
                                         float& s1, float& s2, int& a1, bool& improved) {
    int p = rb >> 1;
    const int pe = (re + 1) >> 1;
    if (p >= pe) return;
    unsigned axy = S.xy2 + 16u * (unsigned)p, ac = S.c2 + 8u * (unsigned)p;
    float4 q = lds128(axy); float2 c = lds64(ac);
#pragma unroll 1
    for (; p + 1 < pe; p += 2) {                          // two pairs per trip, registers ping-pong (no moves)
        const float4 qn = lds128(axy + 16u); const float2 cn = lds64(ac + 8u);
        top2_step(pair_scores(q, c, g), p, s1, s2, a1, improved);
        axy += 32u; ac += 16u;
        q = lds128(axy); c = lds64(ac);
        top2_step(pair_scores(qn, cn, g), p + 1, s1, s2, a1, improved);
    }
    if (p < pe) top2_step(pair_scores(q, c, g), p, s1, s2, a1, improved);
}
// pair index -> record index of the better of the two (ties inside the pair leave s2 == s1: the pixel is not resolved)
__device__ __forceinline__ int best_of_pair(const FcShared& S, int p, const PixelGeom& g) {
    if (p < 0) return -1;
    const float2 sc = score_pair(S, p, g);
    return 2 * p + (sc.y < sc.x ? 1 : 0);
}

// bits j0..j1 (inclusive) of a 64-column occupancy word
__device__ __forceinline__ unsigned long long col_mask(int j0, int j1) {
    return ((~0ull) >> (63 - j1)) & ((~0ull) << j0);
}

// exact value of record s at pixel g: the reference expression, pv recomputed from the point itself
struct ExactCtx {
    const float* xyz; int64_t lo; const float* R; const float* pose; bool posed;
    float zmin, range, negk; bool f16;
};
__device__ __forceinline__ float exact_value(const FcShared& S, const ExactCtx& E, int s, const PixelGeom& g, int& id) {
    id = *rec_idx(S.rec, s);
    const float* q = E.xyz + (size_t)(E.lo + id) * 3;
    float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
    if (E.posed) apply_pose(E.pose, x, y, z);
    const float pv = depth_weight_m(rot_row(E.R + 6, x, y, z), E.zmin, E.range, E.f16);
    const float2 r = record_xy(S, s);
    return pair_value(pair_d2(r.x, r.y, g.cx, g.cy), pv, E.negk);
}

// Disc scan around pixel g: 4-row blocks of the cell grid outwards from the pixel's row, an occupancy
// bit-mask test per block and per row (empty space costs a few instructions per 4 rows), then the
// records of the row's column span from the circle equation.
//   EXACT == false (score domain): on entry s1 is an upper bound of the best score (or +inf) and r2 the
//     largest squared radius that can matter; the radius shrinks to the best score (+margin) as records
//     are found.  On exit s1/a1 = best score / its record, s2 = runner-up score.
//   EXACT == true: every record with score <= r2 is evaluated with the reference expression and the
//     (value, lowest index) order is applied directly to best/arg.
template <bool EXACT>
__device__ __forceinline__ void disc_scan(const FcShared& S, const ExactCtx& E, const PixelGeom& g, float r2, float inv_k,
                                          float& s1, float& s2, int& a1, float& best, int& arg)
{
    const int ci = (int)g.uy;
    const int bi = ci >> 2;
    const float s_thr = r2;
    float rc = radius_cells(r2);
    if (!EXACT) { s1 = INFINITY; s2 = INFINITY; a1 = -1; }
    for (int db = 0; db < FC_G / 4; ++db) {
        bool live = false;
        for (int side = 0; side < 2; ++side) {
            if (side == 1 && db == 0) break;
            const int B = side ? bi + db : bi - db;
            if (B < 0 || B >= FC_G / 4) continue;
            const float bdist = db == 0 ? 0.0f : (side ? __fsub_rn((float)(4 * B), g.uy) : __fsub_rn(g.uy, (float)(4 * B + 4)));
            if (bdist > rc) continue;
            live = true;
            const int jlo = max(0, (int)floorf(g.ux - rc)), jhi = min(FC_G - 1, (int)floorf(g.ux + rc));
            if (jlo > jhi) continue;
            const unsigned long long cm = col_mask(jlo, jhi);
            if ((S.blockmask[B] & cm) == 0ull) continue;
            for (int t = 0; t < 4; ++t) {
                const int i = side ? 4 * B + t : 4 * B + 3 - t;               // nearest row of the block first
                const float dist = i == ci ? 0.0f : (i < ci ? __fsub_rn(g.uy, (float)(i + 1)) : __fsub_rn((float)i, g.uy));
                if (dist > rc) continue;
                if ((S.rowmask[i] & cm) == 0ull) continue;
                const float h2 = __fmaf_rn(rc, rc, -__fmul_rn(dist, dist));
                const float hw = sqrtf(fmaxf(h2, 0.0f)) + 1e-3f;
                const int j0 = max(0, (int)floorf(g.ux - hw)), j1 = min(FC_G - 1, (int)floorf(g.ux + hw));
                if (j0 > j1) continue;
                const int rb = S.start[i * FC_G + j0], re = S.start[i * FC_G + j1 + 1];
                if (EXACT) {
#pragma unroll 1
                    for (int p = rb >> 1; p < (re + 1) >> 1; ++p) {
                        const float2 sc = score_pair(S, p, g);
                        if (fminf(sc.x, sc.y) <= s_thr) {
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                if ((h ? sc.y : sc.x) <= s_thr) {
                                    int id;
                                    const float val = exact_value(S, E, 2 * p + h, g, id);
                                    if (val > best || (val == best && id < arg)) { best = val; arg = id; }
                                }
                            }
                        }
                    }
                } else {
                    bool improved = false;
                    int rb2 = rb;
                    if (s1 == INFINITY) {
                        // no bound yet: take the occupied cell of this row nearest to the pixel's column first,
                        // so the rest of the (possibly long) span is cut by the radius it gives
                        const unsigned long long rm = S.rowmask[i] & col_mask(j0, j1);
                        const int cj = min(max((int)g.ux, j0), j1);
                        const unsigned long long ml = rm & ((2ull << cj) - 1ull), mr = rm >> cj;
                        int jn = -1;
                        if (ml) jn = 63 - __clzll((long long)ml);
                        if (mr) { const int jr = cj + __ffsll((long long)mr) - 1; if (jn < 0 || jr - cj < cj - jn) jn = jr; }
                        if (jn >= 0) {
                            const int nb = S.start[i * FC_G + jn], ne = S.start[i * FC_G + jn + 1];
#pragma unroll 1
                            for (int p = nb >> 1; p < (ne + 1) >> 1; ++p) {
                                const float2 sc = score_pair(S, p, g);
                                s1 = fmin3f(s1, sc.x, sc.y);
                            }
                            // radius from this bound; the full span below re-visits the cell (s1/s2/a1 stay consistent)
                            const float r2n = fminf(r2, s1 + score_margin(s1, inv_k));
                            const float rcn = radius_cells(r2n);
                            const float h2n = __fmaf_rn(rcn, rcn, -__fmul_rn(dist, dist));
                            const float hwn = sqrtf(fmaxf(h2n, 0.0f)) + 1e-3f;
                            const int j0n = max(j0, (int)floorf(g.ux - hwn)), j1n = min(j1, (int)floorf(g.ux + hwn));
                            rb2 = S.start[i * FC_G + j0n];
                            const int re2 = S.start[i * FC_G + j1n + 1];
                            s1 = INFINITY;
                            scan_run(S, rb2, re2, g, s1, s2, a1, improved);
                            rb2 = re;                     // span done
                        }
                    }
                    scan_run(S, rb2, re, g, s1, s2, a1, improved);
                    if (improved) {
                        r2 = fminf(r2, s1 + score_margin(s1, inv_k));
                        rc = radius_cells(r2);
                    }
                }
            }
        }
        if (!live) break;
    }
    if (!EXACT) a1 = best_of_pair(S, a1, g);             // the search tracked the best PAIR
}

// One grid row of the disc of radius rc (cells) around pixel g, fixed radius: the records of the row's column span, the
// expressions of disc_scan.  The warp-cooperative scan (small grids, see the kernel) gives every lane rows of its own.
template <bool EXACT>
__device__ __forceinline__ void row_scan(const FcShared& S, const ExactCtx& E, const PixelGeom& g, int i, float rc, float s_thr,
                                         float& s1, float& s2, int& a1, float& best, int& arg)
{
    const int ci = (int)g.uy;
    const float dist = i == ci ? 0.0f : (i < ci ? __fsub_rn(g.uy, (float)(i + 1)) : __fsub_rn((float)i, g.uy));
    if (dist > rc) return;
    const float h2 = __fmaf_rn(rc, rc, -__fmul_rn(dist, dist));
    float hw;                                            // sqrt.approx (2 ulp) + 1e-3 cells: still an upper bound of the half width
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(hw) : "f"(fmaxf(h2, 0.0f)));
    hw += 1e-3f;
    const int j0 = max(0, (int)floorf(g.ux - hw)), j1 = min(FC_G - 1, (int)floorf(g.ux + hw));
    if (j0 > j1 || (S.rowmask[i] & col_mask(j0, j1)) == 0ull) return;
    const int rb = S.start[i * FC_G + j0], re = S.start[i * FC_G + j1 + 1];
    if (EXACT) {
#pragma unroll 1
        for (int p = rb >> 1; p < (re + 1) >> 1; ++p) {
            const float2 sc = score_pair(S, p, g);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if ((h ? sc.y : sc.x) <= s_thr) {
                    int id;
                    const float val = exact_value(S, E, 2 * p + h, g, id);
                    if (val > best || (val == best && id < arg)) { best = val; arg = id; }
                }
            }
        }
    } else {
        bool improved = false;
        scan_run(S, rb, re, g, s1, s2, a1, improved);
    }
}
// minimum over the warp of scores: non-negative floats or +inf (never NaN: the searches fold scores with fminf), whose bit
// patterns order like unsigned integers -- one REDUX instead of five shuffle steps.  A NaN operand (a record whose depth
// weight rounded below zero) is mapped to +inf first.
__device__ __forceinline__ float warp_min(float v) {
    const unsigned b = (v == v) ? __float_as_uint(v) & 0x7FFFFFFFu : 0x7F800000u;       // -0 -> +0
    return __uint_as_float(__reduce_min_sync(0xFFFFFFFFu, b));
}

// A cloud of a few points (cnt <= FC_LINEAR_MAX), whole image (forward_tiny_kernel): every pixel scores all records, then the
// resolution rules of the disc scans (scan_pixel in the kernel): runner-up beyond the rounding margin -> the winner's exact
// value; else every record within the margin evaluated exactly, (value, lowest index).  Records in point order, compact pair
// arrays: xy2 [pairs] {x0, x1, y0, y1}, c2 [pairs] {c0, c1}; record s is point s of the cloud.
constexpr int FT_THREADS = 256, FT_SLICE = 4096;         // forward_tiny_kernel: threads per CTA, pixels per CTA (w % 64 == 0)
__device__ __forceinline__ void linear_batch(const float4* xy2f, const float2* c2f, const float* xyz, int64_t lo, const float* R,
                                             const float* pose, bool posed, float zmin, float range, float negk, bool f16, int cnt, int w,
                                             float r2cap, float s_zero, float inv_k, float* imgs, int32_t* winner, bool keep_raw, int p_begin, int p_end)
{
    FcShared S; S.xy2 = (unsigned)__cvta_generic_to_shared(xy2f); S.c2 = (unsigned)__cvta_generic_to_shared(c2f);
    S.rec = nullptr; S.start = nullptr; S.rowmask = nullptr; S.blockmask = nullptr;
    const int half = w / 2;
    auto exact = [&](int s, const PixelGeom& g) {           // exact_value() on the compact arrays
        const float* q = xyz + (size_t)(lo + s) * 3;
        float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
        if (posed) apply_pose(pose, x, y, z);
        const float pv = depth_weight_m(rot_row(R + 6, x, y, z), zmin, range, f16);
        const float* f = reinterpret_cast<const float*>(xy2f) + ((s >> 1) << 2) + (s & 1);
        return pair_value(pair_d2(f[0], f[2], g.cx, g.cy), pv, negk);
    };
    for (int p = p_begin + threadIdx.x; p < p_end; p += FT_THREADS) {
        const int row = p / w, col = p - row * w;
        PixelGeom g;
        g.cx = pixel_centre(col, half); g.cy = pixel_centre(row, half);
        g.ux = __fmul_rn(__fadd_rn(g.cx, 1.0f), FC_CELLS_PER_NDC);
        g.uy = __fmul_rn(__fadd_rn(g.cy, 1.0f), FC_CELLS_PER_NDC);
        float s1 = INFINITY, s2 = INFINITY; int a1 = -1;
        bool found = false;
        scan_run(S, 0, cnt, g, s1, s2, a1, found);
        a1 = best_of_pair(S, a1, g);
        float val = 0.0f; int id = 0;
        if (a1 >= 0 && s1 <= r2cap) {
            float s_thr = s1 + score_margin(s1, inv_k);
            const float e1 = __fmul_rn(s1, -negk);
            if (e1 >= FC_EXP_DENORM) {
                const float lu = e1 - 101.2f;
                s_thr = lu >= -0.7f ? s_zero : s_thr + 1.5f * expf(lu) * inv_k;
            }
            if (s2 > s_thr) { id = a1; val = exact(a1, g); }
            else {
                const float thr = fminf(s_thr, s_zero);
#pragma unroll 1
                for (int q = 0; q < (cnt + 1) >> 1; ++q) {
                    const float2 sc = score_pair(S, q, g);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int s = 2 * q + h;
                        if (s < cnt && (h ? sc.y : sc.x) <= thr) {
                            const float ev = exact(s, g);
                            if (ev > val || (ev == val && s < id)) { val = ev; id = s; }
                        }
                    }
                }
            }
            if (val == 0.0f) id = 0;
        }
        imgs[p] = keep_raw ? val : __fmul_rn(__fsub_rn(val, 0.5f), 2.0f);
        if (winner) winner[p] = id;
    }
}

// Visiting order of a tile's 64 pixel rows: from the top and bottom edges towards the middle (the pixels farthest from a
// centred cloud, i.e. the longest disc scans, first).  lpt_row: visiting index -> row; lpt_vis: its inverse.
__device__ __forceinline__ int lpt_row(int vis) { return vis < 16 ? vis : (vis < 32 ? 79 - vis : (vis < 48 ? vis - 16 : 95 - vis)); }
__device__ __forceinline__ int lpt_vis(int row) { return row < 16 ? row : (row < 32 ? row + 16 : (row < 48 ? 95 - row : 79 - row)); }

// a pixel whose nearest occupied cell is `ring` cells away (Chebyshev) has no point nearer than (ring - 1 + edge) cells:
// beyond the radius that can change the output it is background
__device__ __forceinline__ bool blank_pixel(int ring, float edge, float r2cap_all) {
    const float dlb = __fmul_rn((float)(ring - 1) + edge, 1.0f / FC_CELLS_PER_NDC);
    return ring >= 2 && __fmul_rn(__fmul_rn(dlb, dlb), 0.99999f) > r2cap_all;
}
// ring / blank of pixel (row, col): the expressions of phase A
__device__ __forceinline__ void pixel_class(int row, int col, int half, float r2cap_all, const unsigned char* s_ring,
                                            int& ring, bool& blank) {
    const float ux = __fmul_rn(__fadd_rn(pixel_centre(col, half), 1.0f), FC_CELLS_PER_NDC);
    const float uy = __fmul_rn(__fadd_rn(pixel_centre(row, half), 1.0f), FC_CELLS_PER_NDC);
    const int ci = (int)uy, cj = (int)ux;
    const float fx = ux - (float)cj, fy = uy - (float)ci;
    const float edge = fminf(fminf(fx, 1.0f - fx), fminf(fy, 1.0f - fy));
    ring = s_ring[ci * FC_G + cj];
    blank = blank_pixel(ring, edge, r2cap_all);
}

constexpr unsigned FC_FLAG_DUAL = 32u;                   // SCREAM_F_DUAL
constexpr unsigned FC_FLAG_NOWIN1 = 128u;                // SCREAM_F_DUAL_NOWIN1
constexpr unsigned FC_FLAG_HALF = 64u;                   // SCREAM_F_HALF: fp16 depth-weight chain (common.cuh)

// Phase timeline of a CTA (diagnostic builds only: tools/diag_cfg1_timeline.py compiles with -DFC_TIMELINE): SM cycle
// counter at the phase boundaries (thread 0, behind the phase's barrier), queue lengths and per-task times.
#ifdef FC_TIMELINE
__device__ long long g_fc_marks[1024][24];
#define FC_MARK(k) do { if (threadIdx.x == 0) g_fc_marks[blockIdx.x & 1023][(k)] = clock64(); } while (0)
#define FC_COUNT(k, n) atomicAdd(reinterpret_cast<unsigned long long*>(&g_fc_marks[blockIdx.x & 1023][(k)]), (unsigned long long)(n))
#define FC_MAXOF(k, n) atomicMax(reinterpret_cast<unsigned long long*>(&g_fc_marks[blockIdx.x & 1023][(k)]), (unsigned long long)(n))
#define FC_CLOCK() clock64()
#else
#define FC_MARK(k) do { } while (0)
#define FC_COUNT(k, n) do { } while (0)
#define FC_MAXOF(k, n) do { } while (0)
#define FC_CLOCK() 0ll
#endif

// PSTEP 1: every pixel of the image (forward_culled_kernel).  PSTEP 2: only the pixels (2 r + P.prow0, 2 c + P.pcol0) of a
// w == 128 image, one of its four 64 x 64 lattices, as ONE tile (forward_culled_lattice_kernel: giant clouds, one launch per
// lattice pass of the pre-filter).  A template, so that the common kernel keeps its register allocation.
template <int PSTEP>
__global__ void __launch_bounds__(FC_THREADS, 1)
forward_culled_kernel_t(FwdParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* s_rec = smem_raw + FC_OFF_REC;
    unsigned short* s_start = reinterpret_cast<unsigned short*>(smem_raw + FC_OFF_START);
    unsigned* s_hist = reinterpret_cast<unsigned*>(smem_raw + FC_OFF_SCRATCH);
    unsigned short* s_cand = reinterpret_cast<unsigned short*>(smem_raw + FC_OFF_SCRATCH);
    unsigned short* s_queue = s_cand + FC_TPIX;
    __shared__ float s_R[9], s_pose[12];
    __shared__ unsigned s_warp_tot[FC_THREADS / 32];
    __shared__ unsigned s_qfar, s_qnear, s_qhead, s_work, s_zr[2];
    __shared__ int s_kmax;
    __shared__ unsigned long long s_rowmask[FC_G], s_blockmask[FC_G / 4];
    __shared__ __align__(16) unsigned char s_ring[FC_CELLS];            // Chebyshev distance (cells) to the nearest occupied cell

    const int tid = threadIdx.x, lane = tid & 31;
    const int w = P.w, npix = w * w, half = w / 2;
#ifdef FC_TIMELINE
    if (tid < 24) g_fc_marks[blockIdx.x & 1023][tid] = 0;
    __syncthreads();
#endif
    FC_MARK(0);
    // Work of this CTA.  Plain call: image (b, v, c).  SCREAM_F_DUAL (the "fake" and the "real" render of a
    // training step in one launch): outputs carry a leading pose dimension [2][B]... -- 0 = clouds as given, 1 = source
    // posed -- and the CTAs of (b, v) are: source as given, source posed, and ONE for the target, which sorts the target
    // once and searches it twice: its two images differ only through the depth range (render.py:30-31 takes it over
    // source + target), i.e. through the score offsets c.
    const bool dual = (P.flags & FC_FLAG_DUAL) != 0u;
    int bid = blockIdx.x, c, pass0 = 0, pass1 = 1;
    // nparts CTAs share an image (small grids: one pair per call is 12 images on 148 SMs): every one of them sorts the whole
    // cloud, then takes the 32-pixel chunks c with c % nparts == part -- the sort is a fifth of the work, the search scales
    const int nparts = P.nparts, part = bid % nparts;
    bid /= nparts;
    if (!dual) { c = bid % 2; bid /= 2; }
    else {
        const int job = bid % 3; bid /= 3;
        c = job == 2;
        pass0 = job == 1; pass1 = job == 0 ? 1 : 2;
    }
    const int v = bid % P.V;
    const int b = bid / P.V;

    if (tid < 9) s_R[tid] = P.views[(size_t)v * 9 + tid];
    if (P.pose != nullptr && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];
    if (tid == 64) { s_zr[0] = 0xFFFFFFFFu; s_zr[1] = 0u; }

    const ItemSpan sp = item_span(P.offs, b, P.pair_n, P.pair_m);
    const float* xyz_c = (c && P.xyz_t != nullptr) ? P.xyz_t : P.xyz;      // base of this CTA's cloud
    const int64_t lo = c ? sp.tgt0 : sp.src0, hi = c ? sp.end : sp.tgt0;
    // giant clouds (forward_stream.cuh): render the survivor list of this image instead of the whole cloud,
    // unless the list overflowed.  npts = number of points to render, pidx(j) = cloud-local index of the j-th.
    const int* surv = nullptr;
    int64_t npts = hi - lo;
    if (P.surv != nullptr) {
        const size_t img = ((size_t)b * P.V + v) * 2 + c;
        const unsigned ns = P.surv_cnt[img];
        if (ns <= (unsigned)P.surv_cap) { surv = P.surv + img * (size_t)P.surv_cap; npts = (int64_t)ns; }
        else if (P.surv1 != nullptr) {
            const unsigned n1 = P.surv1_cnt[img];
            if (n1 <= (unsigned)P.surv1_cap) { surv = P.surv1 + img * (size_t)P.surv1_cap; npts = (int64_t)n1; }
        }
    }
    // a cloud of a few points is rendered by forward_tiny_kernel (launched next to this one when P.tiny is set)
    if (P.tiny && npts <= (int64_t)FC_LINEAR_MAX && surv == nullptr) return;
    const bool raw = P.flags & 1u;
    const float inv_k = -P.inv_negk;                      // 1/k, k = rho^2/2
    const float s_zero = __fmul_rn(FC_EXP_ZERO, inv_k);   // scores beyond: value exactly 0
    // scores beyond s_far cannot change the OUTPUT: all of them when the value is exactly zero; with the
    // normalised output and no winner map, already when val <= 2^-26 ((val-0.5)/0.5 rounds to -1).
    __syncthreads();

    FcShared S; S.rec = s_rec; S.start = s_start;
    S.xy2 = (unsigned)__cvta_generic_to_shared(s_rec); S.c2 = S.xy2 + (unsigned)(FC_OFF_C2 - FC_OFF_REC);
    S.rowmask = s_rowmask; S.blockmask = s_blockmask;

  for (int pass = pass0; pass < pass1; ++pass) {
    // outputs and depth range of this pass (dual: item pass * B + b of the pose-major outputs)
    const size_t ob = dual ? (size_t)pass * P.B + b : (size_t)b;
    // winner map of this pass: SCREAM_F_DUAL_NOWIN1 leaves the posed ("real") render of a dual launch without one -- nothing
    // differentiates through it (train_3d_match.py:198-201), and without it the -1 cap below applies to that render too
    int32_t* const winner = (dual && pass == 1 && (P.flags & FC_FLAG_NOWIN1)) ? nullptr : P.winner;
    const bool minus1_ok = !raw && winner == nullptr;
    const float r2cap_all = minus1_ok ? fminf(s_zero, __fmul_rn(FC_EXP_MINUS1, inv_k)) : s_zero;
    const bool posed = (dual ? pass == 1 : P.pose != nullptr) && c == 0;
    if (P.fuse_zrange) {
        // render.py:30-31 inside this kernel: min / max of the rotated depth over BOTH clouds of the item (order-preserving
        // integer codes: exact in any order, so every CTA of the item finds the same bits)
        unsigned zl = 0xFFFFFFFFu, zh = 0u;
        for (int64_t i0 = sp.src0 + tid; i0 < sp.end; i0 += FC_ZFLIGHT * FC_THREADS) {      // FC_ZFLIGHT points in flight per thread
            float x[FC_ZFLIGHT], y[FC_ZFLIGHT], z[FC_ZFLIGHT];
#pragma unroll
            for (int u = 0; u < FC_ZFLIGHT; ++u) {
                const int64_t i = i0 + u * FC_THREADS < sp.end ? i0 + u * FC_THREADS : i0;      // tail: the first point again
                const float* q = ((i >= sp.tgt0 && P.xyz_t != nullptr) ? P.xyz_t : P.xyz) + (size_t)i * 3;
                x[u] = __ldg(q); y[u] = __ldg(q + 1); z[u] = __ldg(q + 2);
            }
#pragma unroll
            for (int u = 0; u < FC_ZFLIGHT; ++u) {
                const int64_t i = i0 + u * FC_THREADS < sp.end ? i0 + u * FC_THREADS : i0;
                if (P.pose != nullptr && i < sp.tgt0) apply_pose(s_pose, x[u], y[u], z[u]);
                const unsigned e = enc_f32(rot_row(s_R + 6, x[u], y[u], z[u]));
                zl = min(zl, e); zh = max(zh, e);
            }
        }
        zl = __reduce_min_sync(0xFFFFFFFFu, zl); zh = __reduce_max_sync(0xFFFFFFFFu, zh);
        if (lane == 0) { atomicMin(&s_zr[0], zl); atomicMax(&s_zr[1], zh); }
        __syncthreads();
        FC_MARK(1);
    }
    const float zmin = dec_f32(P.fuse_zrange ? s_zr[0] : P.zenc[(ob * P.V + v) * 2]);
    const float zmax = dec_f32(P.fuse_zrange ? s_zr[1] : P.zenc[(ob * P.V + v) * 2 + 1]);
    const float range = z_range(zmin, zmax);
    const size_t img_off = ((ob * P.V + v) * 2 + c) * npix;
    if (P.zrange && c == 0 && part == 0 && tid == 0) {
        P.zrange[(ob * P.V + v) * 2] = zmin;
        P.zrange[(ob * P.V + v) * 2 + 1] = zmax;
    }
    if (zmin != zmin || zmax != zmax) {
        // a NaN coordinate anywhere in the pair: torch.min / torch.max propagate it (render.py:30), every pv of render.py:31
        // is NaN, and torch.max(dim=0) of an all-NaN column is NaN at index 0 (render.py:46,49) -- both clouds, every pixel
        for (int p = tid; p < npix; p += FC_THREADS) {
            P.imgs[img_off + p] = __int_as_float(0x7FC00000);
            if (winner) winner[img_off + p] = hi > lo ? 0 : -1;
        }
        continue;                                         // uniform over the CTA
    }
    ExactCtx E; E.xyz = xyz_c; E.lo = lo; E.R = s_R; E.pose = s_pose; E.posed = posed;
    E.zmin = zmin; E.range = range; E.negk = P.negk; E.f16 = (P.flags & FC_FLAG_HALF) != 0;
    // the target's second pass keeps the sorted records of the first (one batch only) and renews the score offsets
    const bool resort = !(pass > pass0 && npts <= (int64_t)FC_CAP);

    int batch = 0;
    int64_t c0 = 0;                                      // position in the point list of this image
    do {                                                 // at least one pass, so empty clouds still write
        const int cnt = (int)max((int64_t)0, min((int64_t)FC_CAP, npts - c0));
        const bool last = c0 + cnt >= npts;
        E.lo = lo;                                        // record indices are cloud-local

        if (resort) {
        // ---------------- P1: transform + histogram -------------------------------------------
        for (int t = tid; t < FC_CELLS + 4; t += FC_THREADS) s_hist[t] = 0u;
        __syncthreads();
        unsigned packed[FC_PPT];                          // (cell << 16) | rank inside the cell
        // the loads of points k + 1 .. k + FC_AHEAD are in flight while point k is binned (a thread that walks its points
        // strictly one by one exposes 12 serial L2 round trips here and 12 more in P3)
        auto fetch = [&](int k, float& x, float& y, float& z, int& id) {
            const int j = k * FC_THREADS + tid;
            if (j < cnt) {
                const int64_t pi = surv ? (int64_t)__ldg(surv + c0 + j) : c0 + j;
                const float* q = xyz_c + (size_t)(lo + pi) * 3;
                x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2); id = (int)pi;
            }
        };
        {
            float nx[FC_AHEAD], ny[FC_AHEAD], nz[FC_AHEAD]; int nid[FC_AHEAD];
#pragma unroll
            for (int a = 0; a < FC_AHEAD; ++a) { nx[a] = ny[a] = nz[a] = 0.f; nid[a] = 0; fetch(a, nx[a], ny[a], nz[a], nid[a]); }
#pragma unroll
            for (int k = 0; k < FC_PPT; ++k) {
                float x = nx[k % FC_AHEAD], y = ny[k % FC_AHEAD], z = nz[k % FC_AHEAD];
                if (k + FC_AHEAD < FC_PPT) fetch(k + FC_AHEAD, nx[k % FC_AHEAD], ny[k % FC_AHEAD], nz[k % FC_AHEAD], nid[k % FC_AHEAD]);
                packed[k] = 0xFFFFFFFFu;
                if (k * FC_THREADS + tid < cnt) {
                    if (posed) apply_pose(s_pose, x, y, z);
                    const float xr = rot_row(s_R, x, y, z), yr = rot_row(s_R + 3, x, y, z);
                    const int cell = cell_of(yr) * FC_G + cell_of(xr);
                    const unsigned rank = atomicAdd(&s_hist[cell], 1u);
                    packed[k] = ((unsigned)cell << 16) | (rank & 0xFFFFu);       // cnt <= 12288 < 65536
                }
            }
        }
        __syncthreads();
        FC_MARK(2);

        // ---------------- P2: exclusive scan over the cells -----------------------------------
        {
            constexpr int PER = FC_CELLS / FC_THREADS;                       // 4 consecutive cells per thread
            unsigned loc[PER], sum = 0;
#pragma unroll
            for (int k = 0; k < PER; ++k) { loc[k] = s_hist[tid * PER + k]; sum += loc[k]; }
            unsigned inc = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
                if (lane >= d) inc += o;
            }
            if (lane == 31) s_warp_tot[tid >> 5] = inc;
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
            for (int k = 0; k < PER; ++k) { s_start[tid * PER + k] = (unsigned short)run; run += loc[k]; }
            if (tid == FC_THREADS - 1) s_start[FC_CELLS] = (unsigned short)run;
        }
        __syncthreads();
        FC_MARK(3);

        // ---------------- P3: occupancy masks + scatter of the records into cell order ---------
        if (tid < (FC_G / 4) * 32) {                      // warp wq builds rows 4wq..4wq+3
            const int wq = tid >> 5;
            unsigned long long bm = 0ull;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int i = 4 * wq + t;
                const unsigned short* st = s_start + i * FC_G;
                const unsigned lo32 = __ballot_sync(0xFFFFFFFFu, st[lane + 1] != st[lane]);
                const unsigned hi32 = __ballot_sync(0xFFFFFFFFu, st[lane + 33] != st[lane + 32]);
                const unsigned long long rm = ((unsigned long long)hi32 << 32) | lo32;
                if (lane == 0) s_rowmask[i] = rm;
                bm |= rm;
            }
            if (lane == 0) s_blockmask[wq] = bm;
        }
        // Chebyshev distance (cells) to the nearest occupied cell, by dilation of the row masks: D_k = the cells within k of
        // an occupied one (row above | row | row below, then x | x << 1 | x >> 1).  ONE warp does the steps while the others
        // scatter the records (the steps are serial, ~20 instructions each, and nobody waits for them); the masks go to the
        // histogram's memory, free since P2.  Steps beyond the radius that can matter are not needed: a cell that D_K does
        // not cover is farther than anything that can change the output.
        const int need_ring = (int)ceilf(sqrtf(r2cap_all) * FC_CELLS_PER_NDC) + 2;
        const bool dil = need_ring <= FC_DIL_MAX;         // (a smaller rho reaches farther: the row-wise transform below)
        unsigned long long* s_dil = reinterpret_cast<unsigned long long*>(s_hist);      // [k - 1][row], k = 1 .. FC_DIL_MAX
        if (dil && tid < (FC_G / 4 + 1) * 32) {
            if (tid < (FC_G / 4) * 32) {                  // the mask builders: arrive, do not wait
                __threadfence_block();
                asm volatile("bar.arrive 1, %0;" :: "n"((FC_G / 4 + 1) * 32) : "memory");
            } else {
                asm volatile("bar.sync 1, %0;" :: "n"((FC_G / 4 + 1) * 32) : "memory");
                unsigned long long m0 = s_rowmask[2 * lane], m1 = s_rowmask[2 * lane + 1];
                int kmax = 0;
                for (int k = 1; k <= need_ring; ++k) {
                    unsigned long long up = __shfl_up_sync(0xFFFFFFFFu, m1, 1), dn = __shfl_down_sync(0xFFFFFFFFu, m0, 1);
                    if (lane == 0) up = 0ull;
                    if (lane == 31) dn = 0ull;
                    const unsigned long long v0 = up | m0 | m1, v1 = m0 | m1 | dn;
                    const unsigned long long n0 = v0 | (v0 << 1) | (v0 >> 1), n1 = v1 | (v1 << 1) | (v1 >> 1);
                    const bool same = n0 == m0 && n1 == m1;
                    m0 = n0; m1 = n1;
                    if (__all_sync(0xFFFFFFFFu, same)) break;          // nothing grows any more: all covered, or an empty batch
                    *reinterpret_cast<ulonglong2*>(s_dil + (k - 1) * FC_G + 2 * lane) = make_ulonglong2(m0, m1);
                    kmax = k;
                }
                if (lane == 0) s_kmax = kmax;
            }
        }
        {
            float nx[FC_AHEAD], ny[FC_AHEAD], nz[FC_AHEAD]; int nid[FC_AHEAD];
#pragma unroll
            for (int a = 0; a < FC_AHEAD; ++a) { nx[a] = ny[a] = nz[a] = 0.f; nid[a] = 0; fetch(a, nx[a], ny[a], nz[a], nid[a]); }
#pragma unroll
            for (int k = 0; k < FC_PPT; ++k) {
                float x = nx[k % FC_AHEAD], y = ny[k % FC_AHEAD], z = nz[k % FC_AHEAD]; const int id = nid[k % FC_AHEAD];
                if (k + FC_AHEAD < FC_PPT) fetch(k + FC_AHEAD, nx[k % FC_AHEAD], ny[k % FC_AHEAD], nz[k % FC_AHEAD], nid[k % FC_AHEAD]);
                if (packed[k] != 0xFFFFFFFFu) {
                    if (posed) apply_pose(s_pose, x, y, z);
                    const float xr = rot_row(s_R, x, y, z), yr = rot_row(s_R + 3, x, y, z), zr = rot_row(s_R + 6, x, y, z);
                    const unsigned pos = (unsigned)s_start[packed[k] >> 16] + (packed[k] & 0xFFFFu);
                    float* rx = rec_x(s_rec, (int)pos);
                    rx[0] = xr; rx[2] = yr;
                    // c = -ln(pv)/k >= 0; pv == 0 -> +inf (never wins); pv < 0 (rounding corner) -> NaN (never wins)
                    *rec_c(s_rec, (int)pos) = __fmul_rn(logf(depth_weight_m(zr, zmin, range, (P.flags & FC_FLAG_HALF) != 0)), P.inv_negk);
                    *rec_idx(s_rec, (int)pos) = id;
                }
            }
        }
        if (tid == 0 && (cnt & 1)) {                      // pad behind an odd count: a record that can never win
            float* rx = rec_x(s_rec, cnt);
            rx[0] = 0.0f; rx[2] = 0.0f; *rec_c(s_rec, cnt) = INFINITY; *rec_idx(s_rec, cnt) = 0;
        }
        __syncthreads();
        FC_MARK(4);

        // ---------------- P3b: Chebyshev distance transform of the cell occupancy -------------
        // h[i][j] = distance to the nearest occupied cell of row i (bit tricks on the row mask), then
        // ring[i][j] = min_i' max(|i-i'|, h[i'][j]), rows outwards with early exit.  Lets empty space be
        // classified without touching the records: a pixel whose ring is >= 2 has nothing in its 3 x 3
        // window, and its nearest point is at least (ring - 1 + edge) cells away.
        if (dil) {
            // ring = the number of steps D_0 .. D_kmax that do not cover the cell (D_k grows with k).  A thread counts for four
            // neighbouring cells of a row at once: the four mask bits spread into the four bytes of one counter word.
            const int kmax = s_kmax;
            {
                const int i = tid >> 4, j0 = (tid & 15) * 4;
                unsigned nib = (unsigned)(~s_rowmask[i] >> j0) & 0xFu;
                unsigned cntw = (nib * 0x00204081u) & 0x01010101u;
                for (int k = 1; k <= kmax; ++k) {
                    nib = (unsigned)(~s_dil[(k - 1) * FC_G + i] >> j0) & 0xFu;
                    cntw += (nib * 0x00204081u) & 0x01010101u;
                }
                // still uncovered after the last step: beyond the steps that matter (count = need_ring + 1), or -- the growth
                // stopped early -- an empty batch: nothing anywhere
                if (nib != 0u && kmax < need_ring) {
                    const unsigned far = (nib * 0x00204081u) & 0x01010101u;
                    cntw = (cntw & ~(far * 0xFFu)) | (far * (unsigned)FC_G);
                }
                reinterpret_cast<unsigned*>(s_ring)[tid] = cntw;
            }
            __syncthreads();
            FC_MARK(5);
        } else {
            unsigned char* s_h = reinterpret_cast<unsigned char*>(s_queue);      // scratch until phase A
            for (int cell = tid; cell < FC_CELLS; cell += FC_THREADS) {
                const int i = cell >> 6, j = cell & 63;
                const unsigned long long m = s_rowmask[i];
                const unsigned long long ml = m & ((2ull << j) - 1ull), mr = m >> j;
                int h = FC_G;
                if (ml) h = j - (63 - __clzll((long long)ml));
                if (mr) h = min(h, __ffsll((long long)mr) - 1);
                s_h[cell] = (unsigned char)h;
            }
            __syncthreads();
            for (int cell = tid; cell < FC_CELLS; cell += FC_THREADS) {
                const int i = cell >> 6;
                int best = s_h[cell];
                for (int d = 1; d < best; ++d) {
                    if (i - d >= 0) best = min(best, max(d, (int)s_h[cell - d * FC_G]));
                    if (i + d < FC_G) best = min(best, max(d, (int)s_h[cell + d * FC_G]));
                }
                s_ring[cell] = (unsigned char)best;
            }
            __syncthreads();
            FC_MARK(5);
        }

        } else {
            // same positions, same cell order, same occupancy: only c = -ln(pv)/k follows the new depth range
            for (int pos = tid; pos < cnt; pos += FC_THREADS) {
                const float* q = xyz_c + (size_t)(lo + *rec_idx(s_rec, pos)) * 3;
                const float zr = rot_row(s_R + 6, __ldg(q), __ldg(q + 1), __ldg(q + 2));
                *rec_c(s_rec, pos) = __fmul_rn(logf(depth_weight_m(zr, zmin, range, (P.flags & FC_FLAG_HALF) != 0)), P.inv_negk);
            }
            __syncthreads();
        }

        // ---------------- P4: per-pixel search, one 64 x 64 pixel tile at a time ---------------
        // (lattice mode, P.pstep == 2: one "tile" of 64 x 64 pixels two apart, starting at (P.prow0, P.pcol0))
        constexpr int pstep = PSTEP;
        const int tiles_side = pstep == 2 ? 1 : w / FC_TILE;
        for (int tile = 0; tile < tiles_side * tiles_side; ++tile) {
            const int trow0 = pstep == 2 ? P.prow0 : (tile / tiles_side) * FC_TILE, tcol0 = pstep == 2 ? P.pcol0 : (tile % tiles_side) * FC_TILE;
            if (tid == 0) { s_qfar = 0u; s_qnear = 0u; s_qhead = 0u; s_work = 0u; }
            __syncthreads();

            // write one pixel: merge with what the earlier batches of a large cloud found, (value, lowest index)
            auto put = [&](int row, int col, float val, int id) {
                const size_t o = img_off + (size_t)row * w + col;
                if (batch > 0) {
                    const float pb = P.imgs[o];
                    const int pid = winner ? winner[o] : 0;
                    if (!(val > pb || (val == pb && id < pid))) { val = pb; id = pid; }
                }
                P.imgs[o] = (raw || !last) ? val : __fmul_rn(__fsub_rn(val, 0.5f), 2.0f);
                if (winner) winner[o] = id;
            };
            auto geom = [&](int row, int col) {
                PixelGeom g;
                g.cx = pixel_centre(col, half); g.cy = pixel_centre(row, half);
                g.ux = __fmul_rn(__fadd_rn(g.cx, 1.0f), FC_CELLS_PER_NDC);
                g.uy = __fmul_rn(__fadd_rn(g.cy, 1.0f), FC_CELLS_PER_NDC);
                return g;
            };

            // ---- classification: background pixels are written, "far" pixels (no occupied cell within the 3 x 3 window:
            // the long disc scans) are queued.  Rows are visited from the tile's top and bottom edges towards its middle,
            // so the pixels farthest from a centred cloud enter the queue first: longest tasks first.
            const int own_chunks = (FC_TPIX / 32 - part + nparts - 1) / nparts;      // chunks c < 128 with c % nparts == part
            for (int ol = tid; ol < own_chunks * 32; ol += FC_THREADS) {
                const int vl = ((ol >> 5) * nparts + part) * 32 + (ol & 31);      // the chunks this CTA owns, in visiting order
                const int rl = lpt_row(vl / FC_TILE);
                const int row = trow0 + rl * pstep, col = tcol0 + (vl % FC_TILE) * pstep;
                int ring; bool blank;
                pixel_class(row, col, half, r2cap_all, s_ring, ring, blank);
                if (blank) put(row, col, 0.0f, 0);        // background: value 0 (index 0), or what earlier batches found
                const bool far = ring >= 2 && !blank;
                const unsigned fm = __ballot_sync(0xFFFFFFFFu, far);
                if (fm) {
                    unsigned fb = 0;
                    if (lane == 0) fb = atomicAdd(&s_qfar, (unsigned)__popc(fm));
                    fb = __shfl_sync(0xFFFFFFFFu, fb, 0);
                    if (far) s_queue[fb + __popc(fm & ((1u << lane) - 1u))] = (unsigned short)(rl * FC_TILE + vl % FC_TILE);
                }
            }
            __syncthreads();
            FC_MARK(6);

            // ---- disc scan of one queued pixel (far: no candidate; near: the best record of its 3 x 3 window) ----
            // (defer, lattice renders only: a pixel that needs the exact pass over a band of records is handed back -- the winner's
            // record is returned, nothing is written -- and the caller settles it with the whole warp)
            auto scan_pixel = [&](int pl, int cand, bool defer) -> int {
                const int row = trow0 + (pl / FC_TILE) * pstep, col = tcol0 + (pl % FC_TILE) * pstep;
                const PixelGeom g = geom(row, col);
                // nothing beyond r2cap can change the output of this pixel
                const float r2cap = r2cap_all;
                float s1 = cand == 0xFFFF ? INFINITY : score_of(S, cand, g), s2; int a1;
                float val = 0.0f; int id = 0;
                disc_scan<false>(S, E, g, fminf(r2cap, s1 + score_margin(s1, inv_k)), inv_k, s1, s2, a1, val, id);
                if (a1 >= 0 && s1 <= r2cap) {
                    // scores within s_thr of the best may swap order through rounding.  In the denormal
                    // range the value is a multiple of 2^-149: two values closer than ~8 of those may tie
                    // or swap, i.e. scores within -ln(1 - u)/k <= 1.5 u/k, u = 8 * 2^-149 * e^(k s1);
                    // from u >= 1/2 on (the last ~2.5 e-folds above zero) everything non-zero is evaluated.
                    float s_thr = s1 + score_margin(s1, inv_k);
                    const float e1 = __fmul_rn(s1, -P.negk);
                    if (e1 >= FC_EXP_DENORM) {
                        const float lu = e1 - 101.2f;                 // ln u
                        s_thr = lu >= -0.7f ? s_zero : s_thr + 1.5f * expf(lu) * inv_k;
                    }
                    if (s2 > s_thr) {
                        val = exact_value(S, E, a1, g, id);
                    } else {
                        if (PSTEP == 2 && defer) return a1;
                        disc_scan<true>(S, E, g, fminf(s_thr, s_zero), inv_k, s1, s2, a1, val, id);
                    }
                    if (val == 0.0f) id = 0;              // all-zero pixel: torch.max returns index 0
                }
                put(row, col, val, id);
                return -1;
            };

            // ---- the same for ONE pixel by a whole warp (all lanes call with the same arguments): small grids are bound by
            // latency, a lane's disc scan is a serial chain of ~1500 dependent instructions (3 us), and three lanes of a warp
            // in three different scans run one after the other.  Here every lane takes grid rows of the disc (fixed radius from
            // a first guess), the lanes' (best, runner-up) merge through shuffles, and the exact pass is shared the same way.
            // Returns (warp-uniform) the winner's record when the search settled the pixel -- the caller evaluates and writes it
            // (finish_pixel: one lane per pixel, so the L2 round trips of several pixels overlap) -- or -1: written here.
            auto finish_pixel = [&](int pl, int a) {
                const int row = trow0 + (pl / FC_TILE) * pstep, col = tcol0 + (pl % FC_TILE) * pstep;
                int id;
                float val = exact_value(S, E, a, geom(row, col), id);
                if (val == 0.0f) id = 0;                  // all-zero pixel: torch.max returns index 0
                put(row, col, val, id);
            };
            auto search_pixel_warp = [&](int pl, int cand) -> int {
                const int row = trow0 + (pl / FC_TILE) * pstep, col = tcol0 + (pl % FC_TILE) * pstep;
                const PixelGeom g = geom(row, col);
                const int ci = (int)g.uy, cj = (int)g.ux;
                const float r2cap = r2cap_all;
                float sb = cand == 0xFFFF ? INFINITY : score_of(S, cand, g);
                const long long w_t0 = FC_CLOCK(); (void)w_t0;
                if (cand == 0xFFFF) {
                    // first guess: per grid row the occupied cell nearest to the pixel's column, the nearest of those over the
                    // warp, the best score of its records (any real point bounds the best score from above)
                    unsigned key = 0xFFFFFFFFu;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int i = lane + 32 * h;
                        const unsigned long long m = s_rowmask[i];
                        const unsigned long long ml = m & ((2ull << cj) - 1ull), mr = m >> cj;
                        int jn = -1;
                        if (ml) jn = 63 - __clzll((long long)ml);
                        if (mr) { const int jr = cj + __ffsll((long long)mr) - 1; if (jn < 0 || jr - cj < cj - jn) jn = jr; }
                        if (jn >= 0) key = min(key, ((unsigned)((i - ci) * (i - ci) + (jn - cj) * (jn - cj)) << 16) | (unsigned)(i * FC_G + jn));
                    }
                    key = __reduce_min_sync(0xFFFFFFFFu, key);
                    if (key != 0xFFFFFFFFu) {
                        const int cell = (int)(key & 0xFFFFu), r0 = s_start[cell] + lane;
                        sb = warp_min(r0 < (int)s_start[cell + 1] ? score_of(S, r0, g) : INFINITY);
                    }
                }
                const long long w_t1 = FC_CLOCK(); (void)w_t1;
                const float r2 = fminf(r2cap, sb + score_margin(sb, inv_k));
                const float rc = radius_cells(r2);
                const int reach = (int)fminf(rc, (float)FC_G) + 1;
                const int i_lo = max(0, ci - reach), i_hi = min(FC_G - 1, ci + reach);
                float s1 = INFINITY, s2 = INFINITY; int a1 = -1;
                float val = 0.0f; int id = 0;
                for (int i = i_lo + lane; i <= i_hi; i += 32) row_scan<false>(S, E, g, i, rc, 0.0f, s1, s2, a1, val, id);
                __syncwarp();
                const long long w_t2 = FC_CLOCK(); (void)w_t2;
                a1 = best_of_pair(S, a1, g);
                const float m1 = warp_min(s1);
                const unsigned wm = __ballot_sync(0xFFFFFFFFu, s1 == m1 && a1 >= 0);
                if (wm != 0u && m1 <= r2cap) {
                    const int wl = __ffs(wm) - 1;
                    // runner-up of the disc: the winner lane's own, or the best of any other lane (a second lane with the same
                    // best score leaves s2 == s1: unresolved, as in the serial search)
                    const float m2 = warp_min(lane == wl ? s2 : s1);
                    const int a = __shfl_sync(0xFFFFFFFFu, a1, wl);
                    float s_thr = m1 + score_margin(m1, inv_k);
                    const float e1 = __fmul_rn(m1, -P.negk);
                    if (e1 >= FC_EXP_DENORM) {
                        const float lu = e1 - 101.2f;
                        s_thr = lu >= -0.7f ? s_zero : s_thr + 1.5f * expf(lu) * inv_k;
                    }
                    if (lane == 0) { FC_COUNT(16, w_t1 - w_t0); FC_COUNT(17, w_t2 - w_t1); FC_COUNT(18, FC_CLOCK() - w_t2); FC_COUNT(19, 1); FC_COUNT(20, reach); }
                    if (m2 > s_thr) return a;
                    {
                        const float r2e = fminf(s_thr, s_zero), rce = radius_cells(r2e);
                        const int reach_e = (int)fminf(rce, (float)FC_G) + 1;
                        for (int i = max(0, ci - reach_e) + lane; i <= min(FC_G - 1, ci + reach_e); i += 32)
                            row_scan<true>(S, E, g, i, rce, r2e, s1, s2, a1, val, id);
#pragma unroll
                        for (int d = 16; d > 0; d >>= 1) {                    // (value, lowest index) over the lanes
                            const float ov = __shfl_xor_sync(0xFFFFFFFFu, val, d);
                            const int oid = __shfl_xor_sync(0xFFFFFFFFu, id, d);
                            if (ov > val || (ov == val && oid < id)) { val = ov; id = oid; }
                        }
                    }
                    if (val == 0.0f) id = 0;
                }
                if (lane == 0) put(row, col, val, id);
                return -1;
            };

            // ---- one dynamic work list: the far batches first (long), then the 3 x 3 chunks (short, uniform) fill the
            // gaps -- the heavy scans do not start after a barrier behind the 3 x 3 phase.
            // batch = the queue entries a warp takes at once: 32 when the GPU is full (issue-bound: lanes are what counts);
            // when several CTAs share an image the kernel is latency-bound instead, a warp's time is its slowest lane's
            // serial chain, and the entries are spread over all 32 warps
            // Small grids (coop): a far pixel is a task of its own, searched by the whole warp, and the 3 x 3 chunks -- now the
            // long tasks -- go first.
#ifndef FC_LATTICE_COOP
#define FC_LATTICE_COOP 2
#endif
            // Lattice renders of a giant cloud's survivor lists (PSTEP 2): their corner pixels sit in the last e-folds above zero at
            // rho 36, where EVERY record of the band is evaluated exactly from its point in a 10^8-point cloud -- lane per pixel, a
            // batch of them ran for 2.9 ms of serial DRAM round trips while the CTA waited.  FC_LATTICE_COOP 1: all far pixels by
            // whole warps (1.8 ms per launch, but 3 lanes per instruction on the ordinary ones); 2: lane per pixel, and the pixels
            // that need the exact pass over a band are handed to the whole warp afterwards.
            const bool coop = nparts > 1 || (FC_LATTICE_COOP == 1 && PSTEP == 2);
            const bool coop_near = coop || PSTEP == 2;
            const int nfar = (int)s_qfar;
            const int far_take = coop ? FC_COOP_FAR : FC_FAR_TAKE;
            const int far_batches = (nfar + far_take - 1) / far_take;
            const int nchunks = own_chunks;
            for (;;) {
                int t = 0;
                if (lane == 0) t = (int)atomicAdd(&s_work, 1u);
                t = __shfl_sync(0xFFFFFFFFu, t, 0);
                if (t >= far_batches + nchunks) break;
                const long long task_t0 = FC_CLOCK(); (void)task_t0;
                if (coop) {
                    if (t >= nchunks) {
                        int my_pl = 0, my_a = -1, prev_a = -1;
                        for (int k = 0; k < FC_COOP_FAR; ++k) {
                            const int qi = (t - nchunks) * FC_COOP_FAR + k;
                            if (qi >= nfar) break;
                            const int pl = s_queue[qi];
                            // the winner of the previous pixel (usually its neighbour in the row) is the first guess of this one
                            const int a = search_pixel_warp(pl, prev_a >= 0 ? prev_a : 0xFFFF);
                            prev_a = a;
                            if (a >= 0 && lane == k) { my_pl = pl; my_a = a; }
                        }
                        if (my_a >= 0) finish_pixel(my_pl, my_a);
                        __syncwarp();
                        if (lane == 0) { FC_MAXOF(12, FC_CLOCK() - task_t0); FC_COUNT(14, FC_CLOCK() - task_t0); }
                        continue;
                    }
                    t += far_batches;                    // chunk t of the list below
                } else if (t < far_batches) {
                    const int qi = t * far_take + lane;
                    int dpl = 0, dcand = -1;
                    if (lane < far_take && qi < nfar) { dpl = s_queue[qi]; dcand = scan_pixel(dpl, 0xFFFF, true); }
                    __syncwarp();
                    if (PSTEP == 2) {
                        unsigned dm = __ballot_sync(0xFFFFFFFFu, dcand >= 0);
                        while (dm) {
                            const int src = __ffs(dm) - 1;
                            dm &= dm - 1u;
                            const int fpl = __shfl_sync(0xFFFFFFFFu, dpl, src);
                            const int fa = search_pixel_warp(fpl, __shfl_sync(0xFFFFFFFFu, dcand, src));
                            if (fa >= 0 && lane == 0) finish_pixel(fpl, fa);
                        }
                    }
                    if (lane == 0) { FC_MAXOF(12, ((FC_CLOCK() - task_t0) << 20) | ((long long)batch << 16) | (long long)s_queue[t * far_take]); FC_COUNT(14, FC_CLOCK() - task_t0); }
                    continue;
                }
                // ---- 3 x 3 cells, lane per pixel ----
                const int vl = ((t - far_batches) * nparts + part) * 32 + lane;   // same chunks, same visiting order: a pixel
                const int pl = lpt_row(vl / FC_TILE) * FC_TILE + vl % FC_TILE;    // belongs to ONE CTA in every batch of a cloud
                const int row = trow0 + (pl / FC_TILE) * pstep, col = tcol0 + (pl % FC_TILE) * pstep;
                const PixelGeom g = geom(row, col);
                const int ci = (int)g.uy, cj = (int)g.ux;
                const float fx = g.ux - (float)cj, fy = g.uy - (float)ci;
                const float edge = fminf(fminf(fx, 1.0f - fx), fminf(fy, 1.0f - fy));
                const float cov = __fmul_rn(1.0f + edge, 1.0f / FC_CELLS_PER_NDC);
                const float cov2 = __fmul_rn(__fmul_rn(cov, cov), 0.99999f);
                const int j0 = max(cj - 1, 0), j1 = min(cj + 1, FC_G - 1);
                float s1 = INFINITY, s2 = INFINITY; int a1 = -1;
                const int ring = s_ring[ci * FC_G + cj];
                bool found = false;
                for (int i = max(ci - 1, 0); ring < 2 && i <= min(ci + 1, FC_G - 1); ++i) {
                    const int rb = s_start[i * FC_G + j0], re = s_start[i * FC_G + j1 + 1];
                    scan_run(S, rb, re, g, s1, s2, a1, found);
                }
                a1 = best_of_pair(S, a1, g);
                const float m = score_margin(s1, inv_k);
                const bool resolved = (s1 + m < cov2) && (s2 > s1 + m) && (__fmul_rn(s1, -P.negk) < FC_EXP_DENORM);
                if (resolved) {
                    int id;
                    const float val = exact_value(S, E, a1, g, id);
                    put(row, col, val, id);
                }
                // something in the 3 x 3 window, but not settled by it: the near queue
                const bool near = !resolved && ring < 2;
                const unsigned nm = __ballot_sync(0xFFFFFFFFu, near);
                if (nm) {
                    unsigned nb = 0;
                    if (lane == 0) nb = atomicAdd(&s_qnear, (unsigned)__popc(nm));
                    nb = __shfl_sync(0xFFFFFFFFu, nb, 0);
                    if (near) s_queue[FC_TPIX - 1 - (nb + __popc(nm & ((1u << lane) - 1u)))] = (unsigned short)pl;
                }
                if (lane == 0) { FC_MAXOF(13, FC_CLOCK() - task_t0); FC_COUNT(15, FC_CLOCK() - task_t0); }
            }
            __syncthreads();
            FC_MARK(7);

            // ---- the near queue, 32 pixels at a time with all lanes busy: the 5 x 5 window, five row runs in lockstep.
            // Typical case: the winner is a deep point (score offset c > 2.25 cells^2, i.e. pv < 0.53), so the 3 x 3 radius
            // test can never pass although the winner is a cell away; 17 % of the pixels of a unit-ball view, none of them
            // with a best score beyond the 5 x 5 reach of 6.25 cells^2.  What stays open takes the general disc scan.
            const int nnear = (int)s_qnear;
            if (tid == 0) { FC_COUNT(9, nfar); FC_COUNT(10, nnear); }
            const int near_take = (FC_NEAR_SPREAD || nparts > 1) ? max(1, min(32, (nnear + 31) >> 5)) : 32;
            // Small grids: four pixels per warp, eight lanes each -- five of them take one row run of the 5 x 5 window each, the
            // (best, runner-up) of the rows merge through shuffles, the group's first lane decides and evaluates.  (Lanes in
            // lockstep instead of three lanes walking three different windows one after the other.)
            for (; coop_near;) {
                int q0 = 0;
                if (lane == 0) q0 = (int)atomicAdd(&s_qhead, 4u);
                q0 = __shfl_sync(0xFFFFFFFFu, q0, 0);
                if (q0 >= nnear) break;
                const int grp = lane >> 3, gl = lane & 7, qi = q0 + grp;
                const bool act = qi < nnear;
                const int pl = act ? s_queue[FC_TPIX - 1 - qi] : 0;
                const int row = trow0 + (pl / FC_TILE) * pstep, col = tcol0 + (pl % FC_TILE) * pstep;
                const PixelGeom g = geom(row, col);
                const int ci = (int)g.uy, cj = (int)g.ux;
                float s1 = INFINITY, s2 = INFINITY; int a1 = -1;
                bool found = false;
                const int i = ci - 2 + gl;
                if (act && gl < 5 && i >= 0 && i < FC_G)
                    scan_run(S, s_start[i * FC_G + max(cj - 2, 0)], s_start[i * FC_G + min(cj + 2, FC_G - 1) + 1], g, s1, s2, a1, found);
                a1 = best_of_pair(S, a1, g);
                float m1 = s1;
#pragma unroll
                for (int d = 4; d > 0; d >>= 1) m1 = fminf(m1, __shfl_xor_sync(0xFFFFFFFFu, m1, d));
                const unsigned wm = (__ballot_sync(0xFFFFFFFFu, s1 == m1 && a1 >= 0) >> (8 * grp)) & 0xFFu;
                const int wl = 8 * grp + (wm ? __ffs(wm) - 1 : 0);
                float m2 = (wm && lane == wl) ? s2 : s1;
#pragma unroll
                for (int d = 4; d > 0; d >>= 1) m2 = fminf(m2, __shfl_xor_sync(0xFFFFFFFFu, m2, d));
                const int a = __shfl_sync(0xFFFFFFFFu, a1, wl);
                int fb_cand = -2;
                if (act && gl == 0) {
                    const float fx = g.ux - (float)cj, fy = g.uy - (float)ci;
                    const float edge = fminf(fminf(fx, 1.0f - fx), fminf(fy, 1.0f - fy));
                    const float cov = __fmul_rn(2.0f + edge, 1.0f / FC_CELLS_PER_NDC);
                    const float cov2 = __fmul_rn(__fmul_rn(cov, cov), 0.99999f);
                    const float m = score_margin(m1, inv_k);
                    if (wm && (m1 + m < cov2) && (m2 > m1 + m) && (__fmul_rn(m1, -P.negk) < FC_EXP_DENORM)) {
                        int id;
                        const float val = exact_value(S, E, a, g, id);
                        put(row, col, val, id);
                    } else {
                        FC_COUNT(11, 1);
                        fb_cand = wm ? a : 0xFFFF;
                    }
                }
                unsigned fbm = __ballot_sync(0xFFFFFFFFu, fb_cand != -2);
                while (fbm) {
                    const int src = __ffs(fbm) - 1;
                    fbm &= fbm - 1u;
                    const int fpl = __shfl_sync(0xFFFFFFFFu, pl, src);
                    const int fa = search_pixel_warp(fpl, __shfl_sync(0xFFFFFFFFu, fb_cand, src));
                    if (fa >= 0 && lane == 0) finish_pixel(fpl, fa);
                }
            }
#ifndef FC_NEAR_DEFER
#define FC_NEAR_DEFER 1
#endif
            // (FC_NEAR_DEFER: the pixels the 5 x 5 window leaves open are settled afterwards by the whole warp, one at a time -- rows of
            // the disc per lane -- instead of by their own lanes inside the lockstep batch, where each is a serial chain of ~1500
            // dependent instructions that the other lanes of the batch, and the barrier behind this phase, wait for)
            for (; !coop_near;) {
                int qi = 0;
                if (lane == 0) qi = (int)atomicAdd(&s_qhead, (unsigned)near_take);
                qi = __shfl_sync(0xFFFFFFFFu, qi, 0) + lane;
                if (qi - lane >= nnear) break;
                const bool act = lane < near_take && qi < nnear;
                if (!FC_NEAR_DEFER && !act) continue;
                int fb_cand = -2, pl = 0;
                if (act) {
                    pl = s_queue[FC_TPIX - 1 - qi];
                    const int row = trow0 + (pl / FC_TILE) * pstep, col = tcol0 + (pl % FC_TILE) * pstep;
                    const PixelGeom g = geom(row, col);
                    const int ci = (int)g.uy, cj = (int)g.ux;
                    const float fx = g.ux - (float)cj, fy = g.uy - (float)ci;
                    const float edge = fminf(fminf(fx, 1.0f - fx), fminf(fy, 1.0f - fy));
                    float s1 = INFINITY, s2 = INFINITY; int a1 = -1;
                    bool found = false;
                    {
                        const int j0 = max(cj - 2, 0), j1 = min(cj + 2, FC_G - 1);
                        for (int i = max(ci - 2, 0); i <= min(ci + 2, FC_G - 1); ++i)
                            scan_run(S, s_start[i * FC_G + j0], s_start[i * FC_G + j1 + 1], g, s1, s2, a1, found);
                    }
                    a1 = best_of_pair(S, a1, g);
                    const float cov = __fmul_rn(2.0f + edge, 1.0f / FC_CELLS_PER_NDC);
                    const float cov2 = __fmul_rn(__fmul_rn(cov, cov), 0.99999f);
                    const float m = score_margin(s1, inv_k);
                    if ((s1 + m < cov2) && (s2 > s1 + m) && (__fmul_rn(s1, -P.negk) < FC_EXP_DENORM)) {
                        int id;
                        const float val = exact_value(S, E, a1, g, id);
                        put(row, col, val, id);
                    } else {
                        FC_COUNT(11, 1);
                        if (FC_NEAR_DEFER) fb_cand = a1 < 0 ? 0xFFFF : a1;
                        else scan_pixel(pl, a1 < 0 ? 0xFFFF : a1, false);
                    }
                }
                if (FC_NEAR_DEFER) {
                    __syncwarp();
                    unsigned fbm = __ballot_sync(0xFFFFFFFFu, fb_cand != -2);
                    while (fbm) {
                        const int src = __ffs(fbm) - 1;
                        fbm &= fbm - 1u;
                        const int fpl = __shfl_sync(0xFFFFFFFFu, pl, src);
                        const int fa = search_pixel_warp(fpl, __shfl_sync(0xFFFFFFFFu, fb_cand, src));
                        if (fa >= 0 && lane == 0) finish_pixel(fpl, fa);
                    }
                }
            }
            __syncthreads();
            FC_MARK(8);
        }
        c0 += cnt; ++batch;
    } while (c0 < npts);

    // empty cloud: the reference would raise; the batched C ABI writes the background and winner -1
    if (hi <= lo && winner) {
        for (int p = tid; p < npix; p += FC_THREADS) {
            const int vl = lpt_vis((p / w) % FC_TILE) * FC_TILE + (p % w) % FC_TILE;
            if ((vl >> 5) % nparts == part) winner[img_off + p] = -1;      // the pixels this CTA owns
        }
    }
    __syncthreads();
  }
}
static const auto forward_culled_kernel = &forward_culled_kernel_t<1>;
static const auto forward_culled_lattice_kernel = &forward_culled_kernel_t<2>;

// Clouds of up to FC_LINEAR_MAX points (OpenGG's 6 x 6 DEM against 65536 pixels at w = 256): every pixel of such an image
// would be a "far" pixel walking the rows of an almost empty grid (~1500 instructions); scoring all records costs ~8 each.
// One small CTA per (item, view, cloud) image, launched next to forward_culled_kernel (which returns at once for these
// images, as this kernel does for all others: 2 KB of shared memory, so thousands of them come and go in microseconds):
// records in point order (no sort), then linear_batch.  A kernel of its own so that the common path keeps its register
// allocation (the same code inlined, or called, inside it cost cfg2 2.5 %).
__global__ void __launch_bounds__(FT_THREADS)
forward_tiny_kernel(FwdParams P)
{
    __shared__ float4 s_xy2[FC_LINEAR_MAX / 2 + 1];
    __shared__ float2 s_c2[FC_LINEAR_MAX / 2 + 1];
    __shared__ float s_R[9], s_pose[12];
    // grid: (image, slice of FT_SLICE pixels) -- an image of w = 256 is 16 CTAs
    const int tid = threadIdx.x, w = P.w, npix = w * w, slices = npix / FT_SLICE;
    int bid = blockIdx.x;
    const int slice = bid % slices; bid /= slices;
    const int c = bid % 2; bid /= 2;
    const int v = bid % P.V, b = bid / P.V;
    const ItemSpan sp = item_span(P.offs, b, P.pair_n, P.pair_m);
    const int64_t lo = c ? sp.tgt0 : sp.src0, hi = c ? sp.end : sp.tgt0;
    const int64_t npts = hi - lo;
    if (npts > (int64_t)FC_LINEAR_MAX) return;            // forward_culled_kernel's image
    const int p_begin = slice * FT_SLICE, p_end = p_begin + FT_SLICE;
    if (tid < 9) s_R[tid] = P.views[(size_t)v * 9 + tid];
    const bool posed = P.pose != nullptr && c == 0;
    if (posed && tid >= 32 && tid < 44) s_pose[tid - 32] = P.pose[(size_t)b * 12 + tid - 32];
    __syncthreads();
    const float* xyz_c = (c && P.xyz_t != nullptr) ? P.xyz_t : P.xyz;
    const float zmin = dec_f32(P.zenc[((size_t)b * P.V + v) * 2]), zmax = dec_f32(P.zenc[((size_t)b * P.V + v) * 2 + 1]);
    const float range = z_range(zmin, zmax);
    const size_t img_off = (((size_t)b * P.V + v) * 2 + c) * npix;
    if (P.zrange && c == 0 && slice == 0 && tid == 0) { P.zrange[((size_t)b * P.V + v) * 2] = zmin; P.zrange[((size_t)b * P.V + v) * 2 + 1] = zmax; }
    if (zmin != zmin || zmax != zmax) {                   // a NaN coordinate in the pair: see forward_culled_kernel
        for (int p = p_begin + tid; p < p_end; p += FT_THREADS) {
            P.imgs[img_off + p] = __int_as_float(0x7FC00000);
            if (P.winner) P.winner[img_off + p] = hi > lo ? 0 : -1;
        }
        return;
    }
    const bool raw = P.flags & 1u, f16 = (P.flags & FC_FLAG_HALF) != 0;
    const float inv_k = -P.inv_negk, s_zero = __fmul_rn(FC_EXP_ZERO, inv_k);
    const float r2cap = (!raw && P.winner == nullptr) ? fminf(s_zero, __fmul_rn(FC_EXP_MINUS1, inv_k)) : s_zero;
    const int cnt = (int)npts;
    float* xyf = reinterpret_cast<float*>(s_xy2); float* cf = reinterpret_cast<float*>(s_c2);
    for (int t = tid; t < cnt + (cnt & 1); t += FT_THREADS) {     // (the pad behind an odd count: a record that can never win)
        float xr = 0.0f, yr = 0.0f, cv = INFINITY;
        if (t < cnt) {
            const float* q = xyz_c + (size_t)(lo + t) * 3;
            float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
            if (posed) apply_pose(s_pose, x, y, z);
            xr = rot_row(s_R, x, y, z); yr = rot_row(s_R + 3, x, y, z);
            cv = __fmul_rn(logf(depth_weight_m(rot_row(s_R + 6, x, y, z), zmin, range, f16)), P.inv_negk);
        }
        xyf[((t >> 1) << 2) + (t & 1)] = xr; xyf[((t >> 1) << 2) + (t & 1) + 2] = yr; cf[t] = cv;
    }
    __syncthreads();
    linear_batch(s_xy2, s_c2, xyz_c, lo, s_R, s_pose, posed, zmin, range, P.negk, f16, cnt, w, r2cap, s_zero, inv_k,
                 P.imgs + img_off, P.winner ? P.winner + img_off : nullptr, raw, p_begin, p_end);
    if (hi <= lo && P.winner)                             // empty cloud: the background and winner -1, as the culled kernel does
        for (int p = p_begin + tid; p < p_end; p += FT_THREADS) P.winner[img_off + p] = -1;
}

}  // namespace scream
