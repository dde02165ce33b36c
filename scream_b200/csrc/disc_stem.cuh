// disc_stem.cuh -- the consumer's first layer on the renderer's output (the discriminator half).
//   models/gan.py:39   sequence = [nn.Conv2d(input_nc, ndf, kernel_size=4, stride=2, padding=1), nn.LeakyReLU(0.2, True)]
//   loss.py:50-57      the generator step feeds the rendered views straight into it: g_loss = -mean(D(imgs))
// x [N,C,H,W] is the renderer's own layout ([B,V,2,w,w] = N images of C = 2 channels; no permute, no copy), weight
// [F,C,4,4], bias [F] (F = ndf = 64), y [N,F,H/2,W/2] = leaky_relu(conv(x) + bias, 0.2).  fp32 FFMA, accumulation in
// (c, ky, kx) order -- cuDNN / oneDNN do not define theirs, so parity with the reference is a tolerance (1e-5 relative),
// stated in tests/test_disc_stem.py.  Zero padding is implicit (out-of-image taps contribute nothing).
//   forward   thread per output pixel, the 2 x 4 x 4 window in registers, all F filters from a shared-memory copy of the
//             weights (broadcast LDS.128); stores are coalesced per filter plane.  Output-write bound: 4 F bytes per pixel.
//   backward  dpre = dy * (y > 0 ? 1 : 0.2)  (the reference's LeakyReLU is in-place: autograd differentiates from the OUTPUT)
//             dx    CTA = 8 x 64 input pixels and the 6 x 34 output positions that see them, dpre staged in shared memory
//                   32 filters at a time; thread = two neighbouring pixels
//             dW,db persistent CTAs over chunks of 64 output positions: dpre and the im2col patch in shared memory, thread =
//                   4 filters x 2 taps, four positions per vector load; atomicAdd of the CTA's sums at the end
#pragma once
#include "common.cuh"

namespace scream {

constexpr int DS_THREADS = 256;
constexpr int DS_K = 4, DS_TAPS = DS_K * DS_K;            // kernel 4 x 4, stride 2, padding 1 (gan.py:38-39)
constexpr int DS_MAX_F = 64, DS_MAX_C = 2;
constexpr float DS_SLOPE = 0.2f;

struct DiscStemParams {
    const float* x; const float* weight; const float* bias;
    float* y;                                             // forward output / backward: the saved output
    const float* dy; float* dx; float* dw; float* db;     // backward
    int N, C, F, H, W;                                    // H, W even
};

__global__ void __launch_bounds__(DS_THREADS)
disc_stem_forward_kernel(DiscStemParams P)
{
    __shared__ __align__(16) float s_w[DS_MAX_F * DS_MAX_C * DS_TAPS];
    __shared__ float s_b[DS_MAX_F];
    const int taps = P.C * DS_TAPS;
    for (int t = threadIdx.x; t < P.F * taps; t += DS_THREADS) s_w[t] = P.weight[t];
    for (int t = threadIdx.x; t < P.F; t += DS_THREADS) s_b[t] = P.bias ? P.bias[t] : 0.0f;
    __syncthreads();
    const int Ho = P.H / 2, Wo = P.W / 2;
    const long long total = (long long)P.N * Ho * Wo;
    const long long t = (long long)blockIdx.x * DS_THREADS + threadIdx.x;
    if (t >= total) return;
    const int ox = (int)(t % Wo), oy = (int)((t / Wo) % Ho);
    const long long n = t / ((long long)Wo * Ho);
    float win[DS_MAX_C * DS_TAPS];
#pragma unroll
    for (int c = 0; c < DS_MAX_C; ++c)
#pragma unroll
        for (int ky = 0; ky < DS_K; ++ky)
#pragma unroll
            for (int kx = 0; kx < DS_K; ++kx) {
                const int iy = 2 * oy - 1 + ky, ix = 2 * ox - 1 + kx;
                float v = 0.0f;
                if (c < P.C && iy >= 0 && iy < P.H && ix >= 0 && ix < P.W)
                    v = __ldg(P.x + ((n * P.C + c) * P.H + iy) * P.W + ix);
                win[(c * DS_K + ky) * DS_K + kx] = v;
            }
    float* out = P.y + (n * P.F * Ho + oy) * Wo + ox;
    for (int f = 0; f < P.F; ++f) {
        const float4* wf = reinterpret_cast<const float4*>(s_w + f * taps);
        float acc = 0.0f;
#pragma unroll
        for (int q = 0; q < DS_MAX_C * DS_TAPS / 4; ++q) {
            if (q * 4 < taps) {
                const float4 w4 = wf[q];
                acc = __fmaf_rn(win[4 * q], w4.x, acc); acc = __fmaf_rn(win[4 * q + 1], w4.y, acc);
                acc = __fmaf_rn(win[4 * q + 2], w4.z, acc); acc = __fmaf_rn(win[4 * q + 3], w4.w, acc);
            }
        }
        acc = __fadd_rn(acc, s_b[f]);
        out[(long long)f * Ho * Wo] = acc > 0.0f ? acc : __fmul_rn(acc, DS_SLOPE);
    }
}

// dpre of one output element (0 outside the output): the in-place LeakyReLU differentiates from the saved output
__device__ __forceinline__ float ds_dpre(const DiscStemParams& P, long long n, int f, int oy, int ox, int Ho, int Wo) {
    if (oy < 0 || oy >= Ho || ox < 0 || ox >= Wo) return 0.0f;
    const long long o = ((n * P.F + f) * Ho + oy) * Wo + ox;
    return __ldg(P.dy + o) * (__ldg(P.y + o) > 0.0f ? 1.0f : DS_SLOPE);
}

// dx[n,c,iy,ix] = sum over the output positions (oy, ox) with 2 oy - 1 + ky == iy, 2 ox - 1 + kx == ix and all filters.
// grid (ceil(W / 64), ceil(H / 8), N); CTA = 8 input rows x 64 input columns = the 6 x 34 output positions that see them, 32
// filters at a time in shared memory (each dpre element is read from global memory once per tile instead of once per tap);
// thread = two neighbouring pixels of a row (even, odd column): 3 x 2 output positions, 16 FMAs per filter for 6 scalar and
// 4 vector shared loads.
#ifndef DS_DG_F_N
#define DS_DG_F_N 32
#endif
constexpr int DS_DG_ROWS = 8, DS_DG_COLS = 64, DS_DG_OR = DS_DG_ROWS / 2 + 2, DS_DG_OC = DS_DG_COLS / 2 + 2, DS_DG_F = DS_DG_F_N;
__global__ void __launch_bounds__(DS_THREADS)
disc_stem_dgrad_kernel(DiscStemParams P)
{
    __shared__ float s_d[DS_DG_F][DS_DG_OR][DS_DG_OC + 1];
    __shared__ __align__(16) float s_w[DS_DG_F * DS_MAX_C * DS_TAPS];
    const int Ho = P.H / 2, Wo = P.W / 2, taps = P.C * DS_TAPS;
    const int ix0 = blockIdx.x * DS_DG_COLS, iy0 = blockIdx.y * DS_DG_ROWS;
    const long long n = blockIdx.z;
    const int oyb = iy0 / 2 - 1, oxb = ix0 / 2 - 1;       // output position of the tile's corner
    const int r = threadIdx.x / 32, cp = threadIdx.x % 32;
    const int iy = iy0 + r, ix = ix0 + 2 * cp;
    // the two output rows that see input row iy: oy_a = (iy + 1) / 2 with ky_a = iy + 1 - 2 oy_a, and oy_a - 1 with ky_a + 2
    const int oy_a = (iy + 1) / 2, ky_a = iy + 1 - 2 * oy_a, la = oy_a - oyb;
    float acc[2][DS_MAX_C] = {{0.0f, 0.0f}, {0.0f, 0.0f}};      // [pixel ix, ix + 1][channel]
    for (int f0 = 0; f0 < P.F; f0 += DS_DG_F) {
        const int nf = min(DS_DG_F, P.F - f0);
        __syncthreads();
        for (int t = threadIdx.x; t < DS_DG_F * DS_DG_OR * DS_DG_OC; t += DS_THREADS) {
            const int ff = t / (DS_DG_OR * DS_DG_OC), rr = (t / DS_DG_OC) % DS_DG_OR, cc = t % DS_DG_OC;
            s_d[ff][rr][cc] = ff < nf ? ds_dpre(P, n, f0 + ff, oyb + rr, oxb + cc, Ho, Wo) : 0.0f;
        }
        for (int t = threadIdx.x; t < DS_DG_F * DS_MAX_C * DS_TAPS; t += DS_THREADS) {
            const int ff = t / (DS_MAX_C * DS_TAPS), cj = t % (DS_MAX_C * DS_TAPS);
            s_w[t] = (ff < nf && cj < taps) ? __ldg(P.weight + (size_t)(f0 + ff) * taps + cj) : 0.0f;
        }
        __syncthreads();
#pragma unroll 4
        for (int ff = 0; ff < DS_DG_F; ++ff) {
            // columns cp, cp + 1, cp + 2 of the tile = ox = ix/2 - 1, ix/2, ix/2 + 1
            const float a0 = s_d[ff][la][cp], a1 = s_d[ff][la][cp + 1], a2 = s_d[ff][la][cp + 2];
            const float b0 = s_d[ff][la - 1][cp], b1 = s_d[ff][la - 1][cp + 1], b2 = s_d[ff][la - 1][cp + 2];
#pragma unroll
            for (int c = 0; c < DS_MAX_C; ++c) {
                const float4 wa = *reinterpret_cast<const float4*>(s_w + (ff * DS_MAX_C + c) * DS_TAPS + ky_a * DS_K);
                const float4 wb = *reinterpret_cast<const float4*>(s_w + (ff * DS_MAX_C + c) * DS_TAPS + (ky_a + 2) * DS_K);
                // pixel ix (even): kx = 1 with ox = ix/2, kx = 3 with ox = ix/2 - 1;  pixel ix + 1: kx = 0 with ix/2 + 1, kx = 2 with ix/2
                acc[0][c] = __fmaf_rn(a1, wa.y, acc[0][c]); acc[0][c] = __fmaf_rn(a0, wa.w, acc[0][c]);
                acc[0][c] = __fmaf_rn(b1, wb.y, acc[0][c]); acc[0][c] = __fmaf_rn(b0, wb.w, acc[0][c]);
                acc[1][c] = __fmaf_rn(a2, wa.x, acc[1][c]); acc[1][c] = __fmaf_rn(a1, wa.z, acc[1][c]);
                acc[1][c] = __fmaf_rn(b2, wb.x, acc[1][c]); acc[1][c] = __fmaf_rn(b1, wb.z, acc[1][c]);
            }
        }
    }
    if (iy < P.H) {
#pragma unroll
        for (int c = 0; c < DS_MAX_C; ++c) {
            if (c < P.C) {
                float* o = P.dx + ((n * P.C + c) * P.H + iy) * P.W + ix;
                if (ix < P.W) o[0] = acc[0][c];
                if (ix + 1 < P.W) o[1] = acc[1][c];
            }
        }
    }
}

// dW[f,c,ky,kx] = sum_p dpre[f,p] * x[c, 2 oy - 1 + ky, 2 ox - 1 + kx],  db[f] = sum_p dpre[f,p]  over all output positions p:
// a [F x P] x [P x 32] product.  Persistent CTAs walk chunks of DS_WG_POS positions; per chunk dpre [F][pos] and the im2col
// patch [tap][pos] go to shared memory (y and dy are read from global memory once), thread = 4 filters x 2 taps over the
// positions four at a time (6 vector loads for 32 FMAs); the CTA's sums leave through atomicAdd at the end.
#ifndef DS_WG_POS_N
#define DS_WG_POS_N 32
#endif
#ifndef DS_WG_MINB_N
#define DS_WG_MINB_N 8
#endif
constexpr int DS_WG_MINB = DS_WG_MINB_N;                  // CTAs per SM the register allocation aims at (32-position chunks)
constexpr int DS_WG_POS = DS_WG_POS_N, DS_WG_LD = DS_WG_POS + 4;
constexpr size_t DS_WG_SMEM = (size_t)(DS_MAX_F + DS_MAX_C * DS_TAPS) * DS_WG_LD * sizeof(float) + (size_t)DS_WG_POS * sizeof(int4);
__global__ void __launch_bounds__(DS_THREADS, (DS_WG_POS <= 32 ? DS_WG_MINB : DS_WG_POS <= 64 ? 3 : 2))
disc_stem_wgrad_kernel(DiscStemParams P, long long chunks_per_cta)
{
    extern __shared__ __align__(16) unsigned char s_wg_raw[];
    float (*s_g)[DS_WG_LD] = reinterpret_cast<float (*)[DS_WG_LD]>(s_wg_raw);                                   // [F][pos] dpre of the chunk
    float (*s_x)[DS_WG_LD] = reinterpret_cast<float (*)[DS_WG_LD]>(s_wg_raw + (size_t)DS_MAX_F * DS_WG_LD * sizeof(float));   // [tap][pos] im2col patch
    int4* s_pos = reinterpret_cast<int4*>(s_wg_raw + (size_t)(DS_MAX_F + DS_MAX_C * DS_TAPS) * DS_WG_LD * sizeof(float)); // (image, offset in the plane, oy, ox)
    const int Ho = P.H / 2, Wo = P.W / 2, taps = P.C * DS_TAPS;
    const long long total = (long long)P.N * Ho * Wo;
    const int tf = threadIdx.x / 16, tt = threadIdx.x % 16;                  // filters 4 tf .. 4 tf + 3, taps tt and tt + 16
    float acc[4][2] = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}}, accb[4] = {0.f, 0.f, 0.f, 0.f};
    for (long long ch = 0; ch < chunks_per_cta; ++ch) {
        const long long p0 = ((long long)blockIdx.x * chunks_per_cta + ch) * DS_WG_POS;
        if (p0 >= total) break;
        __syncthreads();
        // where the chunk's positions are: image, offset inside the image's output plane, (oy, ox) -- once per position, not
        // once per element (the 64-bit divisions were most of the fill)
        for (int t = threadIdx.x; t < DS_WG_POS; t += DS_THREADS) {
            const long long p = p0 + t;
            int4 w4 = make_int4(-1, 0, 0, 0);
            if (p < total) {
                const long long n = p / ((long long)Wo * Ho);
                const int rr = (int)(p - n * (long long)Wo * Ho);
                w4 = make_int4((int)n, rr, rr / Wo, rr % Wo);
            }
            s_pos[t] = w4;
        }
        __syncthreads();
        {
            // A thread's elements all belong to the same positions (q = thread % positions), so their place is read once; its
            // loads are issued in groups of up to 16 rows before the group's shared stores (a store between two loads would
            // order them one behind the other).  A warp reads 128 contiguous bytes of one filter's plane per load; the
            // larger the chunk, the longer the contiguous piece per plane (the planes of one position are 4 KB apart).
            constexpr int QPT = DS_WG_POS > DS_THREADS ? DS_WG_POS / DS_THREADS : 1;          // positions per thread
            constexpr int FSTEP = DS_WG_POS >= DS_THREADS ? 1 : DS_THREADS / DS_WG_POS;       // rows between a thread's elements
            constexpr int NG = DS_MAX_F / FSTEP, NX = DS_MAX_C * DS_TAPS / FSTEP, GRP = NG < 16 ? NG : 16;
            static_assert(QPT == 1, "chunks of up to 256 positions");
            const int q = threadIdx.x % DS_WG_POS, r0 = threadIdx.x / DS_WG_POS;
            const int4 w4 = s_pos[q];
#pragma unroll
            for (int g0 = 0; g0 < NG; g0 += GRP) {
                float gv[GRP], yv[GRP];
#pragma unroll
                for (int u = 0; u < GRP; ++u) {
                    const int ff = r0 + FSTEP * (g0 + u);
                    gv[u] = 0.0f; yv[u] = 0.0f;
                    if (w4.x >= 0 && ff < P.F) {
                        const long long o = ((long long)w4.x * P.F + ff) * (Ho * Wo) + w4.y;
                        gv[u] = __ldg(P.dy + o); yv[u] = __ldg(P.y + o);
                    }
                }
#pragma unroll
                for (int u = 0; u < GRP; ++u) s_g[r0 + FSTEP * (g0 + u)][q] = gv[u] * (yv[u] > 0.0f ? 1.0f : DS_SLOPE);
            }
#pragma unroll
            for (int g0 = 0; g0 < NX; g0 += (NX < 16 ? NX : 16)) {
                float xv[(NX < 16 ? NX : 16)];
#pragma unroll
                for (int u = 0; u < (NX < 16 ? NX : 16); ++u) {
                    const int j = r0 + FSTEP * (g0 + u);
                    xv[u] = 0.0f;
                    if (w4.x >= 0 && j < taps) {
                        const int c = j / DS_TAPS, ky = (j % DS_TAPS) / DS_K, kx = j % DS_K;
                        const int iy = 2 * w4.z - 1 + ky, ix = 2 * w4.w - 1 + kx;
                        if (iy >= 0 && iy < P.H && ix >= 0 && ix < P.W) xv[u] = __ldg(P.x + (((long long)w4.x * P.C + c) * P.H + iy) * P.W + ix);
                    }
                }
#pragma unroll
                for (int u = 0; u < (NX < 16 ? NX : 16); ++u) s_x[r0 + FSTEP * (g0 + u)][q] = xv[u];
            }
        }
        __syncthreads();
#pragma unroll 2
        for (int q = 0; q < DS_WG_POS; q += 4) {
            const float4 x0 = *reinterpret_cast<const float4*>(&s_x[tt][q]), x1 = *reinterpret_cast<const float4*>(&s_x[tt + 16][q]);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 g = *reinterpret_cast<const float4*>(&s_g[4 * tf + u][q]);
                acc[u][0] = __fmaf_rn(g.x, x0.x, acc[u][0]); acc[u][0] = __fmaf_rn(g.y, x0.y, acc[u][0]);
                acc[u][0] = __fmaf_rn(g.z, x0.z, acc[u][0]); acc[u][0] = __fmaf_rn(g.w, x0.w, acc[u][0]);
                acc[u][1] = __fmaf_rn(g.x, x1.x, acc[u][1]); acc[u][1] = __fmaf_rn(g.y, x1.y, acc[u][1]);
                acc[u][1] = __fmaf_rn(g.z, x1.z, acc[u][1]); acc[u][1] = __fmaf_rn(g.w, x1.w, acc[u][1]);
                if (tt == 0) accb[u] += (g.x + g.y) + (g.z + g.w);
            }
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int f = 4 * tf + u;
        if (f < P.F) {
            if (tt < taps) atomicAdd(P.dw + f * taps + tt, acc[u][0]);
            if (tt + 16 < taps) atomicAdd(P.dw + f * taps + tt + 16, acc[u][1]);
            if (tt == 0 && P.db) atomicAdd(P.db + f, accb[u]);
        }
    }
}

}  // namespace scream
