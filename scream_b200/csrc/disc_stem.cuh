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
//             dx    thread per input pixel: the 2 x 2 output positions that see it x F filters
//             dW,db one CTA per chunk of output positions: a 64 x 33 outer-product tile accumulated in registers
//                   (thread = 1 filter x 8 taps), then atomicAdd of the CTA's partial sums
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

// dx[n,c,iy,ix] = sum over the output positions (oy, ox) with 2 oy - 1 + ky == iy, 2 ox - 1 + kx == ix and all filters
__global__ void __launch_bounds__(DS_THREADS)
disc_stem_dgrad_kernel(DiscStemParams P)
{
    __shared__ float s_w[DS_MAX_F * DS_MAX_C * DS_TAPS];
    const int taps = P.C * DS_TAPS;
    for (int t = threadIdx.x; t < P.F * taps; t += DS_THREADS) s_w[t] = P.weight[t];
    __syncthreads();
    const int Ho = P.H / 2, Wo = P.W / 2;
    const long long total = (long long)P.N * P.H * P.W;
    const long long t = (long long)blockIdx.x * DS_THREADS + threadIdx.x;
    if (t >= total) return;
    const int ix = (int)(t % P.W), iy = (int)((t / P.W) % P.H);
    const long long n = t / ((long long)P.W * P.H);
    float acc[DS_MAX_C] = {0.0f, 0.0f};
    // ky = iy + 1 - 2 oy in [0, 4)  ->  oy in {(iy + 1) / 2, (iy + 1) / 2 - 1} (when inside the output)
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        const int oy = (iy + 1) / 2 - a, ky = iy + 1 - 2 * oy;
        if (oy < 0 || oy >= Ho) continue;
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int ox = (ix + 1) / 2 - b, kx = ix + 1 - 2 * ox;
            if (ox < 0 || ox >= Wo) continue;
            const float* yp = P.y + (n * P.F * Ho + oy) * Wo + ox;
            const float* gp = P.dy + (n * P.F * Ho + oy) * Wo + ox;
            for (int f = 0; f < P.F; ++f) {
                const float g = __ldg(gp + (long long)f * Ho * Wo) * (__ldg(yp + (long long)f * Ho * Wo) > 0.0f ? 1.0f : DS_SLOPE);
#pragma unroll
                for (int c = 0; c < DS_MAX_C; ++c)
                    if (c < P.C) acc[c] = __fmaf_rn(g, s_w[(f * P.C + c) * DS_TAPS + ky * DS_K + kx], acc[c]);
            }
        }
    }
    for (int c = 0; c < P.C; ++c) P.dx[((n * P.C + c) * P.H + iy) * P.W + ix] = acc[c];
}

// dW[f,c,ky,kx] = sum_p dpre[f,p] * x[c, 2 oy - 1 + ky, 2 ox - 1 + kx],  db[f] = sum_p dpre[f,p]  over all output positions p.
// CTA = DS_WG_POS consecutive output positions of one image row block; thread = filter f (tid / 4) x a quarter of the taps.
constexpr int DS_WG_POS = 64;
__global__ void __launch_bounds__(DS_THREADS)
disc_stem_wgrad_kernel(DiscStemParams P, long long chunks_per_cta)
{
    __shared__ float s_g[DS_MAX_F][DS_WG_POS + 1];        // dpre of the chunk
    __shared__ float s_x[DS_MAX_C * DS_TAPS][DS_WG_POS + 1];   // im2col patch of the chunk
    const int Ho = P.H / 2, Wo = P.W / 2, taps = P.C * DS_TAPS;
    const long long total = (long long)P.N * Ho * Wo;
    const int f = threadIdx.x >> 2, part = threadIdx.x & 3;          // 64 filters x 4 tap groups of 8
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, accb = 0.0f;
    for (long long ch = 0; ch < chunks_per_cta; ++ch) {
        const long long p0 = ((long long)blockIdx.x * chunks_per_cta + ch) * DS_WG_POS;
        if (p0 >= total) break;
        __syncthreads();
        for (int t = threadIdx.x; t < P.F * DS_WG_POS; t += DS_THREADS) {
            const int ff = t / DS_WG_POS, q = t % DS_WG_POS;
            const long long p = p0 + q;
            float g = 0.0f;
            if (p < total) {
                const long long n = p / ((long long)Wo * Ho), r = p % ((long long)Wo * Ho);
                const long long o = (n * P.F + ff) * Ho * Wo + r;
                g = __ldg(P.dy + o) * (__ldg(P.y + o) > 0.0f ? 1.0f : DS_SLOPE);
            }
            s_g[ff][q] = g;
        }
        for (int t = threadIdx.x; t < taps * DS_WG_POS; t += DS_THREADS) {
            const int j = t / DS_WG_POS, q = t % DS_WG_POS;
            const long long p = p0 + q;
            float v = 0.0f;
            if (p < total) {
                const int ox = (int)(p % Wo), oy = (int)((p / Wo) % Ho);
                const long long n = p / ((long long)Wo * Ho);
                const int c = j / DS_TAPS, ky = (j % DS_TAPS) / DS_K, kx = j % DS_K;
                const int iy = 2 * oy - 1 + ky, ix = 2 * ox - 1 + kx;
                if (iy >= 0 && iy < P.H && ix >= 0 && ix < P.W) v = __ldg(P.x + ((n * P.C + c) * P.H + iy) * P.W + ix);
            }
            s_x[j][q] = v;
        }
        __syncthreads();
        if (f < P.F) {
            for (int q = 0; q < DS_WG_POS; ++q) {
                const float g = s_g[f][q];
                if (part == 0) accb += g;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int j = part * 8 + u;
                    if (j < taps) acc[u] = __fmaf_rn(g, s_x[j][q], acc[u]);
                }
            }
        }
    }
    if (f < P.F) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int j = part * 8 + u;
            if (j < taps) atomicAdd(P.dw + f * taps + j, acc[u]);
        }
        if (part == 0 && P.db) atomicAdd(P.db + f, accb);
    }
}

}  // namespace scream
