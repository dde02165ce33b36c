// common.cuh -- shared device helpers of libscream_render.so (sm_100a).
//
// Arithmetic contract (mirrored scalar-for-scalar by oracle/render_oracle.c):
// every step that the reference performs as a separate ATen kernel is rounded to fp32 separately
// here, with explicit __f*_rn intrinsics (never contracted into FMA, independent of -fmad); the
// rotation is the FMA chain of MKL sgemm, fmaf(r2, z, fmaf(r1, y, r0 * x)).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <math.h>

namespace scream {

// ---- order-preserving float <-> uint encoding (for integer min/max atomics / redux.sync) -------
// NaN encodes to the top of the order (explicitly: ptxas turns b | 0x80000000 into FADD -|x|, which re-canonicalises a NaN
// to 0x7FFFFFFF, the middle of the order), so the maximum of a range that saw a NaN depth decodes to NaN -- torch.min /
// torch.max propagate NaN (render.py:30) and the forward kernels poison the images from it.
__device__ __forceinline__ unsigned enc_f32(float f) {
    unsigned b = __float_as_uint(f);
    const unsigned e = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    return (f != f) ? 0xFFFFFFFFu : e;
}
__device__ __forceinline__ float dec_f32(unsigned e) {
    unsigned b = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
    return __uint_as_float(b);
}

// ---- rotation row: render.py:68 (MKL sgemm op order for N >= 15 columns) ------------------------
__device__ __forceinline__ float rot_row(const float* __restrict__ r, float x, float y, float z) {
    return __fmaf_rn(r[2], z, __fmaf_rn(r[1], y, __fmul_rn(r[0], x)));
}

// caller-side rigid transform of the source cloud, train_3d_match.py:198-199:  fl(R p) then fl(+ t)
__device__ __forceinline__ void apply_pose(const float* __restrict__ P, float& x, float& y, float& z) {
    float a = __fadd_rn(rot_row(P + 0, x, y, z), P[3]);
    float b = __fadd_rn(rot_row(P + 4, x, y, z), P[7]);
    float c = __fadd_rn(rot_row(P + 8, x, y, z), P[11]);
    x = a; y = b; z = c;
}

// render.py:31  range = fl32( double(zmax) - double(zmin) )   (Python-double subtraction of .item()s)
__device__ __forceinline__ float z_range(float zmin, float zmax) {
    return __double2float_rn(__dsub_rn((double)zmax, (double)zmin));
}
// render.py:31  pv = 1 - (z - zmin) / range      (true fp32 division on the CPU reference)
__device__ __forceinline__ float depth_weight(float z, float zmin, float range) {
    return __fsub_rn(1.0f, __fdiv_rn(__fsub_rn(z, zmin), range));
}

// ---- fp16 (autocast) input mode, SURVEY 8 f2: train_kitti.py:154-169 wraps the generator in torch.cuda.amp.autocast(), so
// render.py:68's matmul runs in fp16 (rotated coordinates rounded to fp16) and render.py:31 is a chain of fp16 tensor (op)
// Python-scalar kernels, each computed in fp32 and rounded to fp16; from render.py:42 on the fp32 pix_xy promotes everything
// back to fp32.  Pinned against the reference executed under torch.autocast("cpu", dtype=float16) (oracle/make_golden.py).
__device__ __forceinline__ float round_f16(float v) { return __half2float(__float2half_rn(v)); }
__device__ __forceinline__ float depth_weight_f16(float z, float zmin, float range) {
    return round_f16(__fsub_rn(1.0f, round_f16(__fdiv_rn(round_f16(__fsub_rn(z, zmin)), range))));
}
__device__ __forceinline__ float depth_weight_m(float z, float zmin, float range, bool f16) {
    return f16 ? depth_weight_f16(z, zmin, range) : depth_weight(z, zmin, range);
}
// rotated coordinate of the fp16 matmul: fp16 operands, fp32 accumulation (the products are exact in fp32), fp16 result
__device__ __forceinline__ float rot_row_f16(const float* __restrict__ r, float x, float y, float z) {
    return round_f16(__fmaf_rn(round_f16(r[2]), round_f16(z), __fmaf_rn(round_f16(r[1]), round_f16(y),
                     __fmul_rn(round_f16(r[0]), round_f16(x)))));
}

// render.py:14-16  pixel centre ((j - w//2) + 0.5) / (w//2)  -- exact in fp32 (w/2 is a power of two
// times at most a 7-bit integer; all intermediate values are multiples of 0.5 below 2^24)
__device__ __forceinline__ float pixel_centre(int j, int half) {
    const float num = __fadd_rn(__fsub_rn((float)j, (float)half), 0.5f);
    // half a power of two (w = 64, 128, 256, ...): scaling by 2^-e is exact, so the product with the reciprocal -- exponent
    // arithmetic on the bits of (float)half -- has the bits of the IEEE quotient, without the division's ten instructions
    if ((half & (half - 1)) == 0) return __fmul_rn(num, __uint_as_float(0x7F000000u - __float_as_uint((float)half)));
    return __fdiv_rn(num, (float)half);
}

// render.py:42-43,46  value of one (point, pixel) pair.  negk = -(float)(rho*rho) / 2 ; since /2 is
// exact, fl(fl(-d2/2) * rho2) == fl(d2 * negk).
__device__ __forceinline__ float pair_d2(float x, float y, float cx, float cy) {
    float dx = __fsub_rn(x, cx), dy = __fsub_rn(y, cy);
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}
__device__ __forceinline__ float pair_value(float d2, float pv, float negk) {
    return __fmul_rn(pv, expf(__fmul_rn(d2, negk)));
}

// Conservative culling radius (squared) for a pixel whose current best value is `best`:
// a pair with d2 > cull_d2(best) has value strictly below best, because
//   value = fl(pv * expf(a)) <= expf(a) * (1 + 2^-23) (pv <= 1),  a = fl(d2 * negk) <= d2*negk*(1 - 2^-23),
// and we demand a < log(best) - 1e-4, a margin ~1000x larger than every rounding error involved.
// best <= 0 (nothing found yet) or non-finite  ->  +inf (no culling).
__device__ __forceinline__ float cull_d2(float best, float inv_negk) {
    if (!(best > 0.0f)) return INFINITY;
    float la = __fsub_rn(logf(best), 1e-4f);          // <= 0 - margin
    float t = __fmul_rn(la, inv_negk);                // >= 0   (inv_negk = 1/negk < 0)
    return __fmaf_rn(t, 1e-5f, t) + 1e-30f;           // inflate: safe side
}

struct ItemSpan { int64_t src0, tgt0, end; };
__device__ __forceinline__ ItemSpan item_span(const int64_t* __restrict__ offs, int b) {
    ItemSpan s; s.src0 = offs[2 * b]; s.tgt0 = offs[2 * b + 1]; s.end = offs[2 * b + 2]; return s;
}
// pair mode (one item, the two clouds in separate buffers, sizes by value): offs == nullptr
__device__ __forceinline__ ItemSpan item_span(const int64_t* __restrict__ offs, int b, int64_t pair_n, int64_t pair_m) {
    if (offs != nullptr) return item_span(offs, b);
    ItemSpan s; s.src0 = 0; s.tgt0 = pair_n; s.end = pair_n + pair_m; return s;
}

}  // namespace scream
