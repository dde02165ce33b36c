// half_rotate.cuh -- fp16 (autocast) input mode of the renderer, SURVEY.md 8 f2.
//   train_kitti.py:154-169  the generator runs under torch.cuda.amp.autocast(): render.py:68's torch.matmul casts the view
//                           matrix and the [n+m,3] cloud to fp16 and returns fp16 rotated coordinates; render.py:31 then
//                           works on fp16 depths (common.cuh: depth_weight_f16); render.py:42 onwards is fp32 again.
// Layout: the rotated clouds are written once, item-major / view-minor, as fp32 words holding fp16 values
//   xr[(V * offs[2b] + v * len_b + l) * 3 + k],  len_b = n_b + m_b,  l = point index inside item b,
// with a CSR offs_out[2*B*V + 1] that presents them as B*V items of ONE (identity) view: the product kernels then run
// unchanged (a rotation by the identity is exact), with SCREAM_F_HALF switching the depth-weight chain, and their
// [B*V,1,2,w,w] outputs ARE the [B,V,2,w,w] outputs.  HBM: 12 B read + 12 V B written per point (forward), the reverse
// for the gradient.  One thread per point, all views in a loop; grid (chunks, B).
#pragma once
#include "common.cuh"

namespace scream {

constexpr int HR_THREADS = 256;

__global__ void __launch_bounds__(HR_THREADS)
half_rotate_kernel(const float* __restrict__ xyz, const int64_t* __restrict__ offs, const float* __restrict__ views,
                   int B, int V, float* __restrict__ xr, int64_t* __restrict__ offs_out)
{
    const int b = blockIdx.y;
    const ItemSpan sp = item_span(offs, b);
    const int64_t len = sp.end - sp.src0, out0 = (int64_t)V * sp.src0;
    if (blockIdx.x == 0) {
        for (int v = threadIdx.x; v < V; v += HR_THREADS) {
            const int64_t j = (int64_t)b * V + v;
            offs_out[2 * j] = out0 + v * len;
            offs_out[2 * j + 1] = out0 + v * len + (sp.tgt0 - sp.src0);
        }
        if (b == B - 1 && threadIdx.x == 0) offs_out[2 * (int64_t)B * V] = (int64_t)V * sp.end;
    }
    for (int64_t l = (int64_t)blockIdx.x * HR_THREADS + threadIdx.x; l < len; l += (int64_t)gridDim.x * HR_THREADS) {
        const float* q = xyz + (size_t)(sp.src0 + l) * 3;
        const float x = __ldg(q), y = __ldg(q + 1), z = __ldg(q + 2);
        for (int v = 0; v < V; ++v) {
            const float* R = views + (size_t)v * 9;
            float* o = xr + (size_t)(out0 + v * len + l) * 3;
            o[0] = rot_row_f16(R, x, y, z); o[1] = rot_row_f16(R + 3, x, y, z); o[2] = rot_row_f16(R + 6, x, y, z);
        }
    }
}

// gradient of the above with the casts passed straight through (autograd's own rule for .half()):
//   grad_xyz[i] = sum_v  fp16(R_v)^T  grad_xr[v][i],   fp32, views in ascending order (deterministic)
__global__ void __launch_bounds__(HR_THREADS)
half_rotate_backward_kernel(const float* __restrict__ grad_xr, const int64_t* __restrict__ offs,
                            const float* __restrict__ views, int V, float* __restrict__ grad_xyz)
{
    const int b = blockIdx.y;
    const ItemSpan sp = item_span(offs, b);
    const int64_t len = sp.end - sp.src0, out0 = (int64_t)V * sp.src0;
    for (int64_t l = (int64_t)blockIdx.x * HR_THREADS + threadIdx.x; l < len; l += (int64_t)gridDim.x * HR_THREADS) {
        float g0 = 0.f, g1 = 0.f, g2 = 0.f;
        for (int v = 0; v < V; ++v) {
            const float* R = views + (size_t)v * 9;
            const float* g = grad_xr + (size_t)(out0 + v * len + l) * 3;
            const float gx = __ldg(g), gy = __ldg(g + 1), gz = __ldg(g + 2);
            g0 += round_f16(R[0]) * gx + round_f16(R[3]) * gy + round_f16(R[6]) * gz;
            g1 += round_f16(R[1]) * gx + round_f16(R[4]) * gy + round_f16(R[7]) * gz;
            g2 += round_f16(R[2]) * gx + round_f16(R[5]) * gy + round_f16(R[8]) * gz;
        }
        float* o = grad_xyz + (size_t)(sp.src0 + l) * 3;
        o[0] = g0; o[1] = g1; o[2] = g2;
    }
}

}  // namespace scream
