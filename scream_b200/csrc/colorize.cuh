// colorize.cuh -- depth images -> RGB8 strip.
//   utils.py:81-91  deep_to_img(deep_img[w,w,2]):  d = x*0.5 + 0.5;  (img, ch) = max over the 2 channels (first
//                   channel on ties);  colour (1, 0.706, 0) where channel 1 won, (0, 0.651, 0.929) elsewhere;
//                   uint8(img * colour * 255)  (astype truncates toward zero)
//   train_3d_match.py:212-214 / train_kitti.py:211-213 / visualize_depth_image.py:27-28: the V views are
//                   concatenated along the image width and written with cv2.imwrite (a window with cv2.imshow in the
//                   viewer -- unavailable on this box; the headless product is this device buffer).
// The reference multiplies by the mask (1 or 0), then by the colour, adds the two (one term is 0), then * 255:
// per component  u8( fl( fl(img * col) * 255 ) ), every step rounded to fp32 -- restated with explicit _rn ops.
// Reads the renderer's own NCHW layout [B,V,2,w,w] (the reference permutes to NHWC first: same values).
// One thread per output pixel; HBM stream: 8 B in, 3 B out per pixel.
#pragma once
#include "common.cuh"

namespace scream {

constexpr int CZ_THREADS = 256;

__global__ void __launch_bounds__(CZ_THREADS)
depth_to_rgb8_kernel(const float* __restrict__ imgs, int B, int V, int w, unsigned char* __restrict__ rgb)
{
    const int npix = w * w;
    const long long total = (long long)B * V * npix;
    const long long t = (long long)blockIdx.x * CZ_THREADS + threadIdx.x;
    if (t >= total) return;
    const int p = (int)(t % npix);
    const int v = (int)((t / npix) % V);
    const int b = (int)(t / ((long long)npix * V));
    const float* src = imgs + ((size_t)b * V + v) * 2 * npix;
    const float d0 = __fadd_rn(__fmul_rn(src[p], 0.5f), 0.5f);            // channel 0 (source cloud)
    const float d1 = __fadd_rn(__fmul_rn(src[npix + p], 0.5f), 0.5f);     // channel 1 (target cloud)
    const bool one = d1 > d0;                                             // torch.max: first index on ties; NaN handled below
    float img = one ? d1 : d0;
    if (d0 != d0) { img = d0; }                                           // torch.max propagates NaN (first NaN wins)
    else if (d1 != d1) { img = d1; }
    const bool ch1 = (d0 == d0) && (one || d1 != d1);                     // index 1 chosen
    const float c0 = ch1 ? 1.0f : 0.0f, c1 = ch1 ? 0.706f : 0.651f, c2 = ch1 ? 0.0f : 0.929f;
    const int row = p / w, col = p % w;
    unsigned char* o = rgb + (((size_t)b * w + row) * ((size_t)V * w) + (size_t)v * w + col) * 3;
    // numpy astype(uint8) of a float32: C cast (truncation) for in-range values; NaN / out of range is UB in C --
    // clamp instead (the renderer's images are in [-1, 1), so img is in [0, 1)).
    const float r0 = __fmul_rn(__fmul_rn(img, c0), 255.0f), r1 = __fmul_rn(__fmul_rn(img, c1), 255.0f),
                r2 = __fmul_rn(__fmul_rn(img, c2), 255.0f);
    o[0] = (unsigned char)fminf(fmaxf(r0, 0.0f), 255.0f);
    o[1] = (unsigned char)fminf(fmaxf(r1, 0.0f), 255.0f);
    o[2] = (unsigned char)fminf(fmaxf(r2, 0.0f), 255.0f);
}

}  // namespace scream
