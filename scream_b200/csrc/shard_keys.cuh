// shard_keys.cuh -- point-sharded giant clouds (BASELINE config 5): the exact merge of per-shard winners.
//   render.py:46,49  img = max over ALL points, first index on ties.  A rank that holds a slice of the cloud renders its
//   per-pixel (value, cloud-local winner); the pair is packed into ONE 64-bit key
//       (float bits of value << 32) | (0xFFFFFFFF - global index)
//   -- values are >= 0, so the integer order is the float order, and a LOWER index gives a LARGER key -- and one
//   all_reduce(MAX) over the ranks returns exactly the single-GPU image and winner map.  These two kernels are the
//   pack in front of the collective (straight into its buffer) and the unpack behind it.
#pragma once
#include "common.cuh"

namespace scream {

constexpr int SK_THREADS = 256;

// val / winner [V,2,npix] of one item; offs2 = global index of the first local source / target point; winner < 0 (empty
// local cloud) -> key 0, the smallest
__global__ void __launch_bounds__(SK_THREADS)
pack_keys_kernel(const float* __restrict__ val, const int32_t* __restrict__ winner, long long src_off, long long tgt_off,
                 int npix, long long total, long long* __restrict__ keys)
{
    const long long t = (long long)blockIdx.x * SK_THREADS + threadIdx.x;
    if (t >= total) return;
    const int c = (int)((t / npix) & 1);
    const int wi = winner[t];
    const unsigned long long gi = (unsigned long long)((c ? tgt_off : src_off) + wi);
    const unsigned long long key = ((unsigned long long)__float_as_uint(val[t]) << 32) | (0xFFFFFFFFull - (gi & 0xFFFFFFFFull));
    keys[t] = wi < 0 ? 0ll : (long long)key;
}

// keys -> val (raw), imgs = (val - 0.5) / 0.5 (render.py:69), winner_global int64; key 0 ("nothing anywhere") -> value 0, index 0
__global__ void __launch_bounds__(SK_THREADS)
unpack_keys_kernel(const long long* __restrict__ keys, long long total, float* __restrict__ val, float* __restrict__ imgs,
                   long long* __restrict__ winner_global)
{
    const long long t = (long long)blockIdx.x * SK_THREADS + threadIdx.x;
    if (t >= total) return;
    const unsigned long long k = (unsigned long long)keys[t];
    const float v = __uint_as_float((unsigned)(k >> 32));
    val[t] = v;
    if (imgs) imgs[t] = __fmul_rn(__fsub_rn(v, 0.5f), 2.0f);
    winner_global[t] = k == 0ull ? 0ll : (long long)(0xFFFFFFFFull - (k & 0xFFFFFFFFull));
}

}  // namespace scream
