/* scream_render.h -- C ABI of libscream_render.so (B200 / sm_100a scene renderer).
 *
 * Drop-in boundary for ONE hot path of xujiabo/SCREAM: the scene renderer
 *   /root/reference/models/render.py:8-73  (class RegistrationRender: __init__, render, forward)
 * and the autograd backward it implies (no source lines; defined by torch.max / exp / matmul at
 * render.py:31,42-49,68-69).  The reference has no FFI -- its "operator API" is that nn.Module --
 * so each entry point below cites the reference lines it replaces; the Python module
 * scream_b200/render.py binds them through ctypes and keeps the reference's call signature.
 *
 * Conventions
 *   - every function returns 0 on success and a negative SCREAM_E_* code on failure; nothing throws,
 *     nothing calls exit(); scream_last_error() returns a thread-local message for the last failure.
 *   - all data pointers are DEVICE pointers unless the name ends in _host.
 *   - the caller owns every buffer.  The library allocates nothing; scratch space is a caller-provided
 *     workspace whose size comes from scream_render_workspace_bytes().
 *   - all work is enqueued on the given stream; no call synchronises the device or the host
 *     (the *_host convenience entry points are the exception and say so).
 *   - thread-safe for concurrent calls that use distinct streams and distinct workspaces.
 *
 * Layout
 *   xyz    [total_pts,3] fp32 row-major (12 B per point, the layout of the reference's [n,3] tensors).
 *   offs   [2B+1] int64 CSR offsets into xyz, in points: item b owns the SOURCE cloud
 *          [offs[2b], offs[2b+1]) and the TARGET cloud [offs[2b+1], offs[2b+2]).  This is
 *          torch.cat([src_pred, tgt_pcd]) of render.py:63 for every item of a batch.
 *   pose   [B,3,4] fp32 [R|t] or NULL.  Applied to the SOURCE cloud only, before the view rotation:
 *          the caller-side rigid transform of the "real" render, train_3d_match.py:198-199,
 *          train_kitti.py:193-194 (fl(sgemm-chain) then a separately rounded + t).
 *   views  [V,3,3] fp32 view rotations = torch.Tensor(Rotation.from_euler('zyx', e).as_matrix()),
 *          render.py:67 (built on the host by the Python module with the same scipy call).
 *   imgs   [B,V,2,w,w] fp32: channel 0 = source cloud, 1 = target cloud (render.py:46,49,59).
 *   winner [B,V,2,w,w] int32: cloud-local index of the arg-max point of every pixel (lowest index on
 *          ties, 0 for an all-zero pixel -- torch.max(dim=0) semantics), or NULL.
 *   zrange [B,V,2] fp32: (zmin, zmax) of the rotated depth over BOTH clouds of the item (render.py:30-31).
 */
#ifndef SCREAM_RENDER_H_
#define SCREAM_RENDER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCREAM_ABI_VERSION 9

/* error codes */
#define SCREAM_OK            0
#define SCREAM_E_ARG        -1   /* bad argument (NULL pointer, w % 64 != 0, B/V <= 0, ...) */
#define SCREAM_E_WORKSPACE  -2   /* workspace too small or NULL */
#define SCREAM_E_CUDA       -3   /* a CUDA runtime call or launch failed; see scream_last_error() */
#define SCREAM_E_UNSUPPORTED -4  /* a size beyond what this build handles */

/* flags */
#define SCREAM_F_RAW        1u   /* imgs = un-normalised max in [0,1] (render.py:59, RegistrationRender.render)
                                    instead of (img-0.5)/0.5 (render.py:69, RegistrationRender.forward) */
#define SCREAM_F_DENSE      2u   /* force the dense exact kernel (every point x every pixel); default is the
                                    culled exact kernel, which returns bit-identical results */

#define SCREAM_F_ZRANGE_IN  4u   /* zrange [B,V,2] is an INPUT: the depth range of render.py:30-31 was reduced by the
                                    caller (point-sharded clouds: every shard must normalise with the range of
                                    the WHOLE cloud, scream_b200/shard.py); the z-range kernel is skipped */

#define SCREAM_F_NOSTREAM   8u   /* giant clouds (average item >= 147456 points at w == 64, >= 1048576 at w == 128): skip the streaming pre-filter and
                                    render every batch of the cloud (testing / diagnostics; same bits, slower) */

#define SCREAM_F_THINSEED  16u   /* giant clouds: seed the pre-filter from every 4th chunk only and always run its second
                                    round, as clouds of >= 16 M points do by themselves (testing; same bits) */
#define SCREAM_F_DUAL      32u   /* one launch renders every item twice (train_3d_match.py:186-201 renders the
                                   "fake" pair and the "real" pair, which share the target cloud): once with the clouds as
                                   given and once with the source posed.  pose must not be NULL; imgs / winner / zrange get
                                   a leading dimension of 2 ([2][B][V]...: 0 = as given, 1 = posed); the workspace query
                                   is unchanged.  The target cloud is sorted once and searched twice.  Not combined
                                   with SCREAM_F_DENSE or SCREAM_F_ZRANGE_IN, nor with giant clouds (SCREAM_E_UNSUPPORTED). */

#define SCREAM_F_DUAL_NOWIN1 128u /* with SCREAM_F_DUAL: no winner map for the posed render (winner is [B,V,2,w,w], the render as
                                   given only) -- nothing differentiates through the "real" render, and without a map its
                                   pixels below 2^-26 need no search */

#define SCREAM_F_HALF      64u   /* fp16 (autocast) input mode, SURVEY 8 f2 (train_kitti.py:154-169: under torch.cuda.amp.autocast()
                                   render.py:68's matmul and render.py:31 run in fp16): xyz holds clouds already rotated and
                                   rounded by scream_half_rotate, views is the identity, and the depth weight follows the
                                   fp16 chain h(1 - h(h(z - zmin) / range)).  No pose, no SCREAM_F_DUAL; giant clouds take
                                   the plain culled kernel.  Pass the same flag to scream_render_backward. */

/* opaque CUDA stream handle (cudaStream_t / CUstream) */
typedef void *scream_stream_t;

int scream_abi_version(void);
const char *scream_last_error(void);

/* Bytes of scratch the forward/backward calls need for a batch of B items, V views, image side w and
 * total_pts points in all.  (No reference counterpart: the reference lets torch allocate N x 4096
 * temporaries, render.py:42.)  16 bytes per (item, view) for ordinary batches; giant clouds -- total_pts >= 147456 * B at
 * w == 64, >= 1048576 * B at w == 128 -- take an exact pre-filter (seed images, bound tables, survivor lists) whose scratch is
 * included: query with the SAME (B, V, w, total_pts) as the forward call. */
int scream_render_workspace_bytes(int B, int V, int w, int64_t total_pts, size_t *out_bytes);

/* Forward render of a batch: replaces RegistrationRender.forward / .render, render.py:27-73, for B
 * independent (source, target) pairs in one launch sequence.
 *   rho: the reference's rho (render.py:10,43); the weight is exp(-d2/2 * (float)(rho*rho)).
 *   w:   image side, w % 64 == 0 (render.py:33-34,55).
 *   winner may be NULL (inference / "real" render); zrange may be NULL (output), or, with
 *   SCREAM_F_ZRANGE_IN, must hold the (zmin, zmax) pairs to use.
 *   max_item_pts: an upper bound of n+m over the items (sizes the z-range grid without reading offs on
 *   the host); 0 = unknown, total_pts is used.
 */
int scream_render_forward(const float *xyz, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                          const float *pose, const float *views,
                          int B, int V, int w, double rho, unsigned flags,
                          float *imgs, int32_t *winner, float *zrange,
                          void *ws, size_t ws_bytes, scream_stream_t stream);

/* scream_render_forward without a workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, flags, outputs and bits as scream_render_forward; the scratch (the encoded
 * depth ranges of render.py:30-31) is a static device array of SCREAM_NOALLOC_WS_BYTES inside the library, which every GPU
 * that loads the library holds until exit.  Limits: the call's workspace (scream_render_workspace_bytes) must fit, i.e.
 * B*V <= SCREAM_NOALLOC_WS_BYTES / 16 = 4096 (item, view) pairs and fewer than 147456 points per item on average (the
 * giant-cloud pre-filter needs hundreds of MB: use scream_render_forward); beyond that SCREAM_E_UNSUPPORTED, nothing launched.
 * The function calls no cudaMalloc / cudaFree / allocating routine.  ALL calls on a device share the one array (it keeps the
 * last call's contents; every word a call reads is rewritten by that call first): calls that overlap on the GPU -- two
 * streams, two threads, two processes' worth of streams in one context -- RACE; serialise them on one stream. */
#define SCREAM_NOALLOC_WS_BYTES 65536
int scream_render_forward_noalloc(const float *xyz, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                                  const float *pose, const float *views,
                                  int B, int V, int w, double rho, unsigned flags,
                                  float *imgs, int32_t *winner, float *zrange, scream_stream_t stream);

/* Backward: replaces the autograd graph of render.py:27-73.  Scatters
 * grad_imgs [B,V,2,w,w] through the saved winner map to grad_xyz [total_pts,3] (overwritten: the
 * callee zero-fills it first).  With a pose, the source gradient is w.r.t. the PRE-pose coordinates.
 * flags must be the forward call's (SCREAM_F_RAW changes the 1/0.5 output scale). */
int scream_render_backward(const float *xyz, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                           const float *pose, const float *views,
                           int B, int V, int w, double rho, unsigned flags,
                           const float *grad_imgs, const int32_t *winner, const float *zrange,
                           float *grad_xyz,
                           void *ws, size_t ws_bytes, scream_stream_t stream);

/* One pair per call -- the reference's own call pattern: RegistrationRender.forward(src_pred, tgt_pcd), render.py:61-73, is
 * handed the two clouds as two tensors and pointnet.py:39-40 asserts batch 1.  src [n,3], tgt [m,3] are separate device
 * buffers (no torch.cat of render.py:63, no offsets); pose [3,4] or NULL; imgs [V,2,w,w], winner [V,2,w,w] or NULL, zrange
 * [V,2] or NULL (output).  ONE kernel launch: every CTA reduces the depth range of render.py:30-31 itself, and each image is
 * shared by as many CTAs as fill one wave of the SMs (a 6-view pair: 12 per image, 144 SMs instead of 12).  Same bits as scream_render_forward on the
 * concatenated pair.  flags: SCREAM_F_RAW, SCREAM_F_HALF.  n + m < 147456 (giant pairs: scream_render_forward). */
int scream_render_forward_pair(const float *src, int64_t n, const float *tgt, int64_t m,
                               const float *pose, const float *views, int V, int w, double rho, unsigned flags,
                               float *imgs, int32_t *winner, float *zrange, scream_stream_t stream);

/* ... and its backward: grad_src [n,3], grad_tgt [m,3] (overwritten), the gradient w.r.t. both inputs as the reference's
 * autograd returns it (render.py:63's cat routes it to the two tensors). */
int scream_render_backward_pair(const float *src, int64_t n, const float *tgt, int64_t m,
                                const float *pose, const float *views, int V, int w, double rho, unsigned flags,
                                const float *grad_imgs, const int32_t *winner, const float *zrange,
                                float *grad_src, float *grad_tgt, scream_stream_t stream);

/* fp16 (autocast) input mode, stage 1 (replaces the fp16 torch.matmul of render.py:68 under
 * train_kitti.py:154's autocast): every item's clouds rotated by every view with fp16 operands, fp32 accumulation and an
 * fp16 result, stored as fp32 words.  xr [V*total_pts,3]: item b, view v at rows V*offs[2b] + v*(n_b+m_b) ...;
 * offs_out [2*B*V+1]: CSR offsets presenting xr as B*V items (index b*V+v) for scream_render_forward(..., B*V items,
 * one identity view, flags | SCREAM_F_HALF), whose [B*V,1,2,w,w] outputs are the [B,V,2,w,w] outputs. */
int scream_half_rotate(const float *xyz, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                       const float *views, int B, int V, float *xr, int64_t *offs_out, scream_stream_t stream);

/* ... and its gradient: grad_xyz[i] = sum_v fp16(R_v)^T grad_xr[v][i] (fp32; the casts pass gradients straight through,
 * as autograd does for .half()).  grad_xr in xr's layout = scream_render_backward's grad_xyz of the B*V-item call. */
int scream_half_rotate_backward(const float *grad_xr, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                                const float *views, int B, int V, float *grad_xyz, scream_stream_t stream);

/* Nearest target of every (predicted) source point: replaces square_distance(src_ / s, tgt / s)[0].min(dim=1),
 * pointnet.py:71 + utils.py:72-78 (the reference materialises the n x m matrix).  src [n,3], tgt [m,3] device fp32,
 * s the scale both clouds are divided by; dist [n] fp32 and ind [n] int64 (either may be NULL), bit-identical to the
 * reference (first index on ties).  ws: n * 8 bytes. */
int scream_nearest_target(const float *src, int64_t n, const float *tgt, int64_t m, float s,
                          float *dist, int64_t *ind, void *ws, size_t ws_bytes, scream_stream_t stream);

/* Weighted Kabsch fit: replaces rigid_transform_3d, utils.py:138-178, including its CPU SVD (utils.py:169) -- no host
 * round trip.  A [bs,k,3]; correspondence i of item b pairs A[b,i] with B[b, b_index ? b_index[b,i] : i], B [bs,b_rows,3];
 * weights [bs,k] or NULL (ones); weights below weight_threshold count 0 (utils.py:151); valid_dist [bs,k] or NULL:
 * correspondences with !(valid_dist < valid_thr) count 0 (the boolean mask of pointnet.py:72-74 as a weight, which
 * spares the mask's device->host sync).  T [bs,4,4] row-major homogeneous transforms.  ws: bs * 128 bytes. */
int scream_rigid_transform(const float *A, const float *B, const int64_t *b_index, const float *weights,
                           const float *valid_dist, float valid_thr, float weight_threshold,
                           int bs, int64_t k, int64_t b_rows, float *T, void *ws, size_t ws_bytes, scream_stream_t stream);

/* Point-sharded giant clouds (BASELINE config 5): the exact cross-rank merge of render.py:46,49's max with first
 * index on ties.  val [V,2,w,w] un-normalised values and winner [V,2,w,w] cloud-local indices of THIS rank's slice (the
 * output of scream_render_forward with SCREAM_F_RAW | SCREAM_F_ZRANGE_IN); src_offset / tgt_offset: global index of the
 * slice's first source / target point.  keys [V,2,w,w] int64 = (value bits << 32) | (0xFFFFFFFF - global index), written
 * straight into the buffer of the all_reduce(MAX); scream_unpack_keys turns the reduced keys into val, imgs = (val-0.5)/0.5
 * (NULL to skip) and the global winner map -- bit-identical to the single-GPU render. */
/* The forward render of ONE RANK's slice of a point-sharded giant cloud, in three stages with a collective between them, so
 * that the pre-filter of every rank bounds with the best seeds of ALL ranks (a slice's own seeds are only its local best:
 * its survivor lists, and everything behind them, then do not shrink with the slice).  zrange [B,V,2] is an INPUT (the range
 * of the whole cloud, SCREAM_F_ZRANGE_IN implied); the same workspace must be passed to the three calls.
 *   stage 1  depth range -> seed pass over this rank's share of the 4 M-point sample -> seed table.  Returns device pointers
 *            into the workspace: *seed_keys [n_keys] int64 (value bits << 32 | ~rank)  -> caller: all_reduce(MAX)
 *   stage 2  table entries that are not their cell's best are cleared; *seed_table [n_table_floats] fp32
 *                                                                                      -> caller: all_reduce(SUM)
 *   stage 3  bounds from the pooled seeds, voxel masks, filter, render of the survivors -> imgs / winner of this slice
 * Needs the giant-cloud path (w == 64, total_pts >= 147456 * B) on EVERY rank; flags: SCREAM_F_RAW. */
int scream_render_forward_sharded(const float *xyz, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                                  const float *views, int B, int V, int w, double rho, unsigned flags,
                                  int stage, int rank, int world,
                                  float *imgs, int32_t *winner, float *zrange,
                                  void *ws, size_t ws_bytes,
                                  int64_t **seed_keys, size_t *n_keys, float **seed_table, size_t *n_table_floats,
                                  scream_stream_t stream);

int scream_pack_keys(const float *val, const int32_t *winner, int64_t src_offset, int64_t tgt_offset, int V, int w,
                     int64_t *keys, scream_stream_t stream);
int scream_unpack_keys(const int64_t *keys, int V, int w, float *val, float *imgs, int64_t *winner_global,
                       scream_stream_t stream);

/* Transform stage only (testing / parity of the "index path"): rotated coordinates [B? no: total_pts
 * per view] rotated [V,total_pts,3], pv [V,total_pts] (render.py:31), zrange [B,V,2].  Any output may
 * be NULL.  Bit-exact against the reference for clouds of >= 16 points (sgemm op order). */
int scream_render_transform(const float *xyz, const int64_t *offs, int64_t total_pts, int64_t max_item_pts,
                            const float *pose, const float *views, int B, int V,
                            float *rotated, float *pv, float *zrange,
                            void *ws, size_t ws_bytes, scream_stream_t stream);

/* Depth images -> colour strip: replaces utils.py:81-91 (deep_to_img) applied to every view and the
 * np.concatenate(..., axis=1) of train_3d_match.py:212-213, train_kitti.py:211-212, visualize_depth_image.py:27-28.
 *   imgs [B,V,2,w,w] fp32, the forward call's normalised output;  rgb [B, w, V*w, 3] uint8: view v of item b occupies
 *   columns [v*w, (v+1)*w).  Component order is the reference's (it hands the array to cv2 as is).
 * The reference shows / writes the strip with cv2 (imshow is unavailable on this box); here it stays a device buffer. */
int scream_depth_to_rgb8(const float *imgs, int B, int V, int w, unsigned char *rgb, scream_stream_t stream);

/* The consumer's first layer on the renderer's output: replaces the first two modules of NLayerDiscriminator,
 * models/gan.py:39 -- nn.Conv2d(input_nc, ndf, kernel_size=4, stride=2, padding=1) + nn.LeakyReLU(0.2, True) -- which
 * loss.py:50-57 applies to the rendered views.  x [N,C,H,W] is the renderer's own [B,V,2,w,w] layout (N = B*V, C = 2; no
 * permute), weight [F,C,4,4], bias [F] or NULL, y [N,F,H/2,W/2].  C <= 2, F <= 64, H and W even.  fp32; the reference's
 * cuDNN / oneDNN accumulation order is unspecified, so parity is 1e-5 relative (tests/test_disc_stem.py). */
int scream_disc_stem_forward(const float *x, const float *weight, const float *bias, int N, int C, int F, int H, int W,
                             float *y, scream_stream_t stream);

/* ... and its backward (autograd of the two modules; the in-place LeakyReLU differentiates from the saved OUTPUT y):
 * dx [N,C,H,W], dw [F,C,4,4], db [F]; any of them may be NULL (db only together with dw). */
int scream_disc_stem_backward(const float *x, const float *weight, const float *y, const float *dy, int N, int C, int F, int H, int W,
                              float *dx, float *dw, float *db, scream_stream_t stream);

/* Host-buffer convenience entry point (end-to-end path: H2D copies, forward, D2H copies, one stream
 * synchronise).  All pointers are HOST pointers; device staging is allocated per call and freed.
 * device < 0 uses the current device.  flags: SCREAM_F_RAW, _DENSE, _NOSTREAM, _THINSEED; the others
 * (_DUAL, _ZRANGE_IN, _HALF change the buffer layout or make zrange an input) answer SCREAM_E_UNSUPPORTED. */
int scream_render_forward_host(const float *xyz_host, const int64_t *offs_host,
                               const float *pose_host, const float *views_host,
                               int B, int V, int w, double rho, unsigned flags,
                               float *imgs_host, int32_t *winner_host, float *zrange_host, int device);

#ifdef __cplusplus
}
#endif
#endif /* SCREAM_RENDER_H_ */
