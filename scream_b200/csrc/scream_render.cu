// scream_render.cu -- C ABI of libscream_render.so (see include/scream_render.h).
// Build: scream_b200/csrc/build.py  (nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false)
#include "../../include/scream_render.h"
#include "common.cuh"
#include "transform.cuh"
#include "forward_dense.cuh"
#include "forward_culled.cuh"
#include "forward_stream.cuh"
#include "backward.cuh"
#include "colorize.cuh"
#include "half_rotate.cuh"
#include "registration.cuh"
#include "shard_keys.cuh"
#include "disc_stem.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

using namespace scream;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                         \
    do { cudaError_t e_ = (expr);                                                              \
         if (e_ != cudaSuccess) return fail(SCREAM_E_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct WsLayout { size_t zenc_off, seed_off, thr_off, cnt_off, surv_off, surv2_off, vox_off, tab_off, total; bool stream; int surv_cap, surv2_cap, vox_g; };

// The giant-cloud pre-filter (forward_stream.cuh) is used when the AVERAGE item reaches FS_MIN_AVG_PTS points
// (decided from B and total_pts alone, so the workspace query and the forward call agree) and w == 64 -- or w == 128 and
// FS_MIN_AVG_PTS_W128 points: that image is four 64 x 64 pixel lattices, each filtered (a pass over the cloud) and rendered
// from its own survivor lists, one after the other.
constexpr int64_t FS_MIN_AVG_PTS_W128 = 1 << 20;
bool stream_eligible(int B, int w, int64_t total_pts) {
    return (w == 64 && total_pts >= FS_MIN_AVG_PTS * (int64_t)B) || (w == 128 && total_pts >= FS_MIN_AVG_PTS_W128 * (int64_t)B);
}

WsLayout ws_layout(int B, int V, int w, int64_t total_pts) {
    WsLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    L.zenc_off = take((size_t)B * V * 2 * 2 * sizeof(unsigned));        // two sets: SCREAM_F_DUAL keeps the range of both renders
    L.stream = stream_eligible(B, w, total_pts);
    L.seed_off = L.thr_off = L.cnt_off = L.surv_off = L.surv2_off = L.vox_off = L.tab_off = 0;
    L.vox_g = total_pts / B >= FS_VOX_BIG_MIN ? FV_G : FV_G / 2;   // the mask kernel costs g^3 x V; the finer grid pays from FS_VOX_BIG_MIN points on
    L.surv2_cap = 1 << 18;
    // rim cells of a solid cloud pass the cell-level test wholesale (~2-3 % of the points): keep room for 1/16
    L.surv_cap = (int)std::min<int64_t>(std::max<int64_t>(FS_SURV_MIN, total_pts / 16), 1 << 26);
    if (L.stream) {
        const size_t imgs = (size_t)B * V * 2;
        L.seed_off = take(imgs * FC_CELLS * 2 * sizeof(unsigned long long));    // (value bits, point) seed keys, then the second round's aim seeds
        L.cnt_off = take(imgs * 2 * sizeof(unsigned) + 64 + (size_t)B * 12 * sizeof(unsigned)); // round-1 and round-2 counters, 3 diagnostic u64, bounding-box keys; directly after seed: one memset clears all
        L.thr_off = take(imgs * (size_t)FC_CELLS * (36 + FS_SUB * FS_SUB) * sizeof(float));     // per cell: one 128-byte line of the per-point test, then lpv, U, sub-cell lpv
        L.surv_off = take(imgs * (size_t)L.surv_cap * sizeof(int));
        L.surv2_off = take(imgs * (size_t)L.surv2_cap * sizeof(int));
        L.vox_off = take((size_t)B * ((V + FS_VPC_FILTER - 1) / FS_VPC_FILTER) * 2 * (size_t)L.vox_g * L.vox_g * L.vox_g * sizeof(unsigned long long));
        L.tab_off = take(imgs * (size_t)FC_CELLS * sizeof(float4));              // point-sharded clouds: the cross-rank seed table
    }
    L.total = off;
    return L;
}

int check_common(const void* xyz, const void* offs, const void* views, int B, int V, int w, double rho,
                 int64_t total_pts) {
    if (!xyz || !offs || !views) return fail(SCREAM_E_ARG, "xyz/offs/views must not be NULL");
    if (B <= 0 || V <= 0) return fail(SCREAM_E_ARG, "B and V must be positive (B=%d V=%d)", B, V);
    if (w < 64 || w % 64 != 0) return fail(SCREAM_E_ARG, "w must be a positive multiple of 64 (render.py:33-34), got %d", w);
    if (w > 4096) return fail(SCREAM_E_UNSUPPORTED, "w=%d exceeds 4096", w);
    if (!(rho == rho) || rho == 0.0) return fail(SCREAM_E_ARG, "rho must be a non-zero number");
    if (total_pts <= 0) return fail(SCREAM_E_ARG, "total_pts must be positive");
    if ((long long)B * V * 2 * ((long long)w * w / 256) > 2147483647LL)
        return fail(SCREAM_E_UNSUPPORTED, "B*V*w*w too large for one launch");
    return 0;
}

// CTAs per image of the culled kernel: as many as fill one wave of the 148 SMs (1 CTA / SM) -- every CTA sorts the whole cloud,
// which is free on SMs that would otherwise idle, and searches 1 / parts of the pixels.  At most 64 (two 32-pixel chunks each).
int culled_parts(long long images) {
    if (const char* e = getenv("SCREAM_MAX_PARTS")) { const int m = atoi(e); if (m >= 1) return (int)std::max<long long>(1, std::min<long long>(m, 148 / std::max<long long>(images, 1))); }
    return (int)std::max<long long>(1, std::min<long long>(64, 148 / std::max<long long>(images, 1)));
}

// stage 1: z range of every (item, view) into ws (encoded)
int launch_zrange(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts, const float* pose,
                  const float* views, int B, int V, unsigned* zenc, cudaStream_t st) {
    const int pairs = B * V;
    zrange_init_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(zenc, pairs);
    if (max_item_pts <= 0 || max_item_pts > total_pts) max_item_pts = total_pts;
    if (B > 65535) return fail(SCREAM_E_UNSUPPORTED, "batch too large for the z-range grid");
    if (max_item_pts >= ZW_MIN_PTS && (V + ZW_VPG - 1) / ZW_VPG <= 65535) {
        const int64_t chunks = (max_item_pts + ZW_PTS_PER_CTA - 1) / ZW_PTS_PER_CTA;
        zrange_wide_kernel<<<dim3((unsigned)chunks, (unsigned)((V + ZW_VPG - 1) / ZW_VPG), (unsigned)B), ZW_THREADS, 0, st>>>(
            xyz, offs, pose, views, V, zenc);
    } else {
        const int64_t chunks = (max_item_pts + ZR_PTS_PER_CTA - 1) / ZR_PTS_PER_CTA;
        if (chunks > 2147483647LL) return fail(SCREAM_E_UNSUPPORTED, "cloud too large for the z-range grid");
        zrange_kernel<<<dim3((unsigned)chunks, (unsigned)B), ZR_THREADS, 0, st>>>(xyz, offs, pose, views, V, zenc);
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int scream_abi_version(void) { return SCREAM_ABI_VERSION; }
const char* scream_last_error(void) { return g_err; }

int scream_render_workspace_bytes(int B, int V, int w, int64_t total_pts, size_t* out_bytes) {
    if (!out_bytes) return fail(SCREAM_E_ARG, "out_bytes is NULL");
    if (B <= 0 || V <= 0 || w <= 0 || total_pts < 0) return fail(SCREAM_E_ARG, "bad sizes");
    *out_bytes = ws_layout(B, V, w, total_pts).total;
    return 0;
}

}  // extern "C"

namespace {
// the culled kernels' dynamic shared memory (once per device and thread)
int culled_attr() {
    static thread_local int attr_dev = -1;
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        CUDA_TRY(cudaFuncSetAttribute(forward_culled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FC_SMEM));
        CUDA_TRY(cudaFuncSetAttribute(forward_culled_lattice_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FC_SMEM));
        attr_dev = dev;
    }
    return 0;
}

// Exchange points of the point-sharded protocol (scream_render_forward_sharded): the whole call, or one of its three stages.
struct Shard { int stage = 0, rank = 0, world = 1; int64_t** keys = nullptr; size_t* n_keys = nullptr; float** table = nullptr; size_t* n_table = nullptr; };

int forward_impl(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                 const float* pose, const float* views,
                 int B, int V, int w, double rho, unsigned flags,
                 float* imgs, int32_t* winner, float* zrange,
                 void* ws, size_t ws_bytes, scream_stream_t stream, const Shard& sh)
{
    if (int rc = check_common(xyz, offs, views, B, V, w, rho, total_pts)) return rc;
    if (!imgs && sh.stage != 1 && sh.stage != 2) return fail(SCREAM_E_ARG, "imgs is NULL");
    const WsLayout L = ws_layout(B, V, w, total_pts);
    if (!ws || ws_bytes < L.total) return fail(SCREAM_E_WORKSPACE, "workspace needs %zu bytes, got %zu", L.total, ws_bytes);
    if (flags & SCREAM_F_DUAL) {
        if (!pose) return fail(SCREAM_E_ARG, "SCREAM_F_DUAL needs a pose");
        if (flags & (SCREAM_F_DENSE | SCREAM_F_ZRANGE_IN)) return fail(SCREAM_E_UNSUPPORTED, "SCREAM_F_DUAL with SCREAM_F_DENSE / SCREAM_F_ZRANGE_IN");
        if (L.stream && !(flags & SCREAM_F_NOSTREAM)) return fail(SCREAM_E_UNSUPPORTED, "SCREAM_F_DUAL with giant clouds: render the two poses separately");
        if ((long long)B * V * 3 > 2147483647LL) return fail(SCREAM_E_UNSUPPORTED, "B*V too large for one launch");
    }
    if ((flags & SCREAM_F_DUAL_NOWIN1) && !(flags & SCREAM_F_DUAL)) return fail(SCREAM_E_ARG, "SCREAM_F_DUAL_NOWIN1 without SCREAM_F_DUAL");
    if (flags & SCREAM_F_HALF) {
        if (pose || (flags & SCREAM_F_DUAL)) return fail(SCREAM_E_UNSUPPORTED, "SCREAM_F_HALF takes pre-rotated clouds: no pose, no SCREAM_F_DUAL");
        flags |= SCREAM_F_NOSTREAM;           // the giant-cloud pre-filter bounds the fp32 depth weight
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* zenc = reinterpret_cast<unsigned*>((char*)ws + L.zenc_off);
    if (sh.stage >= 2) {
        // stages 2 and 3 of a sharded call: the encoded range of stage 1 is still in the workspace
    } else if (flags & SCREAM_F_ZRANGE_IN) {
        if (!zrange) return fail(SCREAM_E_ARG, "SCREAM_F_ZRANGE_IN needs a zrange buffer");
        zrange_encode_kernel<<<(B * V * 2 + 255) / 256, 256, 0, st>>>(zrange, zenc, B * V * 2);
    } else if (flags & SCREAM_F_DUAL) {
        if (int rc = launch_zrange(xyz, offs, total_pts, max_item_pts, nullptr, views, B, V, zenc, st)) return rc;
        if (int rc = launch_zrange(xyz, offs, total_pts, max_item_pts, pose, views, B, V, zenc + (size_t)B * V * 2, st)) return rc;
    } else if (int rc = launch_zrange(xyz, offs, total_pts, max_item_pts, pose, views, B, V, zenc, st)) return rc;

    FwdParams P;
    P.xyz = xyz; P.offs = offs; P.pose = pose; P.views = views; P.zenc = zenc;
    P.B = B; P.V = V; P.w = w;
    const float rho2 = (float)(rho * rho);           // tensor * python scalar -> scalar cast to fp32
    P.negk = -rho2 * 0.5f; P.inv_negk = 1.0f / P.negk;
    P.flags = flags; P.imgs = imgs; P.winner = winner; P.zrange = (flags & SCREAM_F_ZRANGE_IN) ? nullptr : zrange;
    P.surv = nullptr; P.surv_cnt = nullptr; P.surv_cap = 0;
    P.surv1 = nullptr; P.surv1_cnt = nullptr; P.surv1_cap = 0;
    P.xyz_t = nullptr; P.pair_n = P.pair_m = 0; P.nparts = 1; P.fuse_zrange = 0; P.tiny = 0;
    // giant clouds: seed z-buffer -> bounds -> filter, then the culled kernel renders the survivors
    const int zero_radius = (int)ceilf(sqrtf(FC_EXP_ZERO / (0.5f * rho2)) * FC_CELLS_PER_NDC) + 1;
    bool lattice_done = false;
    const bool use_stream = L.stream && !(flags & (SCREAM_F_DENSE | SCREAM_F_NOSTREAM)) && zero_radius <= 32;
    if (sh.stage != 0 && !use_stream) return fail(SCREAM_E_UNSUPPORTED, "the sharded stages need the giant-cloud path (w == 64, >= 147456 points per item, no DENSE / NOSTREAM)");
    if (use_stream) {
        StreamParams S;
        S.seedtab = sh.stage != 0 ? reinterpret_cast<float4*>((char*)ws + L.tab_off) : nullptr; S.rank = sh.rank;
        const size_t n_cells = (size_t)B * V * 2 * FC_CELLS;
        S.xyz = xyz; S.offs = offs; S.pose = pose; S.views = views; S.zenc = zenc; S.B = B; S.V = V; S.negk = P.negk;
        S.seed = reinterpret_cast<unsigned long long*>((char*)ws + L.seed_off);
        S.win = reinterpret_cast<float4*>((char*)ws + L.thr_off);
        S.lpv = reinterpret_cast<float*>(S.win + (size_t)B * V * 2 * FC_CELLS * 8);
        S.uimg = S.lpv + (size_t)B * V * 2 * FC_CELLS;
        S.lpv9 = S.uimg + (size_t)B * V * 2 * FC_CELLS;
        S.aim = reinterpret_cast<float2*>(S.lpv9 + (size_t)B * V * 2 * FC_CELLS * FS_SUB * FS_SUB);
        S.seed2 = S.seed + (size_t)B * V * 2 * FC_CELLS;
        S.surv_cnt = reinterpret_cast<unsigned*>((char*)ws + L.cnt_off);
        S.surv = reinterpret_cast<int*>((char*)ws + L.surv_off); S.surv_cap = L.surv_cap;
        S.surv2_cnt = S.surv_cnt + (size_t)B * V * 2;
        S.stats = reinterpret_cast<unsigned long long*>((char*)ws + L.cnt_off + align_up((size_t)B * V * 4 * sizeof(unsigned), 8));
        S.surv2 = reinterpret_cast<int*>((char*)ws + L.surv2_off); S.surv2_cap = L.surv2_cap;
        S.bbox = reinterpret_cast<unsigned*>(S.stats + 8);
        S.vox = reinterpret_cast<unsigned long long*>((char*)ws + L.vox_off); S.vox_g = L.vox_g;
        const int64_t mi = (max_item_pts <= 0 || max_item_pts > total_pts) ? total_pts : max_item_pts;
        // S1 seeds from a subset of the cloud (at most 4 M points per item: ~1000 per pixel): any subset gives valid bounds.
        // Sharded: the ranks' samples are pooled (seed table), so each rank draws its share of the 4 M.
        const int64_t seed_budget = std::max<int64_t>((4 << 20) / std::max(sh.world, 1), 1 << 16);
        S.seed_stride = (int)std::min<int64_t>(std::max<int64_t>((mi + seed_budget - 1) / seed_budget, 1), 64);
        if (flags & SCREAM_F_THINSEED) S.seed_stride = std::max(S.seed_stride, 4);
        const int64_t chunks = ((mi + FS_CHUNK - 1) / FS_CHUNK + S.seed_stride - 1) / S.seed_stride;
        // S3: chunks small enough to fill the GPU (>= ~4 CTAs per SM over all view groups), large enough to amortise staging
        const int vgroups = (V + FS_VPC_FILTER - 1) / FS_VPC_FILTER;
        int64_t fchunk = mi * vgroups * B / (16 * 148);
        fchunk = std::min<int64_t>(std::max<int64_t>(fchunk / FS_FILTER_THREADS * FS_FILTER_THREADS, FS_FILTER_CHUNK_MIN), FS_FILTER_CHUNK_MAX);
        S.filter_chunk = (int)fchunk;
        const int64_t fchunks = (mi + fchunk - 1) / fchunk;
        if (B > 65535 || fchunks * 2 > 65535) return fail(SCREAM_E_UNSUPPORTED, "batch or cloud too large for the stream grid");
        {
            static thread_local int bound_attr_dev = -1;
            int dev = 0;
            CUDA_TRY(cudaGetDevice(&dev));
            if (bound_attr_dev != dev) {
                CUDA_TRY(cudaFuncSetAttribute(stream_bound_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FS_BOUND_SMEM));
                CUDA_TRY(cudaFuncSetAttribute(stream_seed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FS_SEED_SMEM));
                CUDA_TRY(cudaFuncSetAttribute(stream_filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FS_FILTER_SMEM));
                CUDA_TRY(cudaFuncSetAttribute(stream_reseed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FS_RESEED_SMEM));
                bound_attr_dev = dev;
            }
        }
        if (chunks > 2147483647LL || V * 2 > 65535 || (mi + FV_BBOX_PTS - 1) / FV_BBOX_PTS > 2147483647LL) return fail(SCREAM_E_UNSUPPORTED, "cloud too large for the stream grid");
        const bool round2 = S.seed_stride >= 4 && sh.stage == 0;
        // w == 64: one pass in the image's own frame.  w == 128: four passes, one per pixel lattice (StreamParams::offx), each
        // followed by the render of that lattice's pixels from the pass's survivor lists (a pixel belongs to exactly one
        // lattice, so every point that can win it is in that pass's list).
        const int passes = w == 128 ? 4 : 1;
        if (passes > 1 && sh.stage != 0) return fail(SCREAM_E_UNSUPPORTED, "the sharded stages take w == 64 only");
        for (int pass = 0; pass < passes; ++pass) {
            // lattice (a, b) of a 128 x 128 image: pixel 2 q + a has the centre of pixel q of the 64 x 64 image + (a - 1/2) / 64,
            // so the points move by (1/2 - a) / 64 = -+1/128 NDC = -+1/4 cell
            S.offx = S.offy = FC_CELLS_PER_NDC; S.shx = S.shy = 0.0f;
            if (passes > 1) {
                S.shx = (pass & 1) ? -1.0f / 128.0f : 1.0f / 128.0f; S.shy = (pass & 2) ? -1.0f / 128.0f : 1.0f / 128.0f;
                S.offx = FC_CELLS_PER_NDC + S.shx * FC_CELLS_PER_NDC; S.offy = FC_CELLS_PER_NDC + S.shy * FC_CELLS_PER_NDC;
                S.seedtab = nullptr;
            }
            // seed image + survivor counters (+ the clouds' bounding-box keys: computed by the first pass, kept by the others)
            if (sh.stage <= 1) CUDA_TRY(cudaMemsetAsync(S.seed, 0, pass == 0 ? L.thr_off - L.seed_off : (size_t)((char*)S.bbox - (char*)S.seed), st));
            if (sh.stage <= 1)
                stream_seed_kernel<<<dim3((unsigned)((chunks + FS_SEED_CPB - 1) / FS_SEED_CPB), (unsigned)((V + FS_VPC - 1) / FS_VPC), (unsigned)B), FS_THREADS, FS_SEED_SMEM, st>>>(S);
            if (sh.stage == 1 || sh.stage == 2) {
                // stage 1: this rank's seed table + keys (value bits << 32 | ~rank) -> caller: all_reduce(MAX) over the keys
                // stage 2: entries that are not the cell's best are cleared        -> caller: all_reduce(SUM) over the table
                if (sh.stage == 1) stream_seedtab_kernel<<<dim3((unsigned)(V * 2), (unsigned)B), 1024, 0, st>>>(S);
                else stream_seedtab_mask_kernel<<<(unsigned)((n_cells + 255) / 256), 256, 0, st>>>(S, (long long)n_cells);
                CUDA_TRY(cudaGetLastError());
                if (sh.keys) *sh.keys = reinterpret_cast<int64_t*>(S.seed);
                if (sh.n_keys) *sh.n_keys = n_cells;
                if (sh.table) *sh.table = reinterpret_cast<float*>(S.seedtab);
                if (sh.n_table) *sh.n_table = n_cells * 4;
                return 0;
            }
            stream_bound_kernel<<<dim3((unsigned)(V * 2), (unsigned)B), 1024, FS_BOUND_SMEM, st>>>(S);
            S.seedtab = nullptr;                              // a second round seeds from this rank's own survivors again
            if (pass == 0) stream_bbox_kernel<<<dim3((unsigned)((mi + FV_BBOX_PTS - 1) / FV_BBOX_PTS), (unsigned)B), FV_BBOX_THREADS, 0, st>>>(S);
            stream_voxmask_kernel<<<dim3((unsigned)((L.vox_g * L.vox_g * L.vox_g + 255) / 256), (unsigned)(vgroups * 2 * 4), (unsigned)B), 256, 0, st>>>(S);
            stream_filter_kernel<<<dim3((unsigned)vgroups, (unsigned)(fchunks * 2), (unsigned)B), FS_FILTER_THREADS, FS_FILTER_SMEM, st>>>(S);
            // second round: seeds from the survivors (the true per-cell best), bounds again, survivor lists filtered again.
            // Pays once the first round's seeds are a thin sample of the cloud (its survivors grow with the stride).
            // (sharded: the pooled seeds already are the 4 M sample of the whole cloud and a shard's survivor lists are short)
            if (round2) {
                const unsigned r2chunks = (unsigned)((std::min<int64_t>(mi, L.surv_cap) + FS_R2_CHUNK - 1) / FS_R2_CHUNK);
                CUDA_TRY(cudaMemsetAsync(S.seed, 0, L.cnt_off - L.seed_off, st));
                const dim3 g2(r2chunks, (unsigned)(V * 2), (unsigned)B);
                stream_reseed_kernel<<<g2, FS_THREADS, FS_RESEED_SMEM, st>>>(S);
                stream_bound_kernel<<<dim3((unsigned)(V * 2), (unsigned)B), 1024, FS_BOUND_SMEM, st>>>(S);
                stream_refilter_kernel<<<g2, FS_THREADS, 0, st>>>(S);
            }
            if (passes > 1) {
                FwdParams Q = P;
                if (round2) {
                    Q.surv = S.surv2; Q.surv_cnt = S.surv2_cnt; Q.surv_cap = L.surv2_cap;
                    Q.surv1 = S.surv; Q.surv1_cnt = S.surv_cnt; Q.surv1_cap = L.surv_cap;
                } else {
                    Q.surv = S.surv; Q.surv_cnt = S.surv_cnt; Q.surv_cap = L.surv_cap;
                }
                Q.pcol0 = pass & 1; Q.prow0 = (pass >> 1) & 1;
                if (int rc = culled_attr()) return rc;
                forward_culled_lattice_kernel<<<(unsigned)((long long)B * V * 2), FC_THREADS, FC_SMEM, st>>>(Q);
            }
        }
        CUDA_TRY(cudaGetLastError());
        if (passes > 1) {
            lattice_done = true;
        } else if (round2) {
            P.surv = S.surv2; P.surv_cnt = S.surv2_cnt; P.surv_cap = L.surv2_cap;
            P.surv1 = S.surv; P.surv1_cnt = S.surv_cnt; P.surv1_cap = L.surv_cap;
        } else {
            P.surv = S.surv; P.surv_cnt = S.surv_cnt; P.surv_cap = L.surv_cap;
        }
    }
    if (lattice_done) {
        // w == 128 giant cloud: every lattice pass has rendered its pixels
    } else if (flags & SCREAM_F_DENSE) {
        const long long grid = (long long)B * V * 2 * (w * w / FD_TILE);
        forward_dense_kernel<<<(unsigned)grid, FD_THREADS, 0, st>>>(P);
    } else {
        if (int rc = culled_attr()) return rc;
        // small grids: several CTAs per image (forward_culled.cuh: nparts), as many as keep the grid within one wave
        const long long images = (long long)B * V * ((flags & SCREAM_F_DUAL) ? 3 : 2);
        if (!(flags & SCREAM_F_DUAL)) P.nparts = culled_parts(images);
        // grids that fill the GPU: clouds of a few points go to forward_tiny_kernel (one more launch, which returns at once for
        // every other image; small grids keep the single launch: there the launch is what counts)
        P.tiny = !(flags & SCREAM_F_DUAL) && P.nparts == 1 && P.surv == nullptr;
        forward_culled_kernel<<<(unsigned)(images * P.nparts), FC_THREADS, FC_SMEM, st>>>(P);
        if (P.tiny) forward_tiny_kernel<<<(unsigned)(images * ((long long)w * w / FT_SLICE)), FT_THREADS, 0, st>>>(P);
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}
}  // namespace

extern "C" {

int scream_render_forward(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                          const float* pose, const float* views,
                          int B, int V, int w, double rho, unsigned flags,
                          float* imgs, int32_t* winner, float* zrange,
                          void* ws, size_t ws_bytes, scream_stream_t stream)
{
    return forward_impl(xyz, offs, total_pts, max_item_pts, pose, views, B, V, w, rho, flags, imgs, winner, zrange, ws, ws_bytes,
                        stream, Shard());
}

#ifdef FC_TIMELINE
// diagnostic builds only: the phase marks of the last forward_culled_kernel launch (forward_culled.cuh), [1024][24] int64
int scream_debug_fc_marks(long long* host_out) {
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpyFromSymbol(host_out, g_fc_marks, sizeof(long long) * 1024 * 24));
    return 0;
}
#endif

// ---- no-allocation variant (no counterpart in the reference) -------------
// The scratch of scream_render_forward (the encoded depth ranges that the z-range kernels reduce into and the render kernel
// reads) lives in a file-scope device array instead of a caller workspace, for callers that cannot cudaMalloc (CUDA graph
// capture, tight serving loops).  Every GPU that loads the library holds SCREAM_NOALLOC_WS_BYTES until exit.  In host code
// the array's NAME is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static __device__ __align__(256) unsigned char g_noalloc_ws[SCREAM_NOALLOC_WS_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};            // per device, resolved on first use (idempotent: racing threads store the same value)
}

int scream_render_forward_noalloc(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                                  const float* pose, const float* views,
                                  int B, int V, int w, double rho, unsigned flags,
                                  float* imgs, int32_t* winner, float* zrange, scream_stream_t stream)
{
    if (int rc = check_common(xyz, offs, views, B, V, w, rho, total_pts)) return rc;
    const size_t need = ws_layout(B, V, w, total_pts).total;
    if (need > (size_t)SCREAM_NOALLOC_WS_BYTES)
        return fail(SCREAM_E_UNSUPPORTED, "scream_render_forward_noalloc: this call needs %zu bytes of scratch, the static array holds %d "
                    "(B*V <= %d and fewer than %lld points per item on average); use scream_render_forward with a caller workspace",
                    need, SCREAM_NOALLOC_WS_BYTES, SCREAM_NOALLOC_WS_BYTES / 16, (long long)FS_MIN_AVG_PTS);
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(SCREAM_E_UNSUPPORTED, "device ordinal %d beyond %d", dev, NOALLOC_MAX_DEVICES);
    void* ptr = g_noalloc_ptr[dev];
    if (!ptr) {
        CUDA_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_ws));
        if (!ptr) return fail(SCREAM_E_CUDA, "cudaGetSymbolAddress returned NULL");
        g_noalloc_ptr[dev] = ptr;
    }
    // nothing in the array has to start at zero: the z-range stage (zrange_init_kernel / zrange_encode_kernel) rewrites every
    // word the render reads, on the caller's stream, before the render kernel is launched on it
    return forward_impl(xyz, offs, total_pts, max_item_pts, pose, views, B, V, w, rho, flags, imgs, winner, zrange,
                        ptr, (size_t)SCREAM_NOALLOC_WS_BYTES, stream, Shard());
}

int scream_render_forward_sharded(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                                  const float* views, int B, int V, int w, double rho, unsigned flags,
                                  int stage, int rank, int world,
                                  float* imgs, int32_t* winner, float* zrange,
                                  void* ws, size_t ws_bytes,
                                  int64_t** seed_keys, size_t* n_keys, float** seed_table, size_t* n_table_floats,
                                  scream_stream_t stream)
{
    if (stage < 1 || stage > 3) return fail(SCREAM_E_ARG, "stage must be 1, 2 or 3");
    if (rank < 0 || world < 1 || rank >= world) return fail(SCREAM_E_ARG, "bad rank / world (%d / %d)", rank, world);
    if (flags & ~(SCREAM_F_RAW | SCREAM_F_ZRANGE_IN)) return fail(SCREAM_E_UNSUPPORTED, "scream_render_forward_sharded takes SCREAM_F_RAW only (SCREAM_F_ZRANGE_IN is implied)");
    Shard sh; sh.stage = stage; sh.rank = rank; sh.world = world;
    sh.keys = seed_keys; sh.n_keys = n_keys; sh.table = seed_table; sh.n_table = n_table_floats;
    return forward_impl(xyz, offs, total_pts, max_item_pts, nullptr, views, B, V, w, rho, flags | SCREAM_F_ZRANGE_IN, imgs, winner,
                        zrange, ws, ws_bytes, stream, sh);
}

int scream_render_backward(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                           const float* pose, const float* views,
                           int B, int V, int w, double rho, unsigned flags,
                           const float* grad_imgs, const int32_t* winner, const float* zrange,
                           float* grad_xyz, void* /*ws*/, size_t /*ws_bytes*/, scream_stream_t stream)
{
    if (int rc = check_common(xyz, offs, views, B, V, w, rho, total_pts)) return rc;
    if (!grad_imgs || !winner || !zrange || !grad_xyz) return fail(SCREAM_E_ARG, "grad_imgs/winner/zrange/grad_xyz must not be NULL");
    cudaStream_t st = (cudaStream_t)stream;
    BwdParams P;
    P.xyz = xyz; P.offs = offs; P.pose = pose; P.views = views;
    P.B = B; P.V = V; P.w = w;
    P.negk = -(float)(rho * rho) * 0.5f;
    P.flags = flags; P.grad_imgs = grad_imgs; P.winner = winner; P.zrange = zrange; P.grad_xyz = grad_xyz;
    P.xyz_t = nullptr; P.grad_t = nullptr; P.pair_n = P.pair_m = 0;
    // grids that fill the GPU: one CTA per (item, cloud), the atomics in shared memory (backward.cuh)
    const char* no_gather = getenv("SCREAM_BWD_SCATTER");
    if ((long long)B * 2 >= 148 && !(no_gather && no_gather[0] == '1')) {
        const int cap = (int)((max_item_pts > 0 && max_item_pts < BG_CAP) ? max_item_pts : BG_CAP);
        const size_t smem = (size_t)cap * 3 * sizeof(float);
        static thread_local int attr_dev = -1;
        int dev = 0;
        CUDA_TRY(cudaGetDevice(&dev));
        if (attr_dev != dev) {
            CUDA_TRY(cudaFuncSetAttribute(backward_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(BG_CAP * 3 * sizeof(float))));
            attr_dev = dev;
        }
        backward_gather_kernel<<<(unsigned)(B * 2), BG_THREADS, smem, st>>>(P, cap);
        CUDA_TRY(cudaGetLastError());
        return 0;
    }
    CUDA_TRY(cudaMemsetAsync(grad_xyz, 0, (size_t)total_pts * 3 * sizeof(float), st));
    const long long grid = (long long)B * V * 2 * (w * w / BW_THREADS);
    backward_scatter_kernel<<<(unsigned)grid, BW_THREADS, 0, st>>>(P);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---- one pair per call: the reference's own call pattern (pointnet.py:39-40 asserts batch 1; render.py:61 takes the two
// clouds as two tensors).  No torch.cat, no offsets tensor, no z-range launches: ONE kernel launch per forward.
int scream_render_forward_pair(const float* src, int64_t n, const float* tgt, int64_t m,
                               const float* pose, const float* views, int V, int w, double rho, unsigned flags,
                               float* imgs, int32_t* winner, float* zrange, scream_stream_t stream)
{
    if (!src || !tgt) return fail(SCREAM_E_ARG, "src/tgt must not be NULL");
    if (n <= 0 || m <= 0) return fail(SCREAM_E_ARG, "max(): cannot reduce over an empty cloud (render.py:46,49)");
    if (int rc = check_common(src, src, views, 1, V, w, rho, n + m)) return rc;
    if (!imgs) return fail(SCREAM_E_ARG, "imgs is NULL");
    if (flags & ~(SCREAM_F_RAW | SCREAM_F_HALF)) return fail(SCREAM_E_UNSUPPORTED, "scream_render_forward_pair takes SCREAM_F_RAW / SCREAM_F_HALF only");
    if ((flags & SCREAM_F_HALF) && pose) return fail(SCREAM_E_UNSUPPORTED, "SCREAM_F_HALF takes pre-rotated clouds: no pose");
    if (n + m >= FS_MIN_AVG_PTS) return fail(SCREAM_E_UNSUPPORTED, "giant pair: use scream_render_forward (streaming pre-filter)");
    cudaStream_t st = (cudaStream_t)stream;
    FwdParams P;
    P.xyz = src; P.offs = nullptr; P.pose = pose; P.views = views; P.zenc = nullptr;
    P.B = 1; P.V = V; P.w = w;
    const float rho2 = (float)(rho * rho);
    P.negk = -rho2 * 0.5f; P.inv_negk = 1.0f / P.negk;
    P.flags = flags; P.imgs = imgs; P.winner = winner; P.zrange = zrange;
    P.surv = nullptr; P.surv_cnt = nullptr; P.surv_cap = 0;
    P.surv1 = nullptr; P.surv1_cnt = nullptr; P.surv1_cap = 0;
    // virtual base: xyz_t + 3 * (n + i) is target point i (never dereferenced below 3 * n)
    P.xyz_t = reinterpret_cast<const float*>(reinterpret_cast<uintptr_t>(tgt) - (uintptr_t)n * 12u);
    P.pair_n = n; P.pair_m = m; P.fuse_zrange = 1; P.tiny = 0;
    P.nparts = culled_parts((long long)V * 2);
    static thread_local int attr_dev = -1;
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        CUDA_TRY(cudaFuncSetAttribute(forward_culled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FC_SMEM));
        attr_dev = dev;
    }
    forward_culled_kernel<<<(unsigned)((long long)V * 2 * P.nparts), FC_THREADS, FC_SMEM, st>>>(P);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_render_backward_pair(const float* src, int64_t n, const float* tgt, int64_t m,
                                const float* pose, const float* views, int V, int w, double rho, unsigned flags,
                                const float* grad_imgs, const int32_t* winner, const float* zrange,
                                float* grad_src, float* grad_tgt, scream_stream_t stream)
{
    if (!src || !tgt) return fail(SCREAM_E_ARG, "src/tgt must not be NULL");
    if (n <= 0 || m <= 0) return fail(SCREAM_E_ARG, "empty cloud");
    if (int rc = check_common(src, src, views, 1, V, w, rho, n + m)) return rc;
    if (!grad_imgs || !winner || !zrange || !grad_src || !grad_tgt) return fail(SCREAM_E_ARG, "grad_imgs/winner/zrange/grad_src/grad_tgt must not be NULL");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaMemsetAsync(grad_src, 0, (size_t)n * 3 * sizeof(float), st));
    CUDA_TRY(cudaMemsetAsync(grad_tgt, 0, (size_t)m * 3 * sizeof(float), st));
    BwdParams P;
    P.xyz = src; P.offs = nullptr; P.pose = pose; P.views = views;
    P.B = 1; P.V = V; P.w = w;
    P.negk = -(float)(rho * rho) * 0.5f;
    P.flags = flags; P.grad_imgs = grad_imgs; P.winner = winner; P.zrange = zrange; P.grad_xyz = grad_src;
    P.xyz_t = reinterpret_cast<const float*>(reinterpret_cast<uintptr_t>(tgt) - (uintptr_t)n * 12u);
    P.grad_t = reinterpret_cast<float*>(reinterpret_cast<uintptr_t>(grad_tgt) - (uintptr_t)n * 12u);
    P.pair_n = n; P.pair_m = m;
    const long long grid = (long long)V * 2 * (w * w / BW_THREADS);
    backward_scatter_kernel<<<(unsigned)grid, BW_THREADS, 0, st>>>(P);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static int half_rotate_grid(int64_t total_pts, int64_t max_item_pts, int B, dim3* grid) {
    if (B > 65535) return fail(SCREAM_E_UNSUPPORTED, "batch too large for the rotate grid");
    if (max_item_pts <= 0 || max_item_pts > total_pts) max_item_pts = total_pts;
    const int64_t chunks = std::min<int64_t>((max_item_pts + HR_THREADS - 1) / HR_THREADS, 1 << 20);
    *grid = dim3((unsigned)std::max<int64_t>(chunks, 1), (unsigned)B);
    return 0;
}

int scream_half_rotate(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                       const float* views, int B, int V, float* xr, int64_t* offs_out, scream_stream_t stream)
{
    if (!xyz || !offs || !views || !xr || !offs_out) return fail(SCREAM_E_ARG, "xyz/offs/views/xr/offs_out must not be NULL");
    if (B <= 0 || V <= 0 || total_pts <= 0) return fail(SCREAM_E_ARG, "B, V and total_pts must be positive");
    dim3 grid;
    if (int rc = half_rotate_grid(total_pts, max_item_pts, B, &grid)) return rc;
    half_rotate_kernel<<<grid, HR_THREADS, 0, (cudaStream_t)stream>>>(xyz, offs, views, B, V, xr, offs_out);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_half_rotate_backward(const float* grad_xr, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                                const float* views, int B, int V, float* grad_xyz, scream_stream_t stream)
{
    if (!grad_xr || !offs || !views || !grad_xyz) return fail(SCREAM_E_ARG, "grad_xr/offs/views/grad_xyz must not be NULL");
    if (B <= 0 || V <= 0 || total_pts <= 0) return fail(SCREAM_E_ARG, "B, V and total_pts must be positive");
    dim3 grid;
    if (int rc = half_rotate_grid(total_pts, max_item_pts, B, &grid)) return rc;
    half_rotate_backward_kernel<<<grid, HR_THREADS, 0, (cudaStream_t)stream>>>(grad_xr, offs, views, V, grad_xyz);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_nearest_target(const float* src, int64_t n, const float* tgt, int64_t m, float s,
                          float* dist, int64_t* ind, void* ws, size_t ws_bytes, scream_stream_t stream)
{
    if (!src || !tgt || !ws) return fail(SCREAM_E_ARG, "src/tgt/ws must not be NULL");
    if (n <= 0 || m <= 0) return fail(SCREAM_E_ARG, "min(): cannot reduce over an empty cloud (n=%lld m=%lld)", (long long)n, (long long)m);
    if (m > 0xFFFFFFFFLL) return fail(SCREAM_E_UNSUPPORTED, "more than 2^32 targets");
    if (ws_bytes < (size_t)n * 8) return fail(SCREAM_E_WORKSPACE, "workspace needs %zu bytes, got %zu", (size_t)n * 8, ws_bytes);
    const int64_t gx = (n + NN_THREADS - 1) / NN_THREADS, gy = (m + NN_TILE - 1) / NN_TILE;
    const int64_t gs = (n + NN_THREADS * NN_PPT - 1) / (NN_THREADS * NN_PPT);
    if (gx > 2147483647LL || gy > 65535) return fail(SCREAM_E_UNSUPPORTED, "clouds too large for the search grid");
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long* best = reinterpret_cast<unsigned long long*>(ws);
    nn_init_kernel<<<(unsigned)gx, NN_THREADS, 0, st>>>(best, n);
    nn_search_kernel<<<dim3((unsigned)gs, (unsigned)gy), NN_THREADS, 0, st>>>(src, n, tgt, m, s, best);
    nn_finish_kernel<<<(unsigned)gx, NN_THREADS, 0, st>>>(best, n, dist, ind);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_rigid_transform(const float* A, const float* Bp, const int64_t* b_index, const float* weights,
                           const float* valid_dist, float valid_thr, float weight_threshold,
                           int bs, int64_t k, int64_t b_rows, float* T, void* ws, size_t ws_bytes, scream_stream_t stream)
{
    if (!A || !Bp || !T || !ws) return fail(SCREAM_E_ARG, "A/B/T/ws must not be NULL");
    if (bs <= 0 || bs > 65535 || k <= 0 || b_rows <= 0) return fail(SCREAM_E_ARG, "bad sizes (bs=%d k=%lld)", bs, (long long)k);
    if (ws_bytes < (size_t)bs * 16 * sizeof(double)) return fail(SCREAM_E_WORKSPACE, "workspace needs %zu bytes, got %zu", (size_t)bs * 16 * sizeof(double), ws_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    double* acc = reinterpret_cast<double*>(ws);
    CUDA_TRY(cudaMemsetAsync(acc, 0, (size_t)bs * 16 * sizeof(double), st));
    const int64_t gx = std::min<int64_t>((k + KB_THREADS - 1) / KB_THREADS, 592);
    kabsch_accum_kernel<<<dim3((unsigned)gx, (unsigned)bs), KB_THREADS, 0, st>>>(A, Bp, b_index, weights, valid_dist, valid_thr,
                                                                                  weight_threshold, k, b_rows, acc);
    kabsch_solve_kernel<<<(bs + 63) / 64, 64, 0, st>>>(acc, bs, T);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_render_transform(const float* xyz, const int64_t* offs, int64_t total_pts, int64_t max_item_pts,
                            const float* pose, const float* views, int B, int V,
                            float* rotated, float* pv, float* zrange,
                            void* ws, size_t ws_bytes, scream_stream_t stream)
{
    if (int rc = check_common(xyz, offs, views, B, V, 64, 1.0, total_pts)) return rc;
    WsLayout L = ws_layout(B, V, 64, total_pts);
    L.total = align_up((size_t)B * V * 2 * sizeof(unsigned), 256);      // only the encoded z range is used here
    if (!ws || ws_bytes < L.total) return fail(SCREAM_E_WORKSPACE, "workspace needs %zu bytes, got %zu", L.total, ws_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* zenc = reinterpret_cast<unsigned*>((char*)ws + L.zenc_off);
    if (int rc = launch_zrange(xyz, offs, total_pts, max_item_pts, pose, views, B, V, zenc, st)) return rc;
    if (zrange) zrange_decode_kernel<<<(B * V * 2 + 255) / 256, 256, 0, st>>>(zenc, zrange, B * V * 2);
    if (rotated || pv) {
        const int64_t mi = (max_item_pts <= 0 || max_item_pts > total_pts) ? total_pts : max_item_pts;
        const int64_t chunks = (mi + 255) / 256;
        transform_dump_kernel<<<dim3((unsigned)chunks, (unsigned)B), 256, 0, st>>>(
            xyz, offs, total_pts, pose, views, V, zenc, rotated, pv);
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_depth_to_rgb8(const float* imgs, int B, int V, int w, unsigned char* rgb, scream_stream_t stream)
{
    if (!imgs || !rgb) return fail(SCREAM_E_ARG, "imgs/rgb must not be NULL");
    if (B <= 0 || V <= 0 || w <= 0) return fail(SCREAM_E_ARG, "B, V and w must be positive");
    const long long total = (long long)B * V * w * w;
    const long long grid = (total + CZ_THREADS - 1) / CZ_THREADS;
    if (grid > 2147483647LL) return fail(SCREAM_E_UNSUPPORTED, "B*V*w*w too large for one launch");
    depth_to_rgb8_kernel<<<(unsigned)grid, CZ_THREADS, 0, (cudaStream_t)stream>>>(imgs, B, V, w, rgb);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_pack_keys(const float* val, const int32_t* winner, int64_t src_offset, int64_t tgt_offset, int V, int w,
                     int64_t* keys, scream_stream_t stream)
{
    if (!val || !winner || !keys) return fail(SCREAM_E_ARG, "val/winner/keys must not be NULL");
    if (V <= 0 || w <= 0) return fail(SCREAM_E_ARG, "V and w must be positive");
    const long long total = (long long)V * 2 * w * w;
    pack_keys_kernel<<<(unsigned)((total + SK_THREADS - 1) / SK_THREADS), SK_THREADS, 0, (cudaStream_t)stream>>>(
        val, winner, (long long)src_offset, (long long)tgt_offset, w * w, total, reinterpret_cast<long long*>(keys));
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_unpack_keys(const int64_t* keys, int V, int w, float* val, float* imgs, int64_t* winner_global, scream_stream_t stream)
{
    if (!keys || !val || !winner_global) return fail(SCREAM_E_ARG, "keys/val/winner_global must not be NULL");
    if (V <= 0 || w <= 0) return fail(SCREAM_E_ARG, "V and w must be positive");
    const long long total = (long long)V * 2 * w * w;
    unpack_keys_kernel<<<(unsigned)((total + SK_THREADS - 1) / SK_THREADS), SK_THREADS, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const long long*>(keys), total, val, imgs, reinterpret_cast<long long*>(winner_global));
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static int disc_stem_check(int N, int C, int F, int H, int W) {
    if (N <= 0 || C <= 0 || F <= 0 || H <= 0 || W <= 0) return fail(SCREAM_E_ARG, "N, C, F, H, W must be positive");
    if (C > DS_MAX_C || F > DS_MAX_F) return fail(SCREAM_E_UNSUPPORTED, "disc stem: C <= %d and F <= %d (the renderer emits 2 channels, gan.py uses ndf = 64)", DS_MAX_C, DS_MAX_F);
    if ((H | W) & 1) return fail(SCREAM_E_ARG, "H and W must be even (kernel 4, stride 2, padding 1)");
    return 0;
}

int scream_disc_stem_forward(const float* x, const float* weight, const float* bias, int N, int C, int F, int H, int W,
                             float* y, scream_stream_t stream)
{
    if (!x || !weight || !y) return fail(SCREAM_E_ARG, "x/weight/y must not be NULL");
    if (int rc = disc_stem_check(N, C, F, H, W)) return rc;
    DiscStemParams P{};
    P.x = x; P.weight = weight; P.bias = bias; P.y = y; P.N = N; P.C = C; P.F = F; P.H = H; P.W = W;
    const long long total = (long long)N * (H / 2) * (W / 2);
    disc_stem_forward_kernel<<<(unsigned)((total + DS_THREADS - 1) / DS_THREADS), DS_THREADS, 0, (cudaStream_t)stream>>>(P);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_disc_stem_backward(const float* x, const float* weight, const float* y, const float* dy, int N, int C, int F, int H, int W,
                              float* dx, float* dw, float* db, scream_stream_t stream)
{
    if (!x || !weight || !y || !dy) return fail(SCREAM_E_ARG, "x/weight/y/dy must not be NULL");
    if (int rc = disc_stem_check(N, C, F, H, W)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DiscStemParams P{};
    P.x = x; P.weight = weight; P.y = const_cast<float*>(y); P.dy = dy; P.dx = dx; P.dw = dw; P.db = db;
    P.N = N; P.C = C; P.F = F; P.H = H; P.W = W;
    if (dx) {
        if ((H + DS_DG_ROWS - 1) / DS_DG_ROWS > 65535 || N > 65535) return fail(SCREAM_E_UNSUPPORTED, "image count or height beyond the dx grid");
        disc_stem_dgrad_kernel<<<dim3((unsigned)((W + DS_DG_COLS - 1) / DS_DG_COLS), (unsigned)((H + DS_DG_ROWS - 1) / DS_DG_ROWS), (unsigned)N), DS_THREADS, 0, st>>>(P);
    }
    if (dw) {
        CUDA_TRY(cudaMemsetAsync(dw, 0, (size_t)F * C * DS_TAPS * sizeof(float), st));
        if (db) CUDA_TRY(cudaMemsetAsync(db, 0, (size_t)F * sizeof(float), st));
        const long long chunks = ((long long)N * (H / 2) * (W / 2) + DS_WG_POS - 1) / DS_WG_POS;
        const long long ctas = std::min<long long>(chunks, 148 * 8);
        const long long per = (chunks + ctas - 1) / ctas;
        static thread_local int wg_attr_dev = -1;
        int dev = 0;
        CUDA_TRY(cudaGetDevice(&dev));
        if (wg_attr_dev != dev) {
            CUDA_TRY(cudaFuncSetAttribute(disc_stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DS_WG_SMEM));
            wg_attr_dev = dev;
        }
        disc_stem_wgrad_kernel<<<(unsigned)((chunks + per - 1) / per), DS_THREADS, DS_WG_SMEM, st>>>(P, per);
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int scream_render_forward_host(const float* xyz_host, const int64_t* offs_host,
                               const float* pose_host, const float* views_host,
                               int B, int V, int w, double rho, unsigned flags,
                               float* imgs_host, int32_t* winner_host, float* zrange_host, int device)
{
    if (!xyz_host || !offs_host || !views_host || !imgs_host) return fail(SCREAM_E_ARG, "NULL host pointer");
    if (B <= 0 || V <= 0) return fail(SCREAM_E_ARG, "B and V must be positive");
    // the staging slab below is sized for the plain [B,V,2,w,w] layout and zrange is an output only
    if (flags & (SCREAM_F_DUAL | SCREAM_F_ZRANGE_IN | SCREAM_F_HALF))
        return fail(SCREAM_E_UNSUPPORTED, "scream_render_forward_host takes SCREAM_F_RAW / DENSE / NOSTREAM / THINSEED only "
                                           "(DUAL, ZRANGE_IN and HALF need the device entry points)");
    if (device >= 0) CUDA_TRY(cudaSetDevice(device));
    const int64_t total = offs_host[2 * B];
    int64_t max_item = 0;
    for (int b = 0; b < B; ++b) max_item = offs_host[2 * b + 2] - offs_host[2 * b] > max_item ? offs_host[2 * b + 2] - offs_host[2 * b] : max_item;
    if (int rc = check_common(xyz_host, offs_host, views_host, B, V, w, rho, total)) return rc;
    const size_t n_img = (size_t)B * V * 2 * w * w;
    size_t ws_bytes = 0;
    scream_render_workspace_bytes(B, V, w, total, &ws_bytes);
    // one device slab: xyz | offs | pose | views | imgs | winner | zrange | ws
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    const size_t o_xyz = take((size_t)total * 12), o_offs = take((size_t)(2 * B + 1) * 8),
                 o_pose = take((size_t)B * 48), o_views = take((size_t)V * 36), o_imgs = take(n_img * 4),
                 o_win = take(n_img * 4), o_zr = take((size_t)B * V * 8), o_ws = take(ws_bytes);
    char* d = nullptr;
    cudaStream_t st;
    CUDA_TRY(cudaMalloc(&d, off));
    cudaError_t e = cudaStreamCreate(&st);
    if (e != cudaSuccess) { cudaFree(d); return fail(SCREAM_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
    int rc = 0;
    auto run = [&]() -> int {
        CUDA_TRY(cudaMemcpyAsync(d + o_xyz, xyz_host, (size_t)total * 12, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d + o_offs, offs_host, (size_t)(2 * B + 1) * 8, cudaMemcpyHostToDevice, st));
        if (pose_host) CUDA_TRY(cudaMemcpyAsync(d + o_pose, pose_host, (size_t)B * 48, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d + o_views, views_host, (size_t)V * 36, cudaMemcpyHostToDevice, st));
        int r = scream_render_forward((const float*)(d + o_xyz), (const int64_t*)(d + o_offs), total, max_item,
                                      pose_host ? (const float*)(d + o_pose) : nullptr, (const float*)(d + o_views),
                                      B, V, w, rho, flags, (float*)(d + o_imgs),
                                      winner_host ? (int32_t*)(d + o_win) : nullptr, (float*)(d + o_zr),
                                      d + o_ws, ws_bytes, (scream_stream_t)st);
        if (r) return r;
        CUDA_TRY(cudaMemcpyAsync(imgs_host, d + o_imgs, n_img * 4, cudaMemcpyDeviceToHost, st));
        if (winner_host) CUDA_TRY(cudaMemcpyAsync(winner_host, d + o_win, n_img * 4, cudaMemcpyDeviceToHost, st));
        if (zrange_host) CUDA_TRY(cudaMemcpyAsync(zrange_host, d + o_zr, (size_t)B * V * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        return 0;
    };
    rc = run();
    cudaStreamDestroy(st);
    cudaFree(d);
    return rc;
}

}  // extern "C"
