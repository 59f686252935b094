// pvv_internal.cuh -- workspace view and kernel launcher prototypes shared by the .cu files of libpvvote.
#pragma once
#include "pvv_common.cuh"

// Device workspace, carved out of pvv_job.workspace by pvv_capi.cu.  All pointers are device pointers.
struct pvv_ws {
    // --- zeroed at the start of every call (one contiguous memset) ---
    unsigned int *counters;       // [16] see PVV_CTR_*
    long long *img_fg;            // [b]  v3: sum of byte values; eq1: number of ==1 pixels   (K1a, atomics)
    int *counts;                  // [b][vn][hn]  inlier counts (K3, atomics)
    unsigned int *fin_ticket;     // [b][vn]      arrival counter of the K4 pixel-range splits
    // --- written before read ---
    int *img_flags;               // [b]  PVV_IMG_*
    int *tn;                      // [b]  compacted pixels per image (0 when skipped)
    long long *img_off;           // [b]  offset of the image's pixels in the pool
    int *chunk_start;             // [b+1] first voting chunk of each image; [b] = total
    int4 *chunk_tab;              // [cap/256 + b] per voting chunk: {image, POOL index of its first pixel, pixels | first chunk of
                                  //               its image << 15 | image rank << 16, x|y<<16 of the middle pixel}
                                  //               (.x .y .z by K1c, .w by K1d): K3 decodes a work item with one load
    int *img_rank;                // [b]  rank of the image among the images that vote (tn > 0)
    uint32_t *img_org;            // [b]  x|y<<16 of the image's median foreground pixel: origin of its frame in the tensor-pipe K3
    uint4 *tcq;                   // tensor-pipe K3: flagged (hypothesis, 16-pixel group) entries, one slice per CTA of k_count_tc2
    int *tcq_cnt;                 // [tcq_slices] entries pushed per slice (may exceed tcq_slice_cap: the excess was resolved in line)
    int tcq_slices, tcq_slice_cap;
    int want_rank;                // 1: K1c also ranks the images that vote (img_rank, PVV_CTR_NIMG): only k_count_tc2 reads them
    int *tile_cnt;                // [b][ntiles] foreground pixels per compaction tile
    int *tile_off;                // [b][ntiles] exclusive prefix of tile_cnt within the image
    uint32_t *fg_bits;            // [b][ntiles][32] foreground bitmask (bit i of word w = pixel 32w+i of the tile), after
                                  //                 subsampling: the mask itself is read exactly once
    uint32_t *pix_xy;             // [cap]      x | y<<16
    float2 *pix_dir;              // [vn][cap]  raw direction vectors (SoA over keypoints)
    float4 *pix_pq;               // [vn][cap]  frame-independent half-plane constants (tensor-pipe K3 only; written with K2)
    float2 *hyp;                  // [b][vn][hn] hypotheses
    double *fin_part;             // [b][vn][PVV_FIN_SPLIT][5] partial normal equations of the K4 splits
    int32_t *idxs_used;           // optional debug [b][hn][vn][2]
    long long cap;
    int ntiles;
    int hn_pad;                   // hn rounded up to a multiple of 4 (hyp / counts row pitch)
};

enum { PVV_CTR_TICKET = 0, PVV_CTR_WORK = 1, PVV_CTR_STATUS = 2, PVV_CTR_AMB_LO = 4, PVV_CTR_AMB_HI = 5,
       PVV_CTR_NIMG = 6 /* images that vote (tn > 0) */, PVV_CTR_NEED_LO = 8 /* + 9: pool pixels the batch needs (u64) */ };
#define PVV_TCQ_SLICES 148   // = CTAs of the tensor-pipe counting kernel (one per SM of a B200)
enum { PVV_IMG_SKIPPED = 1, PVV_IMG_SUBSAMPLED = 2, PVV_IMG_OVERFLOW = 4 };

struct pvv_dims {
    int b, h, w, vn, hn;
    int fg_rule, mask_dtype;
    int min_num, max_num;
    float thresh;
    unsigned long long seed;
    long long image_id_base;
    long long ms[3];   // mask strides (elements)
    int seg_c;         // PVV_MASK_LOGITS_F32: classes and class stride
    long long seg_cs;
    long long vs[5];   // vertex strides (elements)
};

// Programmatic dependent launch (PDL): the kernels of one call form a chain on the caller's stream; each is launched with
// programmatic stream serialisation, so its launch and CTA scheduling overlap the tail of its predecessor instead of
// following the predecessor's completion by a launch latency.  Contract inside the library: EVERY kernel launched
// through pvv_launch_pdl executes pvv_pdl_enter() as its first statement in every thread -- it waits until the
// predecessor grid has completed and its writes are visible (which, by induction, covers all earlier grids of the
// chain), then lets the next kernel of the chain start launching.
#ifdef __CUDACC__
__device__ __forceinline__ void pvv_pdl_enter()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// Shared-memory carve-out of the library's kernels.  The tensor-pipe counting kernel needs the SM's largest one (224 KB),
// and an SM only changes its carve-out when it is empty: one resident CTA of a kernel that settled for a small carve-out
// (the span-pull kernel of a host-buffer call lives for ~0.2 ms) keeps the counting kernel's CTA for that SM -- and with it
// the whole persistent grid -- waiting (tools/coresidency_probe.cu).  Host-buffer calls, which run sub-batches side by
// side, therefore ask for the largest carve-out in every kernel (-25 us per call, measured); device-buffer calls run one
// chain at a time and keep the default, which is 7 us faster for them (more L1 for the compaction and pre-pass kernels).
// The wish is per calling thread (pvv_host.cu sets it around its sub-batches); the attribute is set when it changes.
inline int &pvv_carveout_wish()
{
    thread_local int wish = cudaSharedmemCarveoutDefault;
    return wish;
}
inline void pvv_hint_carveout(const void *kern)
{
    struct seen_t { const void *fn; int dev, state; };
    thread_local seen_t seen[48];
    thread_local int nseen = 0;
    const int wish = pvv_carveout_wish();
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    seen_t *slot = nullptr;
    for (int i = 0; i < nseen; ++i)
        if (seen[i].fn == kern && seen[i].dev == dev) { slot = &seen[i]; break; }
    if (slot ? slot->state == wish : wish == cudaSharedmemCarveoutDefault && nseen < 48) {
        if (!slot) { seen[nseen].fn = kern; seen[nseen].dev = dev; seen[nseen].state = wish; ++nseen; }
        return;
    }
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, wish);
    cudaGetLastError();
    if (slot) slot->state = wish;
    else if (nseen < 48) { seen[nseen].fn = kern; seen[nseen].dev = dev; seen[nseen].state = wish; ++nseen; }
}

template <typename... KArgs, typename... Args>
inline cudaError_t pvv_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args)
{
    pvv_hint_carveout((const void *)kern);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// K1: compaction (pvv_compact.cu)
cudaError_t pvv_launch_compact(const pvv_dims &d, const void *mask, const float *vertex, const float *selection,
                               const pvv_ws &ws, cudaStream_t st, int *launches);
// K2 + K3: hypotheses and fused inlier counting (pvv_vote.cu)
cudaError_t pvv_launch_hypotheses(const pvv_dims &d, const int32_t *idxs, const pvv_ws &ws, float *hyp_out, int force_exact,
                                  int k3_variant, const float *hyp_inject, cudaStream_t st, int *launches);
cudaError_t pvv_launch_count(const pvv_dims &d, const pvv_ws &ws, int force_exact, int k3_variant, cudaStream_t st,
                             int *launches);
// K4 / K5: finalisation (pvv_finalize.cu)
cudaError_t pvv_launch_v3_finalize(const pvv_dims &d, const pvv_ws &ws, float *out_kp, float *out_conf,
                                   const pvv_debug *dbg, cudaStream_t st, int *launches);
cudaError_t pvv_launch_export_counts(const pvv_dims &d, const pvv_ws &ws, int32_t *counts_out, cudaStream_t st,
                                     int *launches);
cudaError_t pvv_launch_motion_mean(const pvv_dims &d, const pvv_ws &ws, float *out_kp, cudaStream_t st, int *launches);
cudaError_t pvv_launch_cov_finalize(const pvv_dims &d, const pvv_ws &ws, const float *mean, float *out_cov,
                                    cudaStream_t st, int *launches);
cudaError_t pvv_launch_debug_export(const pvv_dims &d, const pvv_ws &ws, const pvv_debug *dbg, cudaStream_t st,
                                    int *launches);
