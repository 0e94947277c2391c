// Internal declarations shared by the kernels and the C-ABI layer of libpillarfront.so (sm_100a only).
#pragma once

#include <stdint.h>

#ifdef PF_HOST_EMUL          // tests/emul: the same kernel bodies on host threads (test infrastructure, never the product)
#include "warp_emul.h"
#else
#include <cuda_runtime.h>
#endif

#include "pillarfront.h"
#include "pf_post_desc.h"

namespace pf {

constexpr int kEmpty = 0x7fffffff;   // grid cell never touched this frame
constexpr int kTileWarps = 8;        // warps per tile CTA in k_tile_rank
constexpr int kTileSteps = 8;        // 32-point steps each warp owns, in point order
constexpr int kTile = kTileWarps * kTileSteps * 32;   // 2048 points per tile
constexpr int kCountCap = 255;       // saturation of the per-(tile,voxel) byte counters (>= max_points)
constexpr int kQPer = 4;             // points per thread in the streaming point kernels (k_quantize, k_reset_cells)

// Geometry and capacities, passed by value to every kernel.
struct Params {
    float off[3];
    float vs[3];
    int nx, ny, nz;
    int cells;        // nx*ny*nz
    int plane;        // nx*ny
    int max_voxels;
    int P;            // max points per voxel
    int F;            // floats per point
    int npts_cap;     // per-frame stride of the per-point scratch arrays (multiple of kTile)
    int tiles_cap;    // npts_cap / kTile
    int vcols;        // row stride of the count matrix (max_voxels rounded up to 16)
    unsigned ny_m, nz_m;   // floor(2^32 / ny), floor(2^32 / nz): magic multipliers of div_small()
    int vbits;        // bits needed for a voxel id: smallest b with 2^b >= max_voxels
};

// Device scratch owned by a handle.  Nothing here is API-visible; none of it counts as algorithmic bytes.
struct Work {
    int32_t *grid;      // [B][cells]   kEmpty | first point index (>=0) | ~voxel_id (<0)
    int32_t *cellid;    // [B][npts_cap] linear cell per point, -1 outside
    int32_t *vr;        // [B][npts_cap] (voxel_id << 8 | in-tile rank), -1 if dropped
    int32_t *tilecnt;   // [B][tiles_cap] look-back status words of k_assign: (state << 30 | creators), zeroed by k_quantize
    int32_t *ticket;    // [B] next tile id to hand out (tiles are numbered in CTA start order), zeroed by k_quantize
    uint8_t *cnt;       // [B][tiles_cap][vcols] points per (tile, voxel), then its exclusive column scan
    int32_t *slotidx;   // [B][max_voxels][P] point index held by each voxel slot
    int32_t *vcell;     // [B][max_voxels] linear cell of each voxel
    int32_t *vcount;    // [B][max_voxels] min(points in voxel, P)
    int32_t *istar;     // [B] cut-off index i*: first point index the sequential loop never reaches
    int32_t *vnum;      // [B] voxels per frame
    int32_t *map;       // [B][plane] cell -> feature row for the stand-alone scatter, -1 = none
    cudaEvent_t *ev;    // host array of kStageEvents events, or nullptr: per-kernel timing (pf_debug_stage_timing)
};

// stage boundaries recorded when Work::ev is set (eager launches only; not capturable into a graph)
enum Stage { kEvStart = 0, kEvQuantize, kEvAssign, kEvTileRank, kEvColumnScan, kEvSlot, kEvGather, kEvFillStart, kEvFill,
             kStageEvents };
#ifndef PF_HOST_EMUL
inline void mark(const Work &w, int k, cudaStream_t s) { if (w.ev) cudaEventRecord(w.ev[k], s); }
#endif

struct VoxelizeArgs {
    const float *points;
    const int32_t *frame_offsets;
    int n_frames;
    int max_frame_points;
    int point_stride;
    int compact;
    int coors_stride;
    float vx, vy, x_off, y_off;
    int keep_cell_map;
    float *voxels;
    int32_t *coors;
    int32_t *num_points;
    int32_t *voxel_num;
    float *features;
    const float *pfn_w, *pfn_scale, *pfn_shift;   // optional fused PFN (pf_voxelize_pfn)
    float *pfn_out;
};

// arguments of the per-voxel gather + decoration kernels (pf_voxelize.cu generic, pf_gather.cu F = 4)
struct GatherArgs {
    const float *points;             // FROM_POINTS
    const int32_t *frame_offsets;    // FROM_POINTS
    int point_stride;
    int compact;
    int coors_stride;
    const float *voxels_in;          // !FROM_POINTS
    const int32_t *num_in;           // !FROM_POINTS
    const int32_t *coors_in;         // !FROM_POINTS
    long long n_voxels;              // !FROM_POINTS
    float vx, vy, x_off, y_off;
    float *voxels;                   // may be null
    int32_t *coors;                  // may be null
    int32_t *num_points;             // may be null
    float *features;                 // may be null
    // fused pillar feature net (eval mode): out[v][c] = max_p relu(scale[c] * <w[c], feature row p> + shift[c]), 64 channels
    const float *pfn_w;              // [64][F+5] = Conv1d weight (64, 9, 1); null = no PFN
    const float *pfn_scale;          // [64] gamma / sqrt(running_var + eps)
    const float *pfn_shift;          // [64] beta - running_mean * scale
    float *pfn_out;                  // [rows][64]
};

// launchers (pf_voxelize.cu / pf_scatter.cu); all asynchronous on `stream`, return the launch error.
cudaError_t launch_fill_i32(int32_t *dst, int32_t value, size_t n, cudaStream_t stream);
// F = 4 fast path of the gather/decorate kernel; returns cudaErrorNotSupported when the shape does not qualify
cudaError_t launch_gather4(const Params &p, const Work &w, const GatherArgs &a, int n_frames, bool from_points,
                           cudaStream_t stream);
cudaError_t launch_voxelize(const Params &p, const Work &w, const VoxelizeArgs &a, cudaStream_t stream);
cudaError_t launch_decorate(const float *voxels, const int32_t *num, const int32_t *coors, int coors_stride,
                            int64_t n_voxels, int P, int F, float vx, float vy, float x_off, float y_off,
                            float *features, cudaStream_t stream, const float *pfn_w = nullptr,
                            const float *pfn_scale = nullptr, const float *pfn_shift = nullptr, float *pfn_out = nullptr);
cudaError_t launch_scatter(const Params &p, const Work &w, const float *feats, int64_t sv, int64_t sc,
                           const int32_t *coors, int coors_stride, int64_t n_voxels, int batch, int n_ch,
                           float *canvas, cudaStream_t stream);
cudaError_t launch_scatter_from_map(const Params &p, const Work &w, const float *feats, int64_t sv, int64_t sc,
                                    int n_frames, int compact, int n_ch, float *canvas, cudaStream_t stream);
// allocation-free scatter: the cell map is a file-scope device array of static_scatter_max_cells() cells in pf_scatter.cu
long long static_scatter_max_cells();
cudaError_t launch_scatter_static(const float *feats, int64_t sv, int64_t sc, const int32_t *coors, int coors_stride,
                                  int64_t n_voxels, int batch, int n_ch, int nx, int ny, float *canvas,
                                  cudaStream_t stream);
cudaError_t launch_scatter_backward(const float *grad_canvas, const int32_t *coors, int coors_stride,
                                    int64_t n_voxels, int batch, int n_ch, int nx, int ny, float *grad_feats,
                                    int64_t sv, int64_t sc, cudaStream_t stream);
cudaError_t launch_anchor_mask(const int32_t *coors, int coors_stride, const int32_t *voxel_num, long long n_rows,
                               int rows_per_frame, int n_frames, int nx, int ny, const int32_t *anchor_coors,
                               long long n_anchors, int32_t *integral, uint8_t *mask, int *table_is_u16,
                               cudaStream_t stream);

size_t gather_smem_needed(int P, int F);   // pf_voxelize.cu
bool pdl_enabled();            // pf_voxelize.cu: launch the chain kernels as programmatic dependents (PF_PDL / pf_debug_set_pdl)
void set_pdl(int on);     // 0 off, 1 on, < 0 auto
int pdl_state();          // 0 off, 1 on, 2 auto
bool pdl_stage(int bit);   // experiments: PF_PDL_MASK selects which launches are dependents

#if defined(__CUDACC__)
// kernel launch with or without the programmatic-dependent-launch attribute
template <class... KArgs, class... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, bool dependent,
                            Args... args)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr, cfg.numAttrs = dependent ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif
int fill_variant();            // pf_scatter.cu: which canvas-fill kernel runs (see pf_debug_set_fill_variant)
void set_fill_variant(int v);

size_t nms_scratch_bytes(int n_seg, int max_seg);
cudaError_t launch_nms(const float *boxes, const float *scores, const int32_t *seg_offsets, int n_seg, long long n_total,
                       int max_seg, float thresh, int post_max, void *scratch, int32_t *keep, int32_t *keep_num,
                       cudaStream_t stream);

void nms_presorted_views(void *scratch, int n_seg, int max_seg, float **sorted_boxes, int32_t **order);
cudaError_t launch_nms_presorted(const int32_t *seg_offsets, int n_seg, long long n_total, int max_seg, float thresh,
                                 int post_max, void *scratch, int32_t *keep, int32_t *keep_num, cudaStream_t stream);
size_t post_scratch_layout(int n_classes, int pre_max, int post_max, int n_frames, long long n_anchors, size_t *offsets);
cudaError_t launch_postprocess(const post::Desc &d, float iou_threshold, const float *cls_preds, const float *box_preds,
                               const float *dir_preds, const uint8_t *anchors_mask, const float *anchors, int n_frames,
                               long long n_anchors, void *scratch, float *det_boxes, float *det_scores, int32_t *det_class,
                               int32_t *det_anchor, int32_t *det_num, cudaStream_t stream);

// ---- small device helpers -------------------------------------------------------------------------------
#if defined(__CUDACC__) || defined(PF_HOST_EMUL)

#ifdef PF_HOST_EMUL
#define PF_DYN_SHARED(T, name) T *name = reinterpret_cast<T *>(pf_emul_dyn_smem)
#else
#define PF_DYN_SHARED(T, name) extern __shared__ __align__(16) T name[]
#endif

// Linear cell of a point, or -1 if outside.  Bit-exact restatement of framework/voxel_generator.py:90-91:
// fp32 subtract, IEEE fp32 divide, floor, and a range test that NaN / +-Inf / out-of-int32 values fail
// (SURVEY.md F4, F7).  The intrinsics pin round-to-nearest and forbid FMA contraction / fast division.
__device__ __forceinline__ int cell_of_point(float x, float y, float z, const Params &p)
{
    const float qx = floorf(__fdiv_rn(__fsub_rn(x, p.off[0]), p.vs[0]));
    const float qy = floorf(__fdiv_rn(__fsub_rn(y, p.off[1]), p.vs[1]));
    const float qz = floorf(__fdiv_rn(__fsub_rn(z, p.off[2]), p.vs[2]));
    const bool in = (qx >= 0.f) && (qx < (float)p.nx) && (qy >= 0.f) && (qy < (float)p.ny) && (qz >= 0.f) &&
                    (qz < (float)p.nz);
    if (!in) return -1;
    return ((int)qx * p.ny + (int)qy) * p.nz + (int)qz;
}

// n / d for 0 <= n < 2^31, d >= 1, m = floor(2^32 / d): the estimate is at most one short, fixed by one compare.
__device__ __forceinline__ int div_small(int n, int d, unsigned m)
{
    if (d == 1) return n;
    int q = (int)__umulhi((unsigned)n, m);
    if (n - q * d >= d) ++q;
    return q;
}

#ifndef PF_HOST_EMUL
__device__ __forceinline__ unsigned lanemask_lt()
{
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// streaming (evict-first) 16-byte store: canvas lines are written once and never re-read by this library
__device__ __forceinline__ void st_cs_f4(float *ptr, float a, float b, float c, float d)
{
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}
#endif

// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-stream-serialization attribute may be
// scheduled as soon as every CTA of the kernel in front of it on the stream has executed pdl_trigger() (or exited); it
// runs its prologue (index arithmetic, shared-memory clears, loads of CALLER inputs) under the predecessor's tail and
// stops at pdl_wait() until the predecessor grid has completed and its writes are visible.  Rules used in this library:
// every chain kernel triggers first thing, and touches nothing an earlier chain kernel writes before its pdl_wait().
// Without the launch attribute both instructions are no-ops.
#ifdef PF_HOST_EMUL
inline void pdl_wait() {}
inline void pdl_trigger() {}
#else
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// flag / status words other CTAs poll or publish, and single words several threads may write or read concurrently BY
// DESIGN (always through L2, never a cached copy; a 32-bit access is indivisible on the device).  The host emulation
// states the same thing to ThreadSanitizer with relaxed atomics.
#ifdef PF_HOST_EMUL
__device__ __forceinline__ int ld_flag(const int32_t *p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
__device__ __forceinline__ void st_flag(int32_t *p, int v) { __atomic_store_n(p, v, __ATOMIC_RELAXED); }
__device__ __forceinline__ void st_same(int32_t *p, int v) { __atomic_store_n(p, v, __ATOMIC_RELAXED); }
#else
// a word several threads store the SAME value to (plain store on the device)
__device__ __forceinline__ void st_same(int32_t *p, int v) { *p = v; }
__device__ __forceinline__ int ld_flag(const int32_t *p) { return *reinterpret_cast<const volatile int32_t *>(p); }
__device__ __forceinline__ void st_flag(int32_t *p, int v) { *reinterpret_cast<volatile int32_t *>(p) = v; }
#endif

#endif  // __CUDACC__ || PF_HOST_EMUL

}  // namespace pf
