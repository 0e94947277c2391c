/*
 * pillarfront.h -- C ABI of libpillarfront.so, the B200 (sm_100a) point-cloud front end.
 *
 * This is the drop-in boundary for the reference's voxelise -> pillar decoration -> BEV scatter path
 * (SURVEY.md §8b).  The reference has no FFI of its own (it is Python + numba); the entry points below are
 * what a ctypes binding of that path needs, one per reference function:
 *
 *   pf_voxelize          <-> framework/voxel_generator.py:28-40, 82-106  (VoxelGenerator.generate,
 *                            numba points_to_voxels), batched over independent frames
 *   pf_decorate          <-> networks/pointpillars8_shared.py:32-54      (PointNet.forward, decoration prefix;
 *                            same code inlined at pointpillars8_export.py:186-199, pointpillars_trt.py:103-116)
 *   pf_voxelize_decorate <-> the two above fused (the pillar never leaves the SM between them)
 *   pf_scatter           <-> networks/pointpillars8_shared.py:76-111     (PointPillarsScatter.forward),
 *                            networks/pointpillars8_export.py:44-51      (Scatter.run, channel-major features),
 *                            networks/pointpillars8_trt.py:139-205       (Scatter_cuda + its CUDA-C kernel)
 *   pf_scatter_backward  <-> autograd of the index_put in pointpillars8_shared.py:100 (a gather)
 *   pf_pfn / pf_voxelize_pfn <-> networks/pointpillars8_shared.py:24-28, 57-60 (pfn_layers + max, eval mode), fused
 *                            behind the decoration ("next" row N1)
 *   pf_anchor_mask       <-> framework/box_np_ops.py:236-257 (fused_get_anchors_mask_gpu = init_map_gpu :168,
 *                            cumx_gpu :175, cumy_gpu :183, get_anchors_mask_gpu :214), called from
 *                            AnchorAssigner.create_mask, framework/anchor_assigner.py:322-335 ("next" row N2)
 *   pf_nms               <-> framework/nms.py:6-44 (nms_gpu: host argsort, nms_kernel :116-147 with iou_device
 *                            :102-113, host nms_postprocess :85-99), called through nms(), framework/inference.py:689-703
 *                            ("next" row N4)
 *   pf_postprocess       <-> framework/inference.py:26-138 (Inference.infer_gpu: anchor mask, sigmoid, score mask,
 *                            top-k, box_np_ops.box_decode :406-424, standup boxes, nms, direction flip, range mask,
 *                            limit_period), batched over frames and classes (row N4, the part around pf_nms)
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer named *device* is CUDA device memory on the handle's
 *     device; `stream` is a cudaStream_t passed as void*; all work is enqueued asynchronously on it and no
 *     call synchronises, allocates or frees (so every call can be captured into a CUDA graph).  Exceptions, all host-side
 *     helpers of the single-frame / multi-GPU drop-in paths and documented where they are declared: pf_create / pf_destroy
 *     (allocate), pf_frame_run and pf_frame_copy_out(synchronize = 1) and pf_runner_wait (wait for the device).
 *   - every function returns 0 on success, a negative pf_status on bad arguments, or a positive
 *     cudaError_t; it never throws.  pf_last_error_string() describes the last failure of the calling thread.
 *   - a handle is bound to one device and is not re-entrant: use one handle per (GPU, stream).
 *   - all outputs are written exactly once per call; nothing needs pre-zeroing by the caller.
 */
#ifndef PILLARFRONT_H_
#define PILLARFRONT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PF_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define PF_API __attribute__((visibility("default")))
#else
#define PF_API
#endif

typedef enum pf_status {
    PF_OK = 0,
    PF_ERR_INVALID_ARGUMENT = -1,
    PF_ERR_CAPACITY = -2,      /* n_frames / points per frame / voxels exceed what the handle was created for */
    PF_ERR_UNSUPPORTED = -3,   /* e.g. max_points > 255, num_features outside 3..16 */
    PF_ERR_OUT_OF_MEMORY = -4,
    PF_ERR_STATE = -5          /* a kept cell map was not consumed (or none is pending): see pf_scatter_from_map */
} pf_status;

/* Grid geometry + caps: the attributes of the reference's VoxelGenerator (framework/voxel_generator.py:17-22). */
typedef struct pf_grid_desc {
    float offset[3];      /* lower corner of cell (0,0,0): VoxelGenerator.offset */
    float voxel_size[3];  /* VoxelGenerator.voxel_size */
    int32_t grid[3];      /* (nx, ny, nz): VoxelGenerator.grid_size; cells are C-ordered [x][y][z] */
    int32_t max_voxels;   /* config['max_voxels'] */
    int32_t max_points;   /* config['max_num_points'] (P), <= 255 */
    int32_t num_features; /* floats per point (F), 3..16; config['num_point_features'] */
} pf_grid_desc;

typedef struct pf_handle pf_handle;

PF_API int pf_version(void);
PF_API const char *pf_last_error_string(void);

/* Create the per-(GPU,stream) workspace: cell grid, per-point scratch, tile counters, slot lists, cell->row map.
 * `device` < 0 means the current device. */
PF_API int pf_create(const pf_grid_desc *desc, int max_frames, int max_points_per_frame, int device,
              pf_handle **out);
PF_API int pf_destroy(pf_handle *h);
PF_API size_t pf_workspace_bytes(const pf_handle *h);

/* Host helper of the single-frame drop-in path (VoxelGenerator.generate, framework/voxel_generator.py:28-40, is handed a
 * pageable numpy array per frame): copies `bytes` from `src` to `dst` -- the caller's PINNED staging buffer -- with up to
 * `threads` host threads (1..16), so that the host -> device copy that follows is one asynchronous DMA instead of the
 * driver's synchronous staged copy.  Pure host code; touches no CUDA state. */
PF_API int pf_stage_copy(void *dst, const void *src, size_t bytes, int threads);

/* Up to three asynchronous copies (direction inferred from the pointers: device -> device, or device -> PINNED host) on
 * `stream` in one call: how VoxelGenerator.generate hands out exactly V rows of voxels / coors / num_points after it has
 * read the voxel count (one library call instead of three framework copies).  A zero size skips that copy. */
PF_API int pf_copy3(void *dst0, const void *src0, size_t bytes0, void *dst1, const void *src1, size_t bytes1,
                    void *dst2, const void *src2, size_t bytes2, void *stream);

/* Host side of ONE single-frame call (VoxelGenerator.generate, framework/voxel_generator.py:28-40; the reference's operating
 * point is one frame per call) in one library call, so that the host language pays one foreign call instead of six:
 * copy `bytes` of `points` (pageable) into `pinned_points` with up to `threads` host threads (streaming stores), write the
 * frame offsets {0, n_points} into `pinned_offsets`, make `stream` wait for `pending_event` (may be NULL: a hand-out copy of
 * the previous call on another stream), launch `graph_exec` -- a cudaGraphExec_t the caller captured once: H2D of the
 * offsets, pf_voxelize*, D2H of the voxel count into `pinned_voxel_num` -- on `stream`, record `done_event` (a cudaEvent_t
 * of the caller), wait for it and return the voxel count.
 *   device_points != NULL: the upload of the points is done HERE, in front of the graph: the frame is staged in up to four
 *                          pieces and each piece is handed to the copy engine as soon as it is staged (the graph then holds
 *                          no copy of the points);
 *   device_points == NULL: the whole frame is staged first and the graph is expected to contain the H2D of the staging buffer.
 * `device` is made current for the call.  This is the only call of the library that synchronises, by design: the row count of
 * the outputs is data dependent. */
PF_API int pf_frame_run(int device, void *graph_exec, void *stream, void *done_event, void *pending_event,
                        void *pinned_points, const void *points, size_t bytes, int threads, void *device_points,
                        int32_t *pinned_offsets, int n_points, const int32_t *pinned_voxel_num, int32_t *voxel_num);
/* The hand-out that follows: pf_copy3, then `record_event` (may be NULL) recorded on `stream`, then, if `synchronize`, a wait
 * for the stream (pinned host destinations that are returned to the caller as numpy arrays). */
PF_API int pf_frame_copy_out(int device, void *dst0, const void *src0, size_t bytes0, void *dst1, const void *src1,
                             size_t bytes1, void *dst2, const void *src2, size_t bytes2, void *stream, void *record_event,
                             int synchronize);

/* Batched multi-frame upload (the collate + H2D of row N3, framework/utils.py:23-48, for a list of per-frame arrays): copies
 * `n_frames` frames back to back into `device_points` on `stream` and writes the frame offsets (n_slots + 1 int32, frames
 * beyond n_frames empty) into `pinned_offsets`, host memory the caller then copies / captures to the device.
 *   kinds[b] = 0: frames[b] is pageable host memory -- copied into `pinned_stage` (same layout as device_points) by up to
 *                 `threads` host threads, then one asynchronous H2D;
 *   kinds[b] = 1: pinned host memory -- one asynchronous H2D straight from it;   kinds[b] = 2: device memory (D2D);
 *   kinds[b] < 0: unknown -- the library asks the driver (cudaPointerGetAttributes).
 * rows[b] = points in frame b; row_bytes = bytes per point row.  One call per batch, so a Python worker thread holds the
 * interpreter lock for microseconds, not for the copies. */
PF_API int pf_upload_frames(void *device_points, void *pinned_stage, const void *const *frames, const int64_t *rows,
                            const int32_t *kinds, int n_frames, int n_slots, int row_bytes, int threads,
                            int32_t *pinned_offsets, void *stream);

/* Native per-GPU runner of the frame-sharded front end (the reference pins everything to cuda:0,
 * train.py:27,190,276).  One host thread per runner; a ring of `n_slots` slots, each prepared once by the host language:
 * device buffer for a batch of frames (frames_per_batch x capacity rows of row_bytes), pinned staging buffer of the same
 * size (may be NULL if only pinned / device frames are submitted), pinned frame offsets (frames_per_batch + 1 int32), a
 * stream, and an instantiated CUDA graph (cudaGraphExec_t) of the step that starts by copying those offsets to the device.
 * pf_runner_submit enqueues one batch (pointers only; the frame memory must stay valid until pf_runner_wait returned) and
 * returns a ticket >= 0 (or a negative pf_status); batches run in ticket order, ticket t in slot t % n_slots, i.e. the
 * outputs of ticket t are overwritten when ticket t + n_slots runs.  pf_runner_wait blocks until the batch has left the GPU
 * and reports its device time (upload + graph, ms; -1 if the slot was already reused).  kinds as in pf_upload_frames. */
typedef struct pf_runner pf_runner;
PF_API int pf_runner_create(int device, int n_slots, int frames_per_batch, int row_bytes, int stage_threads, pf_runner **out);
PF_API int pf_runner_set_slot(pf_runner *r, int slot, void *graph_exec, void *stream, void *device_points,
                              void *pinned_stage, int32_t *pinned_offsets);
PF_API long long pf_runner_submit(pf_runner *r, const void *const *frames, const int64_t *rows, const int32_t *kinds,
                                  int n_frames);
PF_API int pf_runner_wait(pf_runner *r, long long ticket, float *device_ms);
PF_API int pf_runner_destroy(pf_runner *r);
PF_API const char *pf_runner_last_error(void);

/* Diagnostics: with timing enabled every voxelise / fill call of this handle records a CUDA event after each of
 * its kernels (eager launches only -- do not capture such a call into a graph).  pf_debug_stage_times waits for
 * the last call and writes the 8 stage durations in milliseconds (quantize, assign, tile_rank, column_scan, slot,
 * gather, gap between gather and fill, canvas fill; -1 for a stage that did not run); returns the count. */
PF_API int pf_debug_stage_timing(pf_handle *h, int enable);
/* Diagnostics / A-B measurements: which kernel writes the canvas (process-wide; also the environment variable
 * PF_FILL_VARIANT read at first use).  0 = one warp per 128 cells, 16-byte streaming stores; 1 = one warp per 256
 * cells, 32-byte streaming stores (STG.256, the default); 2 = TMA bulk stores of shared-memory tiles (cp.async.bulk,
 * UBLKCP), 512 bytes per copy; 3 = as 2 with an L2 evict-first hint; 4 / 5 = as 2 / 3 with 2 KB per copy; 6 = one warp
 * per 512 cells, two 32-byte stores per lane and channel; -1 = the library default: 6 for canvases of at least 2^25 cells
 * (batch * nx * ny), 0 for canvases of at most 1.25 * 2^20 cells (one or two frames), else 1.  Every variant writes the same bytes.  Returns the
 * variant now in force, or a negative pf_status. */
PF_API int pf_debug_set_fill_variant(int variant);
/* Diagnostics / A-B measurements: programmatic dependent launch of the chain (process-wide; also the environment variables
 * PF_PDL=0|1|2 and PF_PDL_MASK read at first use).  enable = 1: the kernels of pf_voxelize* after the first and the canvas fill
 * of pf_scatter_from_map are launched with the programmatic-stream-serialization attribute: they are scheduled while the
 * kernel in front of them on the stream drains, do their index arithmetic and shared-memory clears, and wait
 * (griddepcontrol.wait) before they read anything that kernel wrote.  enable = 0: plain launches, the default.  enable < 0
 * ("auto", PF_PDL=2): only the second quantize range of a single-frame call is a dependent.  Off by default because it measured
 * slower whenever more than one front end is in flight and within the noise for a lone frame
 * (profiles/r2_chain_experiments.md E5).  Results are identical in every mode; graphs captured while a mode is in force keep
 * its edges.  Returns the state now in force (0 off, 1 on, 2 auto). */
PF_API int pf_debug_set_pdl(int enable);
PF_API int pf_debug_stage_times(pf_handle *h, float *ms, int capacity);

/* Output layouts (both written densely, padding rows zeroed):
 *   slab    (compact == 0): frame f owns rows [f*max_voxels, f*max_voxels + voxel_num[f]) of every output;
 *   compact (compact != 0): frames are concatenated, frame f starts at row sum(voxel_num[0..f)), exactly what
 *                           framework/utils.py:23-48 (merge_second_batch) builds on the host.
 *   voxels     float [rows][P][F]
 *   coors      int32 [rows][coors_stride]  (x, y, z[, frame id]) -- coors_stride is 3 or 4; the frame id goes
 *              in the LAST column like merge_second_batch's pad (framework/utils.py:38-41)
 *   num_points int32 [rows]
 *   voxel_num  int32 [n_frames]   voxels produced per frame (device memory)
 *   features   float [rows][P][F+5] (pf_voxelize_decorate only)
 *
 * points        device float [frame_offsets[n_frames]][point_stride] (first F floats of each row are used)
 * frame_offsets device int32 [n_frames+1], frame f is rows [frame_offsets[f], frame_offsets[f+1])
 * max_frame_points: host-side upper bound of any frame's size in this call (<= the handle's), sizes the grids.  A
 *               frame whose device-side length exceeds it is processed up to this bound (its first
 *               max_frame_points points), deterministically; nothing is read or written beyond it.
 */
PF_API int pf_voxelize(pf_handle *h, const float *points, const int32_t *frame_offsets, int n_frames,
                int max_frame_points, int point_stride, int compact, int coors_stride,
                float *voxels, int32_t *coors, int32_t *num_points, int32_t *voxel_num, void *stream);

/* As pf_voxelize, and additionally writes the decorated pillar features; `voxels` may be NULL when only the
 * features are wanted.  vx, vy, x_off, y_off as PointNet.__init__ forms them (pointpillars8_shared.py:16-19).
 * keep_cell_map != 0 leaves the handle's cell->voxel map populated so that a following
 * pf_scatter_from_map can fill the canvas without re-indexing the coordinates (nz must be 1).  The map must be
 * consumed by exactly one pf_scatter_from_map over the same n_frames (or dropped with pf_discard_cell_map) before
 * the handle voxelises again; otherwise the next voxelise call returns PF_ERR_STATE.  The bookkeeping is host-side
 * and follows call order, so capture such a pair into a CUDA graph together, never one half of it. */
PF_API int pf_voxelize_decorate(pf_handle *h, const float *points, const int32_t *frame_offsets, int n_frames,
                         int max_frame_points, int point_stride, int compact, int coors_stride,
                         float vx, float vy, float x_off, float y_off, int keep_cell_map,
                         float *voxels, int32_t *coors, int32_t *num_points, int32_t *voxel_num,
                         float *features, void *stream);

/* "Next" row N1: voxelise + decorate + the pillar feature net in one pass, eval mode --
 *   pillar_feats[row][c] = max_p relu(scale[c] * <weight[c][0..9), feature row p> + shift[c]),  c < 64
 * i.e. Conv1d(9 -> 64, k = 1, bias = False) + BatchNorm1d (running statistics folded into scale / shift by the
 * caller: scale = gamma / sqrt(var + eps), shift = beta - mean * scale) + ReLU + max over the P slots, exactly the
 * tail of PointNet.forward (networks/pointpillars8_shared.py:24-28, 57-60).  The (rows, P, 9) feature tensor never
 * touches HBM unless `features` is non-NULL.  Needs F == 4, point_stride == 4, n_ch == 64.
 *   pfn_weight  device float [64][9]   (Conv1d weight (64, 9, 1), contiguous)
 *   pillar_feats device float [rows][64]   row layout as the other outputs (slab / compact) */
PF_API int pf_voxelize_pfn(pf_handle *h, const float *points, const int32_t *frame_offsets, int n_frames,
                    int max_frame_points, int point_stride, int compact, int coors_stride,
                    float vx, float vy, float x_off, float y_off, int keep_cell_map,
                    const float *pfn_weight, const float *pfn_scale, const float *pfn_shift, int n_ch,
                    float *voxels, int32_t *coors, int32_t *num_points, int32_t *voxel_num,
                    float *features, float *pillar_feats, void *stream);

/* Stand-alone decorate + PFN of already voxelised pillars: PointNet.forward in eval mode
 * (networks/pointpillars8_shared.py:30-60) -> pillar_feats float [n_voxels][64]. */
PF_API int pf_pfn(const float *voxels, const int32_t *num_points, const int32_t *coors, int coors_stride,
           int64_t n_voxels, int max_points, int num_features, float vx, float vy, float x_off, float y_off,
           const float *pfn_weight, const float *pfn_scale, const float *pfn_shift, int n_ch,
           float *pillar_feats, void *stream);

/* Stand-alone decoration of already voxelised pillars (host-known row count).  max_points 1..4096 and num_features
 * 3..16 as long as one CTA's pillars fit the device's opt-in shared memory (about 8 * max_points * (2 * num_features + 5)
 * * 4 bytes outside the F = 4 fast path; e.g. F = 16 allows max_points <= ~190 on B200); PF_ERR_UNSUPPORTED otherwise.
 * pf_create applies the same rule to (desc.max_points, desc.num_features). */
PF_API int pf_decorate(const float *voxels, const int32_t *num_points, const int32_t *coors, int coors_stride,
                int64_t n_voxels, int max_points, int num_features,
                float vx, float vy, float x_off, float y_off, float *features, void *stream);

/* BEV scatter.  canvas float [batch][n_ch][nx][ny] is written completely (zeros where no pillar).
 * feats: element (v, c) at feats[v*feat_stride_v + c*feat_stride_c]  ((V,C) row-major: C,1; (C,V): 1,V).
 * coors rows (x, y, z[, batch id]); rows with batch id / x / y out of range are ignored; on duplicate cells the
 * highest row wins (the sequential index_put behaviour). */
PF_API int pf_scatter(pf_handle *h, const float *feats, int64_t feat_stride_v, int64_t feat_stride_c,
               const int32_t *coors, int coors_stride, int64_t n_voxels, int batch, int n_ch,
               float *canvas, void *stream);

/* Allocation-free BEV scatter.  This entry point has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  Same arguments, canvas contents and duplicate rule as pf_scatter, but no
 * handle: the cell -> row map pf_scatter keeps in the handle's workspace lives in a file-scope device array of the
 * library instead, PF_SCATTER_STATIC_MAX_CELLS = 2^22 int32 cells (16 MB; every GPU that loads the library holds it until
 * exit).
 *   limit    batch * nx * ny <= PF_SCATTER_STATIC_MAX_CELLS: one frame of every shipped grid (10 cm: 1600 x 1600), six
 *            frames of the 20 cm grid (800 x 800), four of the 16 cm grid, seventeen of nuScenes (512 x 480);
 *            beyond that PF_ERR_CAPACITY with a message naming the limit, nothing is launched
 *   device   runs on the CURRENT device (make the canvas's device current; `stream` must belong to it)
 *   calls    no cudaMalloc / cudaFree / allocating library routine; three kernel launches on `stream` (map reset, map
 *            index, canvas fill).  The array's device address is resolved once per device on the first call
 *            (cudaGetSymbolAddress), so make the first call on a device outside graph capture.
 *   sharing  every call on a device shares the one array, which keeps the last call's contents (each call resets the
 *            cells it is about to use first): calls that overlap on the GPU -- two streams, two host threads -- RACE.
 *            Serialise them on one stream or with events. */
#define PF_SCATTER_STATIC_MAX_CELLS (1 << 22)
PF_API int pf_scatter_static(const float *feats, int64_t feat_stride_v, int64_t feat_stride_c, const int32_t *coors,
                      int coors_stride, int64_t n_voxels, int batch, int n_ch, int nx, int ny, float *canvas,
                      void *stream);

/* Fill the canvas from the cell map left by pf_voxelize_decorate(keep_cell_map=1): feature row of voxel v of
 * frame f is row_base(f) + v with row_base = f*max_voxels (slab) or the running sum of voxel_num (compact). */
PF_API int pf_scatter_from_map(pf_handle *h, const float *feats, int64_t feat_stride_v, int64_t feat_stride_c,
                        int n_frames, int compact, int n_ch, float *canvas, void *stream);

/* Drop a kept cell map without filling a canvas (resets the grid cells of the pending frames); no-op when none is
 * pending. */
PF_API int pf_discard_cell_map(pf_handle *h, void *stream);

/* grad_feats (v, c) = grad_canvas[b][c][x][y]; same stride convention as pf_scatter. */
PF_API int pf_scatter_backward(const float *grad_canvas, const int32_t *coors, int coors_stride, int64_t n_voxels,
                        int batch, int n_ch, int nx, int ny, float *grad_feats, int64_t feat_stride_v,
                        int64_t feat_stride_c, void *stream);

/* Anchor occupancy mask of a batch of frames: mask[f][a] = 1 iff the reference's four-corner sum of the 2-D
 * running-sum pillar-count table of frame f over anchor rectangle a is > 0.
 *   coors         device int32 rows (x, y, ...), `coors_stride` ints per row (>= 2)
 *   voxel_num     device int32 [n_frames] or NULL.
 *                   NULL                   : rows [0, n_rows) are all valid; their frame is column 3 when
 *                                            n_frames > 1 (merge_second_batch layout, needs coors_stride >= 4), else 0
 *                   rows_per_frame  > 0    : slab layout of pf_voxelize (frame f = rows [f*rows_per_frame, +voxel_num[f]))
 *                   rows_per_frame == 0    : compact layout of pf_voxelize (first sum(voxel_num) rows, frame in column 3)
 *   n_rows        rows to scan (an upper bound when voxel_num is given)
 *   anchor_coors  device int32 [n_anchors][4] = (minx, miny, maxx, maxy) from get_anchor_coor
 *                 (framework/box_np_ops.py:288-304); 16-byte aligned
 *   integral      device scratch of n_frames * nx * ny * 4 bytes.  Afterwards it holds the running-sum table
 *                 [n_frames][nx][ny], as uint16 when a frame cannot hold more than 65535 pillars
 *                 (rows_per_frame, or n_rows when rows_per_frame == 0, is <= 65535), else as int32
 *   mask          device uint8 [n_frames][n_anchors], 0 / 1 (numpy bool layout)
 * Pillars with x, y or frame out of range are skipped; a negative corner index counts from the end of its axis
 * (numpy / numba indexing), one past the end is clamped. */
PF_API int pf_anchor_mask(const int32_t *coors, int coors_stride, const int32_t *voxel_num, int64_t n_rows,
                   int rows_per_frame, int n_frames, int nx, int ny, const int32_t *anchor_coors,
                   int64_t n_anchors, int32_t *integral, uint8_t *mask, void *stream);

/* Standup-box NMS of a batch of independent segments (one per class and frame), entirely on the device.
 *   boxes         device float [n_total][4] = (x0, y0, x1, y1), 16-byte aligned
 *   scores        device float [n_total]
 *   seg_offsets   device int32 [n_segments + 1], or NULL for one segment of n_total boxes
 *   max_segment_boxes  upper bound on a segment's size, <= 4096 (the reference's pre_max_size is 1000)
 *   iou_threshold strict `>` test on the reference's IoU (box sides + 1; float64 quotient of float32 differences)
 *   scratch       device, pf_nms_scratch_bytes(n_segments, max_segment_boxes) bytes, 256-byte aligned
 *   keep          device int32 [n_segments][post_max]: indices (within the segment) of the kept boxes in
 *                 descending score order, -1 padded  (= nms(...)[:post_max_size] of the reference)
 *   keep_num      device int32 [n_segments]: number kept (<= post_max); -1 if the segment exceeded the bound
 * Equal scores are ordered higher index first (a stable ascending argsort read backwards). */
PF_API size_t pf_nms_scratch_bytes(int n_segments, int max_segment_boxes);
PF_API int pf_nms(const float *boxes, const float *scores, const int32_t *seg_offsets, int n_segments, int64_t n_total,
           int max_segment_boxes, float iou_threshold, int post_max, void *scratch, int32_t *keep,
           int32_t *keep_num, void *stream);

/* Detection post-processing of a batch of frames, entirely on the device: the per-class loop of
 * Inference.infer_gpu (framework/inference.py:26-138) around pf_nms --
 *   anchor mask (:44-47), sigmoid + max over the class channels (:50-54), score mask (:58-64), top-k (:67-69),
 *   box_decode (framework/box_np_ops.py:406-424), centre boxes -> corners -> standup boxes (box_np_ops.py:81-99,
 *   122-153, 64-78, 899-904), nms (:85-91), direction flip (:97-104), range mask (:106-109), limit_period (:113).
 * The reference does this per frame and class with ~10 device synchronisations, a device->host copy of the top-k
 * rows, numpy decoding, and the numba NMS; here one call handles n_frames * n_classes segments in six launches.
 *
 *   cls_preds    device float [n_frames][n_anchors][n_cls_channels]   (logits)
 *   box_preds    device float [n_frames][n_anchors][7]                (x, y, z, l, w, h, r encodings)
 *   dir_preds    device float [n_frames][n_anchors][2]
 *   anchors_mask device uint8 [n_frames][n_anchors] (numpy bool layout, e.g. the output of pf_anchor_mask), or NULL
 *   anchors      device float [n_anchors][7]
 *   scratch      device, pf_post_scratch_bytes(...) bytes, 256-byte aligned
 *   det_boxes    device float [n_frames][n_classes * post_max][7]   location (3), dimensions (3), rotation_y
 *   det_scores   device float [n_frames][n_classes * post_max]
 *   det_class    device int32 [n_frames][n_classes * post_max]      index into class_range (detect_class order)
 *   det_anchor   device int32 [n_frames][n_classes * post_max] or NULL: the anchor each detection came from
 *   det_num      device int32 [n_frames]: valid rows of each frame; they are ordered like the reference's
 *                concatenated lists (classes in class_range order, descending score inside a class)
 *
 * Order: the reference sorts by the sigmoid score and leaves ties undefined (torch.topk, numpy argsort); this
 * library orders by (logit descending, anchor index ascending), which refines that order and does not depend on an
 * exp() implementation.  x, y, rotation_y are bit-exact restatements of the numpy arithmetic; z and the dimensions
 * go through expf and are a few ulp from numpy's.  NaN logits never pass the score mask (as in the reference). */
typedef struct pf_post_desc {
    int32_t n_classes;          /* 1..8 */
    int32_t class_range[8][2];  /* [start, end) anchor indices of each class: AnchorAssigner.class_masks */
    int32_t n_cls_channels;     /* last dimension of cls_preds (1 in the reference's head) */
    int32_t pre_max;            /* Inference._nms_pre_max_size (1000), 1..4096 */
    int32_t post_max;           /* Inference._nms_post_max_size (300), 1..1024 */
    float score_threshold;      /* Inference._nms_score_threshold (0.05) */
    float iou_threshold;        /* Inference._nms_iou_threshold (0.1) */
    double center_limit[6];     /* config['center_limit'] */
} pf_post_desc;

/* Regions of the scratch buffer (byte offsets from pf_post_scratch_layout): intermediate results a caller or a
 * test may read after the call.  "rows" are indexed seg_offsets[segment] + rank, segment = frame * n_classes + class. */
typedef enum pf_post_region {
    PF_POST_CAND = 0,        /* uint64 candidate keys                                  */
    PF_POST_CAND_COUNT,      /* int32 [segments] anchors that passed mask and threshold */
    PF_POST_SEL_KEY,         /* uint64 [segments][pre_max] selected keys, best first     */
    PF_POST_SEL_NUM,         /* int32 [segments] min(candidates, pre_max)                */
    PF_POST_SEG_OFFSETS,     /* int32 [segments + 1] running sum of SEL_NUM              */
    PF_POST_SEL_ANCHOR,      /* int32 rows: anchor index                                 */
    PF_POST_SEL_SCORE,       /* float rows: sigmoid score                                */
    PF_POST_SEL_BOX,         /* float rows x 7: decoded boxes                            */
    PF_POST_SEL_STANDUP,     /* float rows x 4: standup boxes handed to the NMS          */
    PF_POST_SEL_DIR,         /* int32 rows: direction label                              */
    PF_POST_SEL_RANK,        /* float rows: pre_max - rank (decreasing along the order) */
    PF_POST_KEEP,            /* int32 [segments][post_max] kept ranks, -1 padded         */
    PF_POST_KEEP_NUM,        /* int32 [segments]                                         */
    PF_POST_NMS_SCRATCH,
    PF_POST_REGIONS
} pf_post_region;

PF_API size_t pf_post_scratch_bytes(const pf_post_desc *desc, int n_frames, int64_t n_anchors);
/* offsets: size_t [PF_POST_REGIONS]; returns the total (0 on invalid arguments). */
PF_API size_t pf_post_scratch_layout(const pf_post_desc *desc, int n_frames, int64_t n_anchors, size_t *offsets);
PF_API int pf_postprocess(const pf_post_desc *desc, const float *cls_preds, const float *box_preds, const float *dir_preds,
                   const uint8_t *anchors_mask, const float *anchors, int n_frames, int64_t n_anchors, void *scratch,
                   float *det_boxes, float *det_scores, int32_t *det_class, int32_t *det_anchor, int32_t *det_num,
                   void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PILLARFRONT_H_ */
