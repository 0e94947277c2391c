/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
 *
 * Plain-C, CPU restatement of the reference's point-cloud front end.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library,
 * and only as the checker or the timed CPU baseline -- never as a fallback for the CUDA path.
 *
 * Parity status: PINNED.  tests/golden/ holds outputs of the reference itself
 * (/root/reference, imported in the build container by oracle/make_golden.py); tests/test_oracle.py checks
 * every function below against them bit-for-bit (voxelisation, scatter) or to the stated fp32 tolerance
 * (decoration).
 *
 * Reference lines restated (paths relative to /root/reference):
 *   oracle_points_to_voxels  <- framework/voxel_generator.py:82-106  (numba points_to_voxels)
 *   oracle_generate          <- framework/voxel_generator.py:28-40   (VoxelGenerator.generate)
 *   oracle_decorate          <- networks/pointpillars8_shared.py:32-54 (PointNet.forward prefix)
 *   oracle_scatter           <- networks/pointpillars8_shared.py:76-111 (PointPillarsScatter.forward, B = 1
 *                               index rule x*ny + y applied per batch item; see DESIGN.md on F10)
 *   oracle_scatter_cmajor    <- networks/pointpillars8_export.py:44-51 (Scatter.run, channel-major features)
 *   oracle_nms_iou           <- framework/nms.py:102-113 (iou_device) in the arithmetic numba's CUDA target compiles
 *                               it to (tests/golden/nms_iou_ptx_ops.json, oracle/make_golden_nms_f64.py)
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/build.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* float -> int32 the way the numba loop does it: `fptosi float to i32`, lowered to cvttss2si on x86,
 * which yields INT32_MIN for NaN and for anything outside the int32 range.  Written out
 * explicitly because the plain C cast is undefined for those inputs. */
static inline int32_t f2i_x86(float f)
{
    if (!(f >= -2147483648.0f && f < 2147483648.0f)) return INT32_MIN;
    return (int32_t)f;
}

/* framework/voxel_generator.py:82-106.  `grid` is coor_to_voxelidx, C-order [nx][ny][nz], pre-filled with -1;
 * voxels/num/coors are pre-zeroed by the caller exactly as generate() does (:29-32). */
int32_t oracle_points_to_voxels(const float *points, int64_t n_points, int32_t n_feat,
                                float *voxels, int32_t *num_per_voxel, int32_t *coors, int32_t *grid,
                                const float *voxel_size, const float *offset, const int32_t *grid_size,
                                int32_t max_voxels, int32_t max_points)
{
    int32_t voxel_num = 0;
    const int64_t ny = grid_size[1], nz = grid_size[2];
    for (int64_t i = 0; i < n_points; ++i) {
        const float *p = points + i * n_feat;
        int32_t c[3];
        int inside = 1;
        for (int d = 0; d < 3; ++d) {
            /* :90  fp32 subtract, fp32 divide, floor, cast (F4) */
            float diff = p[d] - offset[d];
            float q = diff / voxel_size[d];
            c[d] = f2i_x86(floorf(q));
            if (c[d] < 0 || c[d] >= grid_size[d]) inside = 0;   /* :91 */
        }
        if (!inside) continue;                                   /* :92-93 */
        int32_t *slot = grid + ((int64_t)c[0] * ny + c[1]) * nz + c[2];
        int32_t vid = *slot;                                     /* :94 */
        if (vid == -1) {
            if (voxel_num >= max_voxels) break;                  /* :96-97  break, not continue (F5) */
            vid = voxel_num;
            *slot = vid;
            coors[3 * vid + 0] = c[0];
            coors[3 * vid + 1] = c[1];
            coors[3 * vid + 2] = c[2];
            voxel_num += 1;
        }
        int32_t k = num_per_voxel[vid];                          /* :102 */
        if (k < max_points) {                                    /* :103  first-come cap (F6) */
            memcpy(voxels + ((int64_t)vid * max_points + k) * n_feat, p, sizeof(float) * (size_t)n_feat);
            num_per_voxel[vid] = k + 1;
        }
    }
    return voxel_num;
}

/* framework/voxel_generator.py:28-40 including its per-call allocations (np.zeros x3, -np.ones grid), so a
 * timing of this function is comparable to timing VoxelGenerator.generate.  Outputs are caller-owned
 * buffers of the max sizes; the scratch grid is allocated and freed here like the reference's temporary. */
int32_t oracle_generate(const float *points, int64_t n_points, int32_t n_feat,
                        float *voxels, int32_t *num_per_voxel, int32_t *coors,
                        const float *voxel_size, const float *offset, const int32_t *grid_size,
                        int32_t max_voxels, int32_t max_points)
{
    const size_t cells = (size_t)grid_size[0] * (size_t)grid_size[1] * (size_t)grid_size[2];
    memset(voxels, 0, sizeof(float) * (size_t)max_voxels * (size_t)max_points * (size_t)n_feat);
    memset(num_per_voxel, 0, sizeof(int32_t) * (size_t)max_voxels);
    memset(coors, 0, sizeof(int32_t) * 3 * (size_t)max_voxels);
    int32_t *grid = (int32_t *)malloc(sizeof(int32_t) * cells);
    if (!grid) return -1;
    memset(grid, 0xff, sizeof(int32_t) * cells);                 /* -1 everywhere */
    int32_t v = oracle_points_to_voxels(points, n_points, n_feat, voxels, num_per_voxel, coors, grid,
                                        voxel_size, offset, grid_size, max_voxels, max_points);
    free(grid);
    return v;
}

/* networks/pointpillars8_shared.py:32-54.  features (V,P,F+5) = [point, xyz - mean, xy - pillar centre],
 * rows p >= num[v] zeroed.  The mean divides the sum over all P slots (padding rows are zeros) by the
 * saturated count (:32).  Sum order here is slot order, fp32; torch's own order is unspecified (F14), so
 * parity with the reference is a tolerance (tests/test_oracle.py).  centre = float(c)*v + off is two
 * separately rounded fp32 ops (F13). */
void oracle_decorate(const float *voxels, const int32_t *num_per_voxel, const int32_t *coors,
                     int32_t coors_stride, int64_t n_voxels, int32_t max_points, int32_t n_feat,
                     float vx, float vy, float x_off, float y_off, float *features)
{
    const int32_t nf_out = n_feat + 5;
    for (int64_t v = 0; v < n_voxels; ++v) {
        const float *vox = voxels + v * max_points * n_feat;
        float *out = features + v * max_points * nf_out;
        float sx = 0.f, sy = 0.f, sz = 0.f;
        for (int32_t p = 0; p < max_points; ++p) {
            sx = sx + vox[p * n_feat + 0];
            sy = sy + vox[p * n_feat + 1];
            sz = sz + vox[p * n_feat + 2];
        }
        const float cnt = (float)num_per_voxel[v];
        float mx = sx / cnt, my = sy / cnt, mz = sz / cnt;
        float px = (float)coors[v * coors_stride + 0] * vx;
        float py = (float)coors[v * coors_stride + 1] * vy;
        float cxm = px + x_off, cym = py + y_off;
        for (int32_t p = 0; p < max_points; ++p) {
            const float *pt = vox + p * n_feat;
            float *o = out + p * nf_out;
            if (p < num_per_voxel[v]) {
                for (int32_t k = 0; k < n_feat; ++k) o[k] = pt[k];
                o[n_feat + 0] = pt[0] - mx;
                o[n_feat + 1] = pt[1] - my;
                o[n_feat + 2] = pt[2] - mz;
                o[n_feat + 3] = pt[0] - cxm;
                o[n_feat + 4] = pt[1] - cym;
            } else {
                for (int32_t k = 0; k < nf_out; ++k) o[k] = 0.f;   /* features *= mask (:47-54) */
            }
        }
    }
}

/* networks/pointpillars8_shared.py:76-111 with row-major (V,C) features.  canvas is (B,C,nx,ny), zeroed
 * here (:82); pillar v of batch item b lands at [b, :, x*ny + y] (:93,:100).  coors_stride 3 => single
 * frame (b = 0); coors_stride 4 => batch id in the LAST column (framework/utils.py:38-41).  Later rows win
 * on duplicate cells, like the sequential torch-CPU index_put. */
void oracle_scatter(const float *feats, const int32_t *coors, int32_t coors_stride, int64_t n_voxels,
                    int32_t n_ch, int32_t nx, int32_t ny, int32_t batch, float *canvas)
{
    const int64_t plane = (int64_t)nx * ny;
    memset(canvas, 0, sizeof(float) * (size_t)batch * (size_t)n_ch * (size_t)plane);
    for (int64_t v = 0; v < n_voxels; ++v) {
        const int32_t x = coors[v * coors_stride + 0], y = coors[v * coors_stride + 1];
        const int32_t b = coors_stride >= 4 ? coors[v * coors_stride + 3] : 0;
        if (x < 0 || x >= nx || y < 0 || y >= ny || b < 0 || b >= batch) continue;
        float *dst = canvas + (int64_t)b * n_ch * plane + (int64_t)x * ny + y;
        for (int32_t c = 0; c < n_ch; ++c) dst[(int64_t)c * plane] = feats[v * n_ch + c];
    }
}

/* networks/pointpillars8_export.py:44-51: channel-major (C,V) features, single frame. */
void oracle_scatter_cmajor(const float *feats_cv, const int32_t *coors, int32_t coors_stride,
                           int64_t n_voxels, int32_t n_ch, int32_t nx, int32_t ny, float *canvas)
{
    const int64_t plane = (int64_t)nx * ny;
    memset(canvas, 0, sizeof(float) * (size_t)n_ch * (size_t)plane);
    for (int64_t v = 0; v < n_voxels; ++v) {
        const int32_t x = coors[v * coors_stride + 0], y = coors[v * coors_stride + 1];
        if (x < 0 || x >= nx || y < 0 || y >= ny) continue;
        for (int32_t c = 0; c < n_ch; ++c)
            canvas[(int64_t)c * plane + (int64_t)x * ny + y] = feats_cv[(int64_t)c * n_voxels + v];
    }
}

/* The whole path for a batch of independent frames, one frame per OpenMP thread: generate -> decorate ->
 * scatter (pillar features for the scatter are supplied, as the PFN between the two is not on the path).
 * This is bench.py's CPU baseline body.  Per-thread scratch is allocated inside, like the reference's
 * per-call temporaries.  Returns the total number of voxels, or -1 on allocation failure. */
int64_t oracle_frontend_batch(const float *points, const int64_t *frame_offsets, int32_t n_frames,
                              int32_t n_feat, const float *voxel_size, const float *offset,
                              const int32_t *grid_size, int32_t max_voxels, int32_t max_points,
                              float vx, float vy, float x_off, float y_off,
                              const float *pillar_feats, int32_t n_ch,
                              float *voxels, int32_t *num_per_voxel, int32_t *coors, int32_t *voxel_num,
                              float *features, float *canvas, int32_t n_threads)
{
    int64_t total = 0;
    int failed = 0;
    const int64_t vox_slab = (int64_t)max_voxels * max_points * n_feat;
    const int64_t feat_slab = (int64_t)max_voxels * max_points * (n_feat + 5);
    const int64_t plane = (int64_t)grid_size[0] * grid_size[1];
#pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads) reduction(+ : total)
    for (int32_t f = 0; f < n_frames; ++f) {
        const float *pts = points + frame_offsets[f] * n_feat;
        const int64_t n = frame_offsets[f + 1] - frame_offsets[f];
        float *vox = voxels + f * vox_slab;
        int32_t *num = num_per_voxel + (int64_t)f * max_voxels;
        int32_t *co = coors + (int64_t)f * max_voxels * 3;
        int32_t v = oracle_generate(pts, n, n_feat, vox, num, co, voxel_size, offset, grid_size,
                                    max_voxels, max_points);
        if (v < 0) { failed = 1; continue; }
        voxel_num[f] = v;
        oracle_decorate(vox, num, co, 3, v, max_points, n_feat, vx, vy, x_off, y_off,
                        features + f * feat_slab);
        oracle_scatter(pillar_feats + (int64_t)f * max_voxels * n_ch, co, 3, v, n_ch,
                       grid_size[0], grid_size[1], 1, canvas + (int64_t)f * n_ch * plane);
        total += v;
    }
    return failed ? -1 : total;
}

/* iou_device(a = box i, b = box j) for all pairs, framework/nms.py:102-113, instruction for instruction as numba 0.65 /
 * NVVM compile the reference's source for the GPU (PTX kept by oracle/make_golden_nms_f64.py): min / max and the edge
 * differences in float32, cvt to float64, `+ 1`, max(., 0), products and the quotient in float64, and ONE contraction --
 * Sa + Sb is fma(aw, ah, Sb) with Sb = bw * bh rounded, so the result is not symmetric in (a, b) at the last bit when a
 * product is inexact.  The kernel calls it with a = the earlier (higher-scored) box (nms.py:139-140). */
void oracle_nms_iou(const float *boxes, int32_t n, double *iou)
{
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < n; ++i) {
        const float *a = boxes + 4 * (int64_t)i;
        const double aw = (double)(a[2] - a[0]) + 1.0, ah = (double)(a[3] - a[1]) + 1.0;
        for (int32_t j = 0; j < n; ++j) {
            const float *b = boxes + 4 * (int64_t)j;
            const float left = fmaxf(a[0], b[0]), right = fminf(a[2], b[2]);
            const float top = fmaxf(a[1], b[1]), bottom = fminf(a[3], b[3]);
            const double width = fmax((double)(right - left) + 1.0, 0.0);
            const double height = fmax((double)(bottom - top) + 1.0, 0.0);
            const double inter = width * height;
            const double sb = ((double)(b[2] - b[0]) + 1.0) * ((double)(b[3] - b[1]) + 1.0);
            const double sab = fma(aw, ah, sb);
            iou[(int64_t)i * n + j] = inter / (sab - inter);
        }
    }
}
