// Launch side of the detection post-processing (pf_postprocess): score -> select -> decode -> NMS (overlap mask + greedy
// walk of pf_nms.cu; its sort is skipped, the decode kernel emits the boxes in NMS order) -> finish.
// Kernel bodies and the reference lines they follow: pf_post_kernels.cuh.  Six launches and one 4-byte-per-segment
// memset per batch, nothing synchronises or allocates, so the whole call can be captured into a CUDA graph.
#include "pf_internal.cuh"
#include "pf_post_kernels.cuh"

namespace pf {

namespace {

inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

// Byte offsets of the scratch regions, in the order of pf_post_region (include/pillarfront.h); returns the total.
size_t post_scratch_layout(int n_classes, int pre_max, int post_max, int n_frames, long long n_anchors, size_t *offsets)
{
    const size_t S = (size_t)n_frames * n_classes, rows = S * (size_t)pre_max;
    const size_t sizes[PF_POST_REGIONS] = {
        (size_t)n_frames * (size_t)n_anchors * 8,   // PF_POST_CAND
        S * 4,                                      // PF_POST_CAND_COUNT
        rows * 8,                                   // PF_POST_SEL_KEY
        S * 4,                                      // PF_POST_SEL_NUM
        (S + 1) * 4,                                // PF_POST_SEG_OFFSETS
        rows * 4,                                   // PF_POST_SEL_ANCHOR
        rows * 4,                                   // PF_POST_SEL_SCORE
        rows * 28,                                  // PF_POST_SEL_BOX
        rows * 16,                                  // PF_POST_SEL_STANDUP
        rows * 4,                                   // PF_POST_SEL_DIR
        rows * 4,                                   // PF_POST_SEL_RANK
        S * (size_t)post_max * 4,                   // PF_POST_KEEP
        S * 4,                                      // PF_POST_KEEP_NUM
        nms_scratch_bytes((int)S, pre_max),         // PF_POST_NMS_SCRATCH
    };
    size_t off = 0;
    for (int r = 0; r < PF_POST_REGIONS; ++r) {
        if (offsets) offsets[r] = off;
        off += up256(sizes[r]);
    }
    return off;
}

cudaError_t launch_postprocess(const post::Desc &d, float iou_threshold, const float *cls_preds, const float *box_preds,
                               const float *dir_preds, const uint8_t *anchors_mask, const float *anchors, int n_frames,
                               long long n_anchors, void *scratch, float *det_boxes, float *det_scores, int32_t *det_class,
                               int32_t *det_anchor, int32_t *det_num, cudaStream_t stream)
{
    size_t o[PF_POST_REGIONS];
    post_scratch_layout(d.n_classes, d.pre_max, d.post_max, n_frames, n_anchors, o);
    char *base = static_cast<char *>(scratch);
    auto at = [&](int r) { return base + o[r]; };
    unsigned long long *cand = reinterpret_cast<unsigned long long *>(at(PF_POST_CAND));
    int *cand_count = reinterpret_cast<int *>(at(PF_POST_CAND_COUNT));
    unsigned long long *sel_key = reinterpret_cast<unsigned long long *>(at(PF_POST_SEL_KEY));
    int *sel_num = reinterpret_cast<int *>(at(PF_POST_SEL_NUM));
    int *seg_offsets = reinterpret_cast<int *>(at(PF_POST_SEG_OFFSETS));
    int *sel_anchor = reinterpret_cast<int *>(at(PF_POST_SEL_ANCHOR));
    float *sel_score = reinterpret_cast<float *>(at(PF_POST_SEL_SCORE));
    float *sel_box = reinterpret_cast<float *>(at(PF_POST_SEL_BOX));
    float *sel_standup = reinterpret_cast<float *>(at(PF_POST_SEL_STANDUP));
    int *sel_dir = reinterpret_cast<int *>(at(PF_POST_SEL_DIR));
    float *sel_rank = reinterpret_cast<float *>(at(PF_POST_SEL_RANK));
    int *keep = reinterpret_cast<int *>(at(PF_POST_KEEP));
    int *keep_num = reinterpret_cast<int *>(at(PF_POST_KEEP_NUM));
    const int S = n_frames * d.n_classes;
    float *nms_boxes;
    int32_t *nms_order;
    nms_presorted_views(at(PF_POST_NMS_SCRATCH), S, d.pre_max, &nms_boxes, &nms_order);

    cudaError_t e = cudaMemsetAsync(cand_count, 0, (size_t)S * 4, stream);
    if (e != cudaSuccess) return e;
    const int per_block = post::kScoreThreads * post::kScorePerThread;
    if (n_anchors > 0) {
        const dim3 grid((unsigned)((n_anchors + per_block - 1) / per_block), (unsigned)n_frames);
        const bool vec = d.n_ch == 1 && n_anchors % 4 == 0 && (reinterpret_cast<uintptr_t>(cls_preds) & 15) == 0 &&
                         (reinterpret_cast<uintptr_t>(anchors_mask) & 3) == 0;
        if (vec)
            post::k_post_score<true><<<grid, post::kScoreThreads, 0, stream>>>(d, cls_preds, anchors_mask, n_anchors, cand, cand_count);
        else
            post::k_post_score<false><<<grid, post::kScoreThreads, 0, stream>>>(d, cls_preds, anchors_mask, n_anchors, cand, cand_count);
    }
    int npad = 1;
    while (npad < d.pre_max) npad <<= 1;
    post::k_post_select<<<S, post::kSelectThreads, (size_t)npad * 8, stream>>>(d, n_anchors, cand, cand_count, sel_key, sel_num);
    const dim3 dgrid((unsigned)((d.pre_max + post::kDecodeThreads - 1) / post::kDecodeThreads), (unsigned)S);
    post::k_post_decode<<<dgrid, post::kDecodeThreads, 0, stream>>>(d, n_anchors, S, sel_key, sel_num, box_preds, dir_preds,
                                                                    anchors, seg_offsets, sel_anchor, sel_score, sel_box,
                                                                    sel_standup, sel_dir, sel_rank, nms_boxes, nms_order);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // the boxes are already in their final order (k_post_decode wrote them where pf_nms's sort would have): mask + walk only
    e = launch_nms_presorted(seg_offsets, S, (long long)S * d.pre_max, d.pre_max, iou_threshold, d.post_max,
                             at(PF_POST_NMS_SCRATCH), keep, keep_num, stream);
    if (e != cudaSuccess) return e;
    post::k_post_finish<<<n_frames, post::kFinishThreads, 0, stream>>>(d, seg_offsets, keep, keep_num, sel_box, sel_score,
                                                                       sel_dir, sel_anchor, det_boxes, det_scores, det_class,
                                                                       det_anchor, det_num);
    return cudaGetLastError();
}

}  // namespace pf
