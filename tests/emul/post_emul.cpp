// TEST INFRASTRUCTURE ONLY.  Runs the post-processing kernel bodies (csrc/pf_post_kernels.cuh) on host threads through
// cuda_emul.h and exposes the four stages to ctypes, so that tests/test_post_emul.py can hold the kernel LOGIC (candidate
// lists, radix select, bitonic sort, gather + decode, ordered compaction) to the oracle without a GPU.  expf / sinf /
// cosf are glibc's here and CUDA's on the device; the GPU tests own the numerics.
#define PF_HOST_EMUL 1
#include "pf_post_kernels.cuh"

using namespace pf::post;

namespace {

Desc make_desc(int n_classes, const int *ranges, int n_ch, int pre_max, int post_max, float thr, const double *limit)
{
    Desc d{};
    d.n_classes = n_classes;
    for (int c = 0; c < n_classes; ++c) d.lo[c] = ranges[2 * c], d.hi[c] = ranges[2 * c + 1];
    d.n_ch = n_ch, d.pre_max = pre_max, d.post_max = post_max, d.score_thr = thr;
    for (int i = 0; i < 6; ++i) d.limit[i] = limit ? limit[i] : 0.0;
    set_prefilter(d);
    return d;
}

}  // namespace

extern "C" {

// score + select + decode; every output array is caller-allocated with the sizes of the device scratch regions
int emul_post_front(int n_classes, const int *ranges, int n_ch, int pre_max, float thr, const float *cls, const float *box,
                    const float *dir, const uint8_t *amask, const float *anchors, int n_frames, long long A,
                    unsigned long long *cand, int *cand_count, unsigned long long *sel_key, int *sel_num, int *seg_offsets,
                    int *sel_anchor, float *sel_score, float *sel_box, float *sel_standup, int *sel_dir, float *sel_rank,
                    float *nms_boxes, int *nms_order, int force_scalar)
{
    const Desc d = make_desc(n_classes, ranges, n_ch, pre_max, 1, thr, nullptr);
    const int S = n_frames * n_classes;
    for (int s = 0; s < S; ++s) cand_count[s] = 0;
    const int per_block = kScoreThreads * kScorePerThread;
    if (A > 0) {
        const dim3 grid((unsigned)((A + per_block - 1) / per_block), (unsigned)n_frames);
        if (n_ch == 1 && A % 4 == 0 && !force_scalar)        // the same choice pf_post.cu makes (host arrays are aligned)
            pf_emul_launch(grid, kScoreThreads, 0, [=] { k_post_score<true>(d, cls, amask, A, cand, cand_count); });
        else
            pf_emul_launch(grid, kScoreThreads, 0, [=] { k_post_score<false>(d, cls, amask, A, cand, cand_count); });
    }
    int npad = 1;
    while (npad < pre_max) npad <<= 1;
    pf_emul_launch(dim3((unsigned)S), kSelectThreads, (size_t)npad * 8, [=] { k_post_select(d, A, cand, cand_count, sel_key, sel_num); });
    pf_emul_launch(dim3((unsigned)((pre_max + kDecodeThreads - 1) / kDecodeThreads), (unsigned)S), kDecodeThreads, 0, [=] {
        k_post_decode(d, A, S, sel_key, sel_num, box, dir, anchors, seg_offsets, sel_anchor, sel_score, sel_box, sel_standup,
                      sel_dir, sel_rank, nms_boxes, nms_order);
    });
    return 0;
}

// k_post_select alone on caller-made candidate lists: one class, segment f owns cand[f * A, f * A + cand_count[f])
int emul_post_select(int pre_max, int n_seg, long long A, const unsigned long long *cand, const int *cand_count,
                     unsigned long long *sel_key, int *sel_num)
{
    const int range[2] = {0, (int)A};
    const Desc d = make_desc(1, range, 1, pre_max, 1, 0.5f, nullptr);
    int npad = 1;
    while (npad < pre_max) npad <<= 1;
    pf_emul_launch(dim3((unsigned)n_seg), kSelectThreads, (size_t)npad * 8, [=] { k_post_select(d, A, cand, cand_count, sel_key, sel_num); });
    return 0;
}

int emul_post_finish(int n_classes, const int *ranges, int pre_max, int post_max, const double *limit, int n_frames,
                     const int *seg_offsets, const int *keep, const int *keep_num, const float *sel_box,
                     const float *sel_score, const int *sel_dir, const int *sel_anchor, float *det_box, float *det_score,
                     int *det_class, int *det_anchor, int *det_num)
{
    const Desc d = make_desc(n_classes, ranges, 1, pre_max, post_max, 0.f, limit);
    pf_emul_launch(dim3((unsigned)n_frames), kFinishThreads, 0, [=] {
        k_post_finish(d, seg_offsets, keep, keep_num, sel_box, sel_score, sel_dir, sel_anchor, det_box, det_score, det_class,
                      det_anchor, det_num);
    });
    return 0;
}

}  // extern "C"
