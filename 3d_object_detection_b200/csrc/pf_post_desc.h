// Launch constants and the by-value parameter block of the post-processing kernels (pf_post_kernels.cuh).  Plain C++:
// shared by the nvcc build and by the host emulation of the kernels under tests/emul/.
#pragma once

namespace pf {
namespace post {

constexpr int kMaxClasses = 8;
constexpr int kScoreThreads = 256;
constexpr int kScorePerThread = 4;
constexpr int kSelectThreads = 1024;
constexpr int kDecodeThreads = 128;
constexpr int kFinishThreads = 1024;   // >= post_max

struct Desc {
    int n_classes;
    int lo[kMaxClasses], hi[kMaxClasses];   // anchor index range of each class (AnchorAssigner.class_masks)
    int n_ch;                               // class channels per anchor (1 in the reference's head)
    int pre_max, post_max;
    float score_thr;
    double limit[6];                        // config['center_limit']
};

}  // namespace post
}  // namespace pf
