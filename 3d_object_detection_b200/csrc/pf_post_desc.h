// Launch constants and the by-value parameter block of the post-processing kernels (pf_post_kernels.cuh).  Plain C++:
// shared by the nvcc build and by the host emulation of the kernels under tests/emul/.
#pragma once

#include <math.h>

namespace pf {
namespace post {

constexpr int kMaxClasses = 8;
constexpr int kScoreThreads = 256;
constexpr int kScorePerThread = 4;
constexpr int kSelectThreads = 1024;
constexpr int kSelectBatch = 8;       // key loads a thread of k_post_select keeps in flight
constexpr int kDecodeThreads = 128;
constexpr int kFinishThreads = 1024;   // >= post_max

struct Desc {
    int n_classes;
    int lo[kMaxClasses], hi[kMaxClasses];   // anchor index range of each class (AnchorAssigner.class_masks)
    int n_ch;                               // class channels per anchor (1 in the reference's head)
    int pre_max, post_max;
    float score_thr;
    float logit_lo, logit_hi;               // logits below / above are decided without evaluating the sigmoid (set_prefilter)
    double limit[6];                        // config['center_limit']
};

// The score test is sigmoid(logit) >= score_thr, evaluated in fp32.  A logit more than 2e-3 away from the threshold's own
// logit moves the sigmoid by thr (1 - thr) 2e-3, orders of magnitude more than the few ulp a computed fp32 sigmoid can be
// off, so it is decided by comparison alone; thresholds too close to 0 or 1 for that argument get no short cut.
inline void set_prefilter(Desc &d)
{
    const double t = (double)d.score_thr;
    if (t > 1e-6 && t < 0.999) {
        const double L = log(t / (1.0 - t));
        d.logit_lo = (float)(L - 2e-3 - 1e-6 * fabs(L));
        d.logit_hi = (float)(L + 2e-3 + 1e-6 * fabs(L));
    } else {
        d.logit_lo = -INFINITY;
        d.logit_hi = INFINITY;
    }
}

}  // namespace post
}  // namespace pf
