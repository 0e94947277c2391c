// Kernels of the detection post-processing front ("next" row N4, the part in front of pf_nms and behind it):
// the per-class loop of Inference.infer_gpu, framework/inference.py:26-138 --
//   anchor mask (:44-47) -> sigmoid + max over the class channels (:50-54) -> score mask (:58-64) -> top-k (:67-69)
//   -> box_decode (framework/box_np_ops.py:406-424) -> centre/size/angle to corners to standup boxes
//   (center_to_corner_box2d :81-99 with corners_nd :122-153 and rotation_2d :64-78; corner_to_standup_nd :899-904)
//   -> [pf_nms] -> direction flip (:100-104), range mask (:106-109), limit_period (:113, box_np_ops.py:102-103).
//
// The file is written against a small vocabulary (threadIdx / blockIdx / __syncthreads / atomicAdd on shared and global
// words, no warp intrinsics) so that tests/emul/ can compile the very same kernel bodies with g++ and run them on host
// threads (PF_HOST_EMUL): the build container has no GPU, and the emulation is how the kernel logic is checked there.
// The emulation is test infrastructure; the product only ever runs the nvcc build.
//
// Order of the selected boxes.  The reference orders by the sigmoid score (torch.topk, then a descending argsort inside
// nms_gpu) and leaves ties undefined.  sigmoid is monotone, so ordering by the raw logit (descending) and then by the
// anchor index (ascending) is one of the orders the reference allows; it is exact integer work that does not depend on
// anybody's exp(), and it is what the oracle states.  Only the threshold test `score >= nms_score_threshold` uses the
// computed score.
#pragma once

#include <stdint.h>

#ifdef PF_HOST_EMUL
#include "cuda_emul.h"
#else
#include <cuda_runtime.h>
#endif

#include "pf_post_desc.h"

namespace pf {
namespace post {

// ---- fp32 arithmetic pinned to IEEE round-to-nearest, no contraction (numpy evaluates one ufunc per operator) --------
#ifdef PF_HOST_EMUL
static inline float f_add(float a, float b) { return a + b; }
static inline float f_sub(float a, float b) { return a - b; }
static inline float f_mul(float a, float b) { return a * b; }
static inline float f_div(float a, float b) { return a / b; }
static inline float f_sqrt(float a) { return sqrtf(a); }
#define PF_DEV static inline
#define PF_DYN_SHARED(T, name) T *name = reinterpret_cast<T *>(pf_emul_dyn_smem)
#else
__device__ __forceinline__ float f_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float f_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float f_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float f_div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float f_sqrt(float a) { return __fsqrt_rn(a); }
#define PF_DEV __device__ __forceinline__
#define PF_DYN_SHARED(T, name) extern __shared__ T name[]
#endif

PF_DEV unsigned int f2u(float f)
{
#ifdef PF_HOST_EMUL
    unsigned int u;
    memcpy(&u, &f, 4);
    return u;
#else
    return __float_as_uint(f);
#endif
}

// monotone map float -> uint32 (a < b <=> orderable(a) < orderable(b), -0 just below +0)
PF_DEV unsigned int orderable(float s)
{
    const unsigned int u = f2u(s);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

PF_DEV float from_orderable(unsigned int o)
{
    const unsigned int u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef PF_HOST_EMUL
    float f;
    memcpy(&f, &u, 4);
    return f;
#else
    return __uint_as_float(u);
#endif
}

// torch.sigmoid in fp32: 1 / (1 + exp(-x))
PF_DEV float sigmoid(float x) { return f_div(1.0f, f_add(1.0f, expf(-x))); }

PF_DEV int class_of(const Desc &d, int a)
{
    for (int c = 0; c < d.n_classes; ++c)
        if (a >= d.lo[c] && a < d.hi[c]) return c;
    return -1;
}

// candidate key: (orderable logit, ~local anchor index) -- larger key = better candidate; keys of a segment are unique
PF_DEV unsigned long long make_key(float logit, int local) { return ((unsigned long long)orderable(logit) << 32) | (unsigned int)~(unsigned int)local; }

// ---- 1. score: anchor mask, max logit over the class channels, score threshold, candidate lists per (frame, class) ----
// grid (ceil(A / (256 * 4)), n_frames); a thread owns 4 consecutive anchors.  VEC (one class channel, A % 4 == 0, aligned
// pointers): one 16-byte load of the logits and one 4-byte load of the mask per thread.  The exact test is
// sigmoid(logit) >= threshold; logits outside [logit_lo, logit_hi] (the threshold's logit -+ 2e-3, far more than the few
// ulp a computed sigmoid can be off) are decided without evaluating it, which is almost all of them.
// cand holds n_frames * A keys: segment (f, c) owns [f * A + lo[c], f * A + hi[c]).  cand_count must be zero on entry.
template <bool VEC>
__global__ void __launch_bounds__(kScoreThreads)
k_post_score(Desc d, const float *__restrict__ cls, const uint8_t *__restrict__ amask, long long A,
             unsigned long long *__restrict__ cand, int *__restrict__ cand_count)
{
    __shared__ int s_cnt[kMaxClasses], s_base[kMaxClasses];
    const int f = blockIdx.y, tid = threadIdx.x;
    if (tid < kMaxClasses) s_cnt[tid] = 0;
    __syncthreads();
    const long long a0 = ((long long)blockIdx.x * kScoreThreads + tid) * kScorePerThread;
    float logit[kScorePerThread];
    bool ok[kScorePerThread];
    if (VEC) {
        const bool in = a0 < A;                       // A % 4 == 0: the four anchors are in range together
        float4 v = {0.f, 0.f, 0.f, 0.f};
        uchar4 m = {1, 1, 1, 1};
        if (in) {
            v = *reinterpret_cast<const float4 *>(cls + (long long)f * A + a0);
            if (amask) m = *reinterpret_cast<const uchar4 *>(amask + (long long)f * A + a0);
        }
        logit[0] = v.x, logit[1] = v.y, logit[2] = v.z, logit[3] = v.w;
        ok[0] = in && m.x, ok[1] = in && m.y, ok[2] = in && m.z, ok[3] = in && m.w;
    } else {
#pragma unroll
        for (int j = 0; j < kScorePerThread; ++j) {
            const long long a = a0 + j;
            ok[j] = a < A && (!amask || amask[(long long)f * A + a]);
            logit[j] = 0.f;
            if (ok[j]) {
                const float *p = cls + ((long long)f * A + a) * d.n_ch;
                float top = p[0];
                bool nan = top != top;
                for (int k = 1; k < d.n_ch; ++k) {
                    const float v = p[k];
                    nan = nan || v != v;
                    if (v > top) top = v;
                }
                logit[j] = top;
                ok[j] = !nan;                        // NaN >= threshold is false in the reference as well
            }
        }
    }
    int cl[kScorePerThread], rk[kScorePerThread];
#pragma unroll
    for (int j = 0; j < kScorePerThread; ++j) {
        cl[j] = -1;
        rk[j] = 0;
        if (!ok[j]) continue;
        const float x = logit[j];
        const bool pass = x > d.logit_hi || (x >= d.logit_lo && sigmoid(x) >= d.score_thr);     // false for NaN
        if (!pass) continue;
        const int c = class_of(d, (int)(a0 + j));
        if (c < 0) continue;
        cl[j] = c;
        rk[j] = atomicAdd(&s_cnt[c], 1);
    }
    __syncthreads();
    if (tid < d.n_classes && s_cnt[tid] > 0) s_base[tid] = atomicAdd(&cand_count[f * d.n_classes + tid], s_cnt[tid]);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kScorePerThread; ++j)
        if (cl[j] >= 0)
            cand[(long long)f * A + d.lo[cl[j]] + s_base[cl[j]] + rk[j]] = make_key(logit[j], (int)(a0 + j) - d.lo[cl[j]]);
}

// ---- 2. select: the pre_max largest keys of every segment, sorted descending -------------------------------------------
// One CTA per segment.  More candidates than pre_max: MSB-first radix select (8 bits a pass) finds the pre_max-th largest
// key T, then the keys >= T (exactly pre_max of them: keys are unique) are collected.  A pass ends the search as soon as
// the boundary bin is needed whole (typically once the bits that tell the logits apart are consumed: 3-4 passes).  The
// upper bits of the keys are nearly constant, so a plain shared-memory histogram would serialise on one bin: each thread
// counts runs of equal digits in registers and issues one atomic per run.  One CTA reads its segment's keys through one SM,
// so every sweep keeps kSelectBatch independent 8-byte loads per thread in flight.  Bitonic sort in shared memory.
// dynamic shared memory: npad * 8 bytes, npad = pre_max rounded up to a power of two.
__global__ void __launch_bounds__(kSelectThreads)
k_post_select(Desc d, long long A, const unsigned long long *__restrict__ cand, const int *__restrict__ cand_count,
              unsigned long long *__restrict__ sel_key, int *__restrict__ sel_num)
{
    PF_DYN_SHARED(unsigned long long, s_keys);
    __shared__ unsigned int s_hist[256];
    __shared__ unsigned int s_grp[16];
    __shared__ unsigned long long s_prefix, s_T;
    __shared__ int s_krem, s_n, s_done;
    const int seg = blockIdx.x, tid = threadIdx.x;
    const int f = seg / d.n_classes, c = seg - f * d.n_classes;
    const unsigned long long *src = cand + (long long)f * A + d.lo[c];
    const int M = cand_count[seg];
    const int k = M < d.pre_max ? M : d.pre_max;
    int npad = 1;
    while (npad < k) npad <<= 1;                     // <= pre_max rounded up: the launch's shared memory covers it
    unsigned long long T = 0;
    if (M > d.pre_max) {
        if (tid == 0) s_prefix = 0, s_krem = k, s_done = 0, s_T = 0;
        for (int pass = 0; pass < 8; ++pass) {
            const int shift = 56 - 8 * pass;
            if (tid < 256) s_hist[tid] = 0;
            __syncthreads();
            const unsigned long long prefix = s_prefix;
            int run_d = -1;
            unsigned int run_n = 0;
            for (int i0 = tid; i0 < M; i0 += kSelectThreads * kSelectBatch) {
                unsigned long long kb[kSelectBatch];          // all loads of a batch in flight before the first use
#pragma unroll
                for (int u = 0; u < kSelectBatch; ++u) {
                    const int i = i0 + u * kSelectThreads;
                    kb[u] = i < M ? src[i] : 0ull;
                }
#pragma unroll
                for (int u = 0; u < kSelectBatch; ++u) {
                    const unsigned long long key = kb[u];
                    if (i0 + u * kSelectThreads < M && (pass == 0 || (key >> (shift + 8)) == prefix)) {
                        const int dg = (int)((key >> shift) & 255ull);
                        if (dg == run_d) ++run_n;
                        else {
                            if (run_n) atomicAdd(&s_hist[run_d], run_n);
                            run_d = dg, run_n = 1;
                        }
                    }
                }
            }
            if (run_n) atomicAdd(&s_hist[run_d], run_n);
            __syncthreads();
            if (tid < 16) {
                unsigned int g = 0;
                for (int b = 0; b < 16; ++b) g += s_hist[tid * 16 + b];
                s_grp[tid] = g;
            }
            __syncthreads();
            if (tid == 0) {
                unsigned int acc = 0, krem = (unsigned int)s_krem;
                int g = 15;
                while (g > 0 && acc + s_grp[g] < krem) acc += s_grp[g], --g;
                int b = g * 16 + 15;
                while (b > g * 16 && acc + s_hist[b] < krem) acc += s_hist[b], --b;
                krem -= acc;
                const unsigned long long next = (prefix << 8) | (unsigned long long)b;
                if (s_hist[b] == krem || shift == 0) s_T = next << shift, s_done = 1;     // the whole boundary bin is wanted
                s_krem = (int)krem;
                s_prefix = next;
            }
            __syncthreads();
            if (s_done) break;
        }
        T = s_T;
    }
    if (tid == 0) s_n = 0;
    for (int i = tid; i < npad; i += kSelectThreads) s_keys[i] = 0;
    __syncthreads();
    for (int i0 = tid; i0 < M; i0 += kSelectThreads * kSelectBatch) {
        unsigned long long kb[kSelectBatch];
#pragma unroll
        for (int u = 0; u < kSelectBatch; ++u) {
            const int i = i0 + u * kSelectThreads;
            kb[u] = i < M ? src[i] : 0ull;
        }
#pragma unroll
        for (int u = 0; u < kSelectBatch; ++u)
            if (i0 + u * kSelectThreads < M && kb[u] >= T) {
                const int pos = atomicAdd(&s_n, 1);
                if (pos < npad) s_keys[pos] = kb[u];
            }
    }
    __syncthreads();
    for (int kk = 2; kk <= npad; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < npad; i += kSelectThreads) {
                const int p = i ^ j;
                if (p > i) {
                    const unsigned long long a = s_keys[i], b = s_keys[p];
                    const bool desc = (i & kk) == 0;
                    if (desc ? a < b : a > b) s_keys[i] = b, s_keys[p] = a;
                }
            }
            __syncthreads();
        }
    for (int i = tid; i < k; i += kSelectThreads) sel_key[(long long)seg * d.pre_max + i] = s_keys[i];
    if (tid == 0) sel_num[seg] = k;
}

// box_decode + standup box of one selected anchor.  enc / anc: 7 floats (x, y, z, l, w, h, r).
PF_DEV void decode_one(const float *enc, const float *anc, float *box, float *standup)
{
    const float xa = anc[0], ya = anc[1], la = anc[3], wa = anc[4], ha = anc[5], ra = anc[6];
    const float za = f_add(anc[2], f_div(ha, 2.0f));
    const float diagonal = f_sqrt(f_add(f_mul(la, la), f_mul(wa, wa)));
    const float xg = f_add(f_mul(enc[0], diagonal), xa);
    const float yg = f_add(f_mul(enc[1], diagonal), ya);
    float zg = f_add(f_mul(enc[2], ha), za);
    const float lg = f_mul(expf(enc[3]), la);
    const float wg = f_mul(expf(enc[4]), wa);
    const float hg = f_mul(expf(enc[5]), ha);
    const float rg = f_add(enc[6], ra);
    zg = f_sub(zg, f_div(hg, 2.0f));
    box[0] = xg, box[1] = yg, box[2] = zg, box[3] = lg, box[4] = wg, box[5] = hg, box[6] = rg;
    // corners_nd: (l, w) * {(-.5,-.5), (-.5,.5), (.5,.5), (.5,-.5)}; rotation_2d: x' = x cos - y sin, y' = x sin + y cos
    const float s = sinf(rg), co = cosf(rg);
    const float hx = f_mul(lg, 0.5f), hy = f_mul(wg, 0.5f);
    const float px[4] = {-hx, -hx, hx, hx}, py[4] = {-hy, hy, hy, -hy};
    float x0 = 0.f, y0 = 0.f, x1 = 0.f, y1 = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float cx = f_add(f_add(f_mul(px[q], co), f_mul(py[q], -s)), xg);
        const float cy = f_add(f_add(f_mul(px[q], s), f_mul(py[q], co)), yg);
        if (q == 0) x0 = x1 = cx, y0 = y1 = cy;
        else {
            x0 = fminf(x0, cx), x1 = fmaxf(x1, cx);
            y0 = fminf(y0, cy), y1 = fmaxf(y1, cy);
        }
    }
    standup[0] = x0, standup[1] = y0, standup[2] = x1, standup[3] = y1;
}

// ---- 3. decode: gather the selected anchors, decode, standup boxes; compact layout for pf_nms ------------------------
// grid (ceil(pre_max / 128), n_segments).  Outputs are indexed seg_offsets[seg] + rank; seg_offsets [n_seg + 1] is
// written by block (0, 0).  The boxes are already in NMS order, so the kernel also writes what pf_nms's sort would produce
// (nms_boxes [seg][pre_max] float4 = the standup boxes, nms_order [seg][pre_max] = the identity) and the sort is skipped;
// out_rank = pre_max - rank is the strictly decreasing "score" that would make pf_nms's own sort keep this order.
__global__ void __launch_bounds__(kDecodeThreads)
k_post_decode(Desc d, long long A, int n_seg, const unsigned long long *__restrict__ sel_key, const int *__restrict__ sel_num,
              const float *__restrict__ box_preds, const float *__restrict__ dir_preds, const float *__restrict__ anchors,
              int *__restrict__ seg_offsets, int *__restrict__ out_anchor, float *__restrict__ out_score,
              float *__restrict__ out_box, float *__restrict__ out_standup, int *__restrict__ out_dir,
              float *__restrict__ out_rank, float *__restrict__ nms_boxes, int *__restrict__ nms_order)
{
    __shared__ int s_off;
    const int seg = blockIdx.y, tid = threadIdx.x;
    if (tid == 0) {
        int off = 0;
        for (int s = 0; s < seg; ++s) off += sel_num[s];
        s_off = off;
        if (blockIdx.x == 0 && seg == 0) {
            int run = 0;
            for (int s = 0; s < n_seg; ++s) seg_offsets[s] = run, run += sel_num[s];
            seg_offsets[n_seg] = run;
        }
    }
    __syncthreads();
    const int i = blockIdx.x * kDecodeThreads + tid;
    if (i >= sel_num[seg]) return;
    const int f = seg / d.n_classes, c = seg - f * d.n_classes;
    const unsigned long long key = sel_key[(long long)seg * d.pre_max + i];
    const int local = (int)~(unsigned int)(key & 0xffffffffull);
    const long long a = (long long)d.lo[c] + local;
    const float logit = from_orderable((unsigned int)(key >> 32));
    float enc[7], anc[7], box[7], su[4];
#pragma unroll
    for (int q = 0; q < 7; ++q) enc[q] = box_preds[((long long)f * A + a) * 7 + q], anc[q] = anchors[a * 7 + q];
    decode_one(enc, anc, box, su);
    const float d0 = dir_preds[((long long)f * A + a) * 2], d1 = dir_preds[((long long)f * A + a) * 2 + 1];
    const long long o = (long long)s_off + i;
    out_anchor[o] = (int)a;
    out_score[o] = sigmoid(logit);
    out_dir[o] = d1 > d0 ? 1 : 0;                 // torch.max(dim=-1)[1]: first maximum on a tie
    out_rank[o] = (float)(d.pre_max - i);
#pragma unroll
    for (int q = 0; q < 7; ++q) out_box[o * 7 + q] = box[q];
#pragma unroll
    for (int q = 0; q < 4; ++q) out_standup[o * 4 + q] = su[q], nms_boxes[((long long)seg * d.pre_max + i) * 4 + q] = su[q];
    nms_order[(long long)seg * d.pre_max + i] = i;
}

// ---- 4. finish: what infer_gpu does with the kept boxes (:97-121), per frame, classes in class_masks order ------------
// One CTA per frame walks the kept boxes of all its classes as one list (class-major, NMS order inside a class), 1024 at
// a time: direction flip, range mask, limit_period, then an ordered compaction through 32 bit-words of pass flags.
// det_* rows of frame f start at f * n_classes * post_max; det_num[f] rows are valid.
__global__ void __launch_bounds__(kFinishThreads)
k_post_finish(Desc d, const int *__restrict__ seg_offsets, const int *__restrict__ keep, const int *__restrict__ keep_num,
              const float *__restrict__ in_box, const float *__restrict__ in_score, const int *__restrict__ in_dir,
              const int *__restrict__ in_anchor, float *__restrict__ det_box, float *__restrict__ det_score,
              int *__restrict__ det_class, int *__restrict__ det_anchor, int *__restrict__ det_num)
{
    __shared__ unsigned char s_flag[kFinishThreads];
    __shared__ unsigned int s_word[kFinishThreads / 32];
    __shared__ int s_start[kMaxClasses + 1];
    const int f = blockIdx.x, tid = threadIdx.x;
    const long long row0 = (long long)f * d.n_classes * d.post_max;
    if (tid == 0) {
        int run = 0;
        for (int c = 0; c < d.n_classes; ++c) {
            int n = keep_num[f * d.n_classes + c];
            n = n < 0 ? 0 : (n > d.post_max ? d.post_max : n);
            s_start[c] = run, run += n;
        }
        s_start[d.n_classes] = run;
    }
    __syncthreads();
    const int total = s_start[d.n_classes];
    int base = 0;
    for (int e0 = 0; e0 < total; e0 += kFinishThreads) {
        const int e = e0 + tid;
        bool pass = false;
        float b[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        long long g = 0;
        int c = 0;
        if (e < total) {
            while (e >= s_start[c + 1]) ++c;
            const int seg = f * d.n_classes + c;
            g = (long long)seg_offsets[seg] + keep[(long long)seg * d.post_max + (e - s_start[c])];
#pragma unroll
            for (int q = 0; q < 7; ++q) b[q] = in_box[g * 7 + q];
            // box_preds[..., -1] += np.where(opp_labels, np.pi, 0.0): float64 addend, result cast back to float32
            const bool opp = (b[6] > 0.f) != (in_dir[g] != 0);
            b[6] = (float)((double)b[6] + (opp ? 3.141592653589793 : 0.0));
            const bool min_mask = (double)b[0] > d.limit[0] || (double)b[1] > d.limit[1] || (double)b[2] > d.limit[2];
            const bool max_mask = (double)b[3] < d.limit[3] || (double)b[4] < d.limit[4] || (double)b[5] < d.limit[5];
            pass = min_mask && max_mask;
            // limit_period(val, offset=0.5, period=2*pi), all float32 (the Python scalars are weak)
            const float period = 6.2831853071795862f;
            b[6] = f_sub(b[6], f_mul(floorf(f_add(f_div(b[6], period), 0.5f)), period));
        }
        s_flag[tid] = pass ? 1 : 0;
        __syncthreads();
        if (tid < kFinishThreads / 32) {
            unsigned int w = 0;
            for (int q = 0; q < 32; ++q) w |= (unsigned int)s_flag[tid * 32 + q] << q;
            s_word[tid] = w;
        }
        __syncthreads();
        int before = 0, all = 0;
        for (int w = 0; w < kFinishThreads / 32; ++w) {
            const int n = __popc(s_word[w]);
            all += n;
            if (w < (tid >> 5)) before += n;
        }
        if (pass) {
            const long long o = row0 + base + before + __popc(s_word[tid >> 5] & ((1u << (tid & 31)) - 1u));
#pragma unroll
            for (int q = 0; q < 7; ++q) det_box[o * 7 + q] = b[q];
            det_score[o] = in_score[g];
            det_class[o] = c;
            if (det_anchor) det_anchor[o] = in_anchor[g];
        }
        base += all;
        __syncthreads();
    }
    if (tid == 0) det_num[f] = base;
}

}  // namespace post
}  // namespace pf
