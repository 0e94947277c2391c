// Standup-box NMS ("next" row N4): replaces nms_gpu of the reference (framework/nms.py:6-44 = host argsort,
// nms_kernel :116-147 with iou_device :102-113, host nms_postprocess :85-99), called per class and frame through
// nms() in framework/inference.py:689-703.
//
// The reference sorts on the host, uploads the boxes, builds an (n x n/64) bit mask on the device, downloads it and
// walks it on the host.  Here nothing leaves the device and a launch handles a batch of independent segments (one
// per class and frame): (1) k_nms_sort: one CTA per segment, bitonic sort of (score, index) keys in shared memory,
// emits the order and the sorted boxes; (2) k_nms_mask: the upper triangle of the 64 x 64-tiled overlap mask;
// (3) k_nms_scan: one CTA per segment walks the mask 64 boxes at a time -- the diagonal word resolves a tile in
// registers, the rows of the boxes it keeps are OR-ed into the running "removed" words by all lanes in parallel.
//
// Arithmetic: exactly the reference's.  Compiled numba types `right - left + 1` as float32 - float32 (a float32
// subtraction) + int64 -> float64, so the IoU is a float64 quotient of float32 edge differences; the comparison
// with the threshold is strict.  Score ties are ordered like a stable ascending argsort read backwards (numpy's
// default argsort is not stable, so the reference itself does not define them).
#include "pf_internal.cuh"

namespace pf {

namespace {

constexpr int kNmsTile = 64;
constexpr int kNmsSortThreads = 1024;
constexpr int kNmsScanThreads = 1024;

__device__ __forceinline__ unsigned int orderable(float s)
{
    const unsigned int u = __float_as_uint(s);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void segment_range(const int32_t *seg_offsets, long long n_total, int seg, int max_seg, int &lo,
                                              int &n, bool &too_big)
{
    lo = seg_offsets ? seg_offsets[seg] : 0;
    const long long m = seg_offsets ? (long long)seg_offsets[seg + 1] - lo : n_total;
    too_big = m > max_seg;
    n = (int)(m < 0 ? 0 : (too_big ? max_seg : m));
}

// keys sorted descending: (orderable score, index) -> highest score first, among equal scores the higher index first
__global__ void __launch_bounds__(kNmsSortThreads)
k_nms_sort(const float *__restrict__ boxes, const float *__restrict__ scores, const int32_t *__restrict__ seg_offsets,
           long long n_total, int max_seg, int32_t *__restrict__ order, float4 *__restrict__ sboxes)
{
    extern __shared__ unsigned long long s_key[];
    const int seg = blockIdx.x;
    int lo, n;
    bool too_big;
    segment_range(seg_offsets, n_total, seg, max_seg, lo, n, too_big);
    int npad = 1;
    while (npad < n) npad <<= 1;
    for (int i = threadIdx.x; i < npad; i += kNmsSortThreads)
        s_key[i] = i < n ? ((unsigned long long)orderable(scores[lo + i]) << 32) | (unsigned int)i : 0ull;
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < npad; i += kNmsSortThreads) {
                const int p = i ^ j;
                if (p > i) {
                    const unsigned long long a = s_key[i], b = s_key[p];
                    const bool desc = (i & k) == 0;
                    if (desc ? a < b : a > b) s_key[i] = b, s_key[p] = a;
                }
            }
            __syncthreads();
        }
    const float4 *b4 = reinterpret_cast<const float4 *>(boxes);
    for (int i = threadIdx.x; i < n; i += kNmsSortThreads) {
        const int src = (int)(s_key[i] & 0xffffffffu);
        order[(size_t)seg * max_seg + i] = src;
        sboxes[(size_t)seg * max_seg + i] = b4[lo + src];
    }
}

// iou_device + the threshold test.  `reject_disjoint` (threshold >= 0): when the boxes are a unit or more apart on
// an axis the reference's width or height is max(d + 1, 0) = 0, the intersection is 0 and 0 / union > t is false for
// every union (positive, zero -> NaN, negative -> -0) -- most pairs end here without touching float64.
__device__ __forceinline__ bool overlaps(const float4 a, const float4 b, double thresh, bool reject_disjoint)
{
    const float left = fmaxf(a.x, b.x), right = fminf(a.z, b.z), top = fmaxf(a.y, b.y), bottom = fminf(a.w, b.w);
    const float dw = __fsub_rn(right, left), dh = __fsub_rn(bottom, top);
    if (reject_disjoint && (dw <= -1.0f || dh <= -1.0f)) return false;
    const double width = fmax((double)dw + 1.0, 0.0);
    const double height = fmax((double)dh + 1.0, 0.0);
    const double inter = __dmul_rn(width, height);
    // Sa + Sb the way numba / NVVM compile the reference's source (tests/golden/nms_iou_ptx_ops.json): Sb rounded, Sa's
    // product contracted into the sum -- fma.rn.f64(aw, ah, Sb) -- with a = the earlier (row) box
    const double sb = __dmul_rn((double)__fsub_rn(b.z, b.x) + 1.0, (double)__fsub_rn(b.w, b.y) + 1.0);
    const double sab = __fma_rn((double)__fsub_rn(a.z, a.x) + 1.0, (double)__fsub_rn(a.w, a.y) + 1.0, sb);
    return __ddiv_rn(inter, __dsub_rn(sab, inter)) > thresh;
}

// mask[seg][i][cb]: bit t set <=> sorted box i overlaps sorted box cb*64 + t (only cb >= i/64, and t > i%64 on the
// diagonal, are ever read by the scan; the tiles below the diagonal are not computed)
__global__ void __launch_bounds__(kNmsTile)
k_nms_mask(const float4 *__restrict__ sboxes, const int32_t *__restrict__ seg_offsets, long long n_total, int max_seg,
           int mask_stride, float thresh, unsigned long long *__restrict__ mask)
{
    __shared__ float4 s_col[kNmsTile];
    const int seg = blockIdx.z, row_b = blockIdx.y, col_b = blockIdx.x;
    if (col_b < row_b) return;
    int lo, n;
    bool too_big;
    segment_range(seg_offsets, n_total, seg, max_seg, lo, n, too_big);
    if (row_b * kNmsTile >= n || col_b * kNmsTile >= n) return;
    const int tx = threadIdx.x;
    const int col_size = min(n - col_b * kNmsTile, kNmsTile), row_size = min(n - row_b * kNmsTile, kNmsTile);
    const float4 *sb = sboxes + (size_t)seg * max_seg;
    if (tx < col_size) s_col[tx] = sb[col_b * kNmsTile + tx];
    __syncthreads();
    if (tx >= row_size) return;
    const int cur = row_b * kNmsTile + tx;
    const float4 a = sb[cur];
    const double t = (double)thresh;
    const bool reject_disjoint = thresh >= 0.0f;
    unsigned long long bits = 0;
    for (int i = (row_b == col_b) ? tx + 1 : 0; i < col_size; ++i)
        if (overlaps(a, s_col[i], t, reject_disjoint)) bits |= 1ull << i;
    mask[((size_t)seg * max_seg + cur) * mask_stride + col_b] = bits;
}

// nms_postprocess on the device.  Warp 0 walks the tiles; the other warps only help to stage the mask in shared
// memory when the segment's triangle fits (n <= ~1500), which takes the global-memory latency out of the walk.
__global__ void __launch_bounds__(kNmsScanThreads)
k_nms_scan(const unsigned long long *__restrict__ mask, const int32_t *__restrict__ order,
           const int32_t *__restrict__ seg_offsets, long long n_total, int max_seg, int mask_stride, int post_max,
           int smem_words, int32_t *__restrict__ keep, int32_t *__restrict__ keep_num)
{
    extern __shared__ unsigned long long s_mask[];
    const int seg = blockIdx.x;
    int lo, n;
    bool too_big;
    segment_range(seg_offsets, n_total, seg, max_seg, lo, n, too_big);
    const int nb = (n + kNmsTile - 1) / kNmsTile;
    const unsigned long long *M = mask + (size_t)seg * max_seg * mask_stride;
    int stride = mask_stride;
    if ((long long)n * nb <= smem_words) {        // stage rows [0, n) x column blocks [0, nb)
        const int words = n * nb;
        for (int w0 = threadIdx.x; w0 < words; w0 += 4 * kNmsScanThreads) {     // four independent loads in flight
            unsigned long long v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int w = w0 + u * kNmsScanThreads;
                const int i = w / nb, j = w - i * nb;
                v[u] = (w < words && j >= i / kNmsTile) ? M[(size_t)i * mask_stride + j] : 0ull;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (w0 + u * kNmsScanThreads < words) s_mask[w0 + u * kNmsScanThreads] = v[u];
        }
        __syncthreads();
        M = s_mask;
        stride = nb;
    }
    if (threadIdx.x >= 32) return;
    const int lane = threadIdx.x;
    const int32_t *ord = order + (size_t)seg * max_seg;
    int32_t *out = keep + (size_t)seg * post_max;
    // removed-words of column blocks lane, lane + 32, ... (max_seg <= 4096 -> at most two per lane)
    unsigned long long remv0 = 0, remv1 = 0;
    int count = 0;
    for (int b = 0; b < nb && count < post_max; ++b) {
        const int base = b * kNmsTile;
        const int valid = min(n - base, kNmsTile);
        // diagonal words of rows base + lane and base + lane + 32
        const unsigned long long d0 = lane < valid ? M[(size_t)(base + lane) * stride + b] : 0ull;
        const unsigned long long d1 = lane + 32 < valid ? M[(size_t)(base + lane + 32) * stride + b] : 0ull;
        // original indices of this tile's boxes: loaded now so that the latency hides under the walk
        const int o0 = lane < valid ? ord[base + lane] : -1;
        const int o1 = lane + 32 < valid ? ord[base + lane + 32] : -1;
        unsigned long long removed = __shfl_sync(0xffffffffu, b < 32 ? remv0 : remv1, b & 31);
        if (valid < kNmsTile) removed |= ~0ull << valid;
        unsigned long long kept = 0;
        while (~removed) {                             // every lane runs the same walk on the same words
            const int t = __ffsll((long long)~removed) - 1;
            kept |= 1ull << t;
            const unsigned long long lo_w = __shfl_sync(0xffffffffu, d0, t & 31);
            const unsigned long long hi_w = __shfl_sync(0xffffffffu, d1, t & 31);
            removed |= (t < 32 ? lo_w : hi_w) | (1ull << t);
        }
        // the rows of the kept boxes suppress later column blocks: lane owns blocks lane and lane + 32
        unsigned long long k = kept;
        while (k) {
            const int t = __ffsll((long long)k) - 1;
            k &= k - 1;
            const unsigned long long *row = M + (size_t)(base + t) * stride;
            if (lane > b && lane < nb) remv0 |= row[lane];
            if (lane + 32 > b && lane + 32 < nb) remv1 |= row[lane + 32];
        }
        // emit kept boxes in order
        const int before_lo = __popcll(kept & ((1ull << lane) - 1ull));
        const int n_lo = __popcll(kept & 0xffffffffull);
        if ((kept >> lane) & 1ull) {
            const int pos = count + before_lo;
            if (pos < post_max) out[pos] = o0;
        }
        if ((kept >> (lane + 32)) & 1ull) {
            const int pos = count + n_lo + __popcll((kept >> 32) & ((1ull << lane) - 1ull));
            if (pos < post_max) out[pos] = o1;
        }
        count += __popcll(kept);
    }
    count = min(count, post_max);
    for (int i = count + lane; i < post_max; i += 32) out[i] = -1;
    if (lane == 0) keep_num[seg] = too_big ? -1 : count;
}

}  // namespace

size_t nms_scratch_bytes(int n_seg, int max_seg)
{
    const size_t rows = (size_t)n_seg * max_seg;
    const size_t cb = (size_t)(max_seg + kNmsTile - 1) / kNmsTile;
    return rows * sizeof(float4) + rows * sizeof(int32_t) + rows * cb * sizeof(unsigned long long) + 512;
}

namespace {

struct NmsViews {
    float4 *sboxes;
    unsigned long long *mask;
    int32_t *order;
    int cb;
};

NmsViews nms_views(void *scratch, int n_seg, int max_seg)
{
    const size_t rows = (size_t)n_seg * max_seg;
    NmsViews v;
    v.cb = (max_seg + kNmsTile - 1) / kNmsTile;
    char *base = static_cast<char *>(scratch);
    v.sboxes = reinterpret_cast<float4 *>(base);
    v.mask = reinterpret_cast<unsigned long long *>(base + ((rows * sizeof(float4) + 255) & ~(size_t)255));
    v.order = reinterpret_cast<int32_t *>(reinterpret_cast<char *>(v.mask) + ((rows * v.cb * 8 + 255) & ~(size_t)255));
    return v;
}

// overlap mask + greedy walk on sorted boxes (v.sboxes / v.order filled by k_nms_sort or by the caller)
cudaError_t launch_mask_scan(const NmsViews &v, const int32_t *seg_offsets, int n_seg, long long n_total, int max_seg,
                             float thresh, int post_max, int32_t *keep, int32_t *keep_num, cudaStream_t stream)
{
    k_nms_mask<<<dim3(v.cb, v.cb, n_seg), kNmsTile, 0, stream>>>(v.sboxes, seg_offsets, n_total, max_seg, v.cb, thresh, v.mask);
    // shared-memory staging of the mask: up to 200 KB (opt-in above 48 KB)
    constexpr int kScanSmem = 200 * 1024;
    cudaError_t e = cudaFuncSetAttribute(k_nms_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, kScanSmem);   // per device
    if (e != cudaSuccess) return e;
    const long long need = (long long)max_seg * v.cb * 8;
    const int smem = (int)(need < kScanSmem ? need : kScanSmem);
    k_nms_scan<<<n_seg, kNmsScanThreads, smem, stream>>>(v.mask, v.order, seg_offsets, n_total, max_seg, v.cb, post_max,
                                                        smem / 8, keep, keep_num);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_nms(const float *boxes, const float *scores, const int32_t *seg_offsets, int n_seg, long long n_total,
                       int max_seg, float thresh, int post_max, void *scratch, int32_t *keep, int32_t *keep_num,
                       cudaStream_t stream)
{
    const NmsViews v = nms_views(scratch, n_seg, max_seg);
    int npad = 1;
    while (npad < max_seg) npad <<= 1;
    k_nms_sort<<<n_seg, kNmsSortThreads, (size_t)npad * 8, stream>>>(boxes, scores, seg_offsets, n_total, max_seg, v.order,
                                                                    v.sboxes);
    return launch_mask_scan(v, seg_offsets, n_seg, n_total, max_seg, thresh, post_max, keep, keep_num, stream);
}

// Where a caller that already holds its boxes in NMS order writes them instead of running the sort: sorted boxes
// [n_seg][max_seg][4] and the index each one reports when kept [n_seg][max_seg].
void nms_presorted_views(void *scratch, int n_seg, int max_seg, float **sorted_boxes, int32_t **order)
{
    const NmsViews v = nms_views(scratch, n_seg, max_seg);
    *sorted_boxes = reinterpret_cast<float *>(v.sboxes);
    *order = v.order;
}

cudaError_t launch_nms_presorted(const int32_t *seg_offsets, int n_seg, long long n_total, int max_seg, float thresh,
                                 int post_max, void *scratch, int32_t *keep, int32_t *keep_num, cudaStream_t stream)
{
    return launch_mask_scan(nms_views(scratch, n_seg, max_seg), seg_offsets, n_seg, n_total, max_seg, thresh, post_max, keep,
                            keep_num, stream);
}

}  // namespace pf
