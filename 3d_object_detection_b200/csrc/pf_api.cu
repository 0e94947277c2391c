// C-ABI layer of libpillarfront.so: handle life-cycle, argument validation, error reporting.
// Declarations and the reference functions each entry point replaces: include/pillarfront.h.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include <omp.h>
#if defined(__x86_64__)
#include <emmintrin.h>   // SSE2 (baseline x86-64): non-temporal stores for the staging copy
#endif

#include <nvtx3/nvToolsExt.h>   // header-only; a no-op unless a profiler (nsys / ncu) is attached

#include "pf_internal.cuh"

struct pf_handle {
    pf::Params p;
    pf::Work w;
    pf_grid_desc desc;
    int device;
    int max_frames;
    int max_points_per_frame;
    size_t bytes;
    void *slab;   // one allocation, carved into the Work arrays
    cudaEvent_t events[pf::kStageEvents];
    bool timing;
    int map_pending;   // frames whose grid still holds the cell -> voxel map of a keep_cell_map call
};

namespace {

thread_local std::string g_err = "";

int fail(int code, const char *fmt, ...) __attribute__((format(printf, 2, 3)));
int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

int cuda_fail(cudaError_t e, const char *what)
{
    g_err = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return (int)e;
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// 0 when the gather / decoration kernel for (P, F) fits the opt-in shared memory of `device` (< 0: current), else the
// PF_ERR_UNSUPPORTED status (message set)
int check_gather_shape(int P, int F, int device, const char *who)
{
    if (device < 0 && cudaGetDevice(&device) != cudaSuccess) return 0;   // no device yet: the launch will report it
    int limit = 0;
    if (cudaDeviceGetAttribute(&limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, device) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    const size_t need = pf::gather_smem_needed(P, F);
    if (need > (size_t)limit)
        return fail(PF_ERR_UNSUPPORTED, "%s: max_points %d with %d floats per point needs %zu bytes of shared memory per CTA, the "
                                        "device allows %d", who, P, F, need, limit);
    return 0;
}

// NVTX range around the enqueue of one C-ABI call (keep the path visible in traces)
struct Range {
    explicit Range(const char *name) { nvtxRangePushA(name); }
    ~Range() { nvtxRangePop(); }
};

struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard()
    {
        if (switched) cudaSetDevice(prev);
    }
};

}  // namespace

extern "C" {

int pf_version(void) { return PF_VERSION; }

const char *pf_last_error_string(void) { return g_err.c_str(); }

int pf_create(const pf_grid_desc *d, int max_frames, int max_points_per_frame, int device, pf_handle **out)
{
    if (!d || !out) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: null argument");
    *out = nullptr;
    if (max_frames < 1 || max_frames > 65535) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: max_frames %d out of [1, 65535]", max_frames);
    if (max_points_per_frame < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: max_points_per_frame must be >= 1");
    for (int k = 0; k < 3; ++k) {
        if (!(d->voxel_size[k] > 0.f)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: voxel_size[%d] must be > 0", k);
        if (d->grid[k] < 1 || d->grid[k] >= (1 << 24)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: grid[%d]=%d out of [1, 2^24)", k, d->grid[k]);
    }
    const long long cells = (long long)d->grid[0] * d->grid[1] * d->grid[2];
    if (cells >= (1ll << 31)) return fail(PF_ERR_UNSUPPORTED, "pf_create: %lld cells do not fit int32", cells);
    if (d->max_voxels < 1 || d->max_voxels >= (1 << 23)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: max_voxels %d out of [1, 2^23)", d->max_voxels);
    if (d->max_points < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_create: max_points must be >= 1");
    if (d->max_points > pf::kCountCap) return fail(PF_ERR_UNSUPPORTED, "pf_create: max_points %d > %d", d->max_points, pf::kCountCap);
    if (d->num_features < 3 || d->num_features > 16) return fail(PF_ERR_UNSUPPORTED, "pf_create: num_features %d outside 3..16", d->num_features);

    if (device < 0) {
        cudaError_t e = cudaGetDevice(&device);
        if (e != cudaSuccess) return cuda_fail(e, "pf_create: cudaGetDevice");
    }
    DeviceGuard guard(device);
    if (int rc = check_gather_shape(d->max_points, d->num_features, device, "pf_create")) return rc;

    pf_handle *h = new (std::nothrow) pf_handle();
    if (!h) return fail(PF_ERR_OUT_OF_MEMORY, "pf_create: host allocation failed");
    h->desc = *d;
    h->device = device;
    h->max_frames = max_frames;
    h->max_points_per_frame = max_points_per_frame;
    pf::Params &p = h->p;
    for (int k = 0; k < 3; ++k) p.off[k] = d->offset[k], p.vs[k] = d->voxel_size[k];
    p.nx = d->grid[0], p.ny = d->grid[1], p.nz = d->grid[2];
    p.cells = (int)cells;
    p.plane = d->grid[0] * d->grid[1];
    p.max_voxels = d->max_voxels;
    p.P = d->max_points;
    p.F = d->num_features;
    p.npts_cap = (int)align_up((size_t)max_points_per_frame, pf::kTile);
    p.tiles_cap = p.npts_cap / pf::kTile;
    p.vcols = (int)align_up((size_t)d->max_voxels, 16);
    p.ny_m = (unsigned)((1ull << 32) / (unsigned)p.ny);   // ny, nz == 1 are short-cut in div_small
    p.nz_m = (unsigned)((1ull << 32) / (unsigned)p.nz);
    p.vbits = 0;
    while ((1 << p.vbits) < d->max_voxels) ++p.vbits;
    if ((size_t)p.vcols > 200 * 1024) { delete h; return fail(PF_ERR_UNSUPPORTED, "pf_create: max_voxels %d too large for the tile table", d->max_voxels); }

    const size_t B = (size_t)max_frames;
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    const size_t o_grid = carve(B * p.cells * 4), o_cell = carve(B * p.npts_cap * 4), o_vr = carve(B * p.npts_cap * 4);
    const size_t o_tc = carve(B * p.tiles_cap * 4), o_tk = carve(B * 4), o_cnt = carve(B * p.tiles_cap * (size_t)p.vcols);
    const size_t o_slot = carve(B * p.max_voxels * (size_t)p.P * 4), o_vcell = carve(B * p.max_voxels * 4);
    const size_t o_vcount = carve(B * p.max_voxels * 4), o_istar = carve(B * 4), o_vnum = carve(B * 4);
    const size_t o_map = carve(B * p.plane * 4);
    h->bytes = off;
    cudaError_t e = cudaMalloc(&h->slab, h->bytes);
    if (e != cudaSuccess) { delete h; cudaGetLastError(); return fail(PF_ERR_OUT_OF_MEMORY, "pf_create: cudaMalloc of %zu bytes failed", off); }
    char *base = static_cast<char *>(h->slab);
    pf::Work &w = h->w;
    w.grid = reinterpret_cast<int32_t *>(base + o_grid);
    w.cellid = reinterpret_cast<int32_t *>(base + o_cell);
    w.vr = reinterpret_cast<int32_t *>(base + o_vr);
    w.tilecnt = reinterpret_cast<int32_t *>(base + o_tc);
    w.ticket = reinterpret_cast<int32_t *>(base + o_tk);
    w.cnt = reinterpret_cast<uint8_t *>(base + o_cnt);
    w.slotidx = reinterpret_cast<int32_t *>(base + o_slot);
    w.vcell = reinterpret_cast<int32_t *>(base + o_vcell);
    w.vcount = reinterpret_cast<int32_t *>(base + o_vcount);
    w.istar = reinterpret_cast<int32_t *>(base + o_istar);
    w.vnum = reinterpret_cast<int32_t *>(base + o_vnum);
    w.map = reinterpret_cast<int32_t *>(base + o_map);
    // one-time initialisation: the kernels put the grid / map back themselves after every call
    e = cudaMemsetAsync(h->slab, 0, h->bytes, 0);
    if (e == cudaSuccess) e = pf::launch_fill_i32(w.grid, pf::kEmpty, B * p.cells, 0);
    if (e == cudaSuccess) e = pf::launch_fill_i32(w.map, -1, B * p.plane, 0);
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);
    if (e != cudaSuccess) { cudaFree(h->slab); delete h; return cuda_fail(e, "pf_create: workspace initialisation"); }
    *out = h;
    return PF_OK;
}

int pf_destroy(pf_handle *h)
{
    if (!h) return PF_OK;
    DeviceGuard guard(h->device);
    if (h->timing)
        for (auto &ev : h->events) cudaEventDestroy(ev);
    cudaError_t e = cudaFree(h->slab);
    delete h;
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_destroy: cudaFree");
}

size_t pf_workspace_bytes(const pf_handle *h) { return h ? h->bytes : 0; }

int pf_debug_stage_timing(pf_handle *h, int enable)
{
    if (!h) return fail(PF_ERR_INVALID_ARGUMENT, "pf_debug_stage_timing: null handle");
    DeviceGuard guard(h->device);
    if (enable && !h->timing) {
        for (auto &ev : h->events) {
            cudaError_t e = cudaEventCreate(&ev);
            if (e != cudaSuccess) return cuda_fail(e, "pf_debug_stage_timing: cudaEventCreate");
        }
        h->timing = true;
    }
    h->w.ev = (enable && h->timing) ? h->events : nullptr;
    return PF_OK;
}

namespace {

// memcpy into a PINNED staging buffer with non-temporal stores.  A plain memcpy leaves the destination lines dirty in the
// caches of the cores that wrote them, and the H2D DMA that follows has to pull every line out of those caches: on the
// pool's hosts the frame graph (H2D of 1.9 MB + kernels) then takes 170-180 us on the device instead of 95
// (profiles/r2_frame_probe.txt).  Streaming stores send the lines straight to memory and skip the read-for-ownership.
// PF_STAGE_NT=0 selects the plain memcpy (A/B measurements).
bool stage_streaming()
{
    static const bool on = [] { const char *e = getenv("PF_STAGE_NT"); return !(e && e[0] == '0'); }();
    return on;
}

void copy_to_pinned(char *dst, const char *src, size_t bytes)
{
#if defined(__x86_64__)
    if (stage_streaming() && bytes >= 4096) {
        size_t head = (16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15;
        memcpy(dst, src, head);
        dst += head, src += head, bytes -= head;
        const size_t blocks = bytes / 64;
        for (size_t i = 0; i < blocks; ++i) {
            const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src));
            const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + 16));
            const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + 32));
            const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + 48));
            _mm_stream_si128(reinterpret_cast<__m128i *>(dst), a);
            _mm_stream_si128(reinterpret_cast<__m128i *>(dst + 16), b);
            _mm_stream_si128(reinterpret_cast<__m128i *>(dst + 32), c);
            _mm_stream_si128(reinterpret_cast<__m128i *>(dst + 48), d);
            src += 64, dst += 64;
        }
        memcpy(dst, src, bytes - blocks * 64);
        _mm_sfence();   // the stores must be globally visible before the copy engine is told to read them
        return;
    }
#endif
    memcpy(dst, src, bytes);
}

}  // namespace

int pf_stage_copy(void *dst, const void *src, size_t bytes, int threads)
{
    if (bytes == 0) return PF_OK;
    if (!dst || !src) return fail(PF_ERR_INVALID_ARGUMENT, "pf_stage_copy: null pointer");
    if (threads < 1) threads = 1;
    if (threads > 16) threads = 16;
    const size_t kMin = 256 * 1024;   // below this a second thread costs more than it saves
    if (threads == 1 || bytes < 2 * kMin) {
        copy_to_pinned(static_cast<char *>(dst), static_cast<const char *>(src), bytes);
        return PF_OK;
    }
    const size_t want = (bytes + kMin - 1) / kMin;
    if ((size_t)threads > want) threads = (int)want;
    const size_t chunk = ((bytes + threads - 1) / threads + 63) & ~(size_t)63;
#pragma omp parallel for num_threads(threads) schedule(static, 1)
    for (int t = 0; t < threads; ++t) {
        const size_t lo = (size_t)t * chunk;
        if (lo < bytes) copy_to_pinned(static_cast<char *>(dst) + lo, static_cast<const char *>(src) + lo, bytes - lo < chunk ? bytes - lo : chunk);
    }
    return PF_OK;
}

int pf_copy3(void *dst0, const void *src0, size_t bytes0, void *dst1, const void *src1, size_t bytes1, void *dst2,
             const void *src2, size_t bytes2, void *stream)
{
    void *d[3] = {dst0, dst1, dst2};
    const void *s[3] = {src0, src1, src2};
    const size_t b[3] = {bytes0, bytes1, bytes2};
    for (int k = 0; k < 3; ++k) {
        if (b[k] == 0) continue;
        if (!d[k] || !s[k]) return fail(PF_ERR_INVALID_ARGUMENT, "pf_copy3: null pointer in copy %d", k);
        cudaError_t e = cudaMemcpyAsync(d[k], s[k], b[k], cudaMemcpyDefault, static_cast<cudaStream_t>(stream));
        if (e != cudaSuccess) return cuda_fail(e, "pf_copy3");
    }
    return PF_OK;
}

// Host side of one VoxelGenerator.generate call in ONE library call (the interpreter costs more than the copies here):
// stage the frame, replay the frame graph, wait, read the voxel count.
int pf_frame_run(int device, void *graph_exec, void *stream, void *done_event, void *pending_event, void *pinned_points,
                 const void *points, size_t bytes, int threads, void *device_points, int32_t *pinned_offsets, int n_points,
                 const int32_t *pinned_voxel_num, int32_t *voxel_num)
{
    if (!graph_exec || !done_event || !pinned_offsets || !pinned_voxel_num || !voxel_num || n_points < 0)
        return fail(PF_ERR_INVALID_ARGUMENT, "pf_frame_run: bad argument");
    if (bytes > 0 && (!pinned_points || !points)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_frame_run: null points / staging buffer");
    pinned_offsets[0] = 0, pinned_offsets[1] = n_points;
    DeviceGuard guard(device);
    Range nvtx_range("pf_frame_run");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = cudaSuccess;
    if (bytes > 0 && device_points) {
        // the upload is not part of the graph: the frame is staged in up to four pieces and every piece goes to the copy
        // engine as soon as it is staged, so all but the first staging copy (and the graph launch) run under the DMA
        size_t chunk = (bytes + 3) / 4;
        if (chunk < (1u << 20)) chunk = 1u << 20;
        chunk = (chunk + 4095) & ~(size_t)4095;
        for (size_t off = 0; off < bytes && e == cudaSuccess; off += chunk) {
            const size_t len = bytes - off < chunk ? bytes - off : chunk;
            if (int rc = pf_stage_copy(static_cast<char *>(pinned_points) + off, static_cast<const char *>(points) + off, len, threads)) return rc;
            e = cudaMemcpyAsync(static_cast<char *>(device_points) + off, static_cast<const char *>(pinned_points) + off, len,
                                cudaMemcpyHostToDevice, st);
        }
        if (e != cudaSuccess) return cuda_fail(e, "pf_frame_run: upload");
    } else if (bytes > 0) {
        if (int rc = pf_stage_copy(pinned_points, points, bytes, threads)) return rc;
    }
    if (pending_event) e = cudaStreamWaitEvent(st, static_cast<cudaEvent_t>(pending_event), 0);
    if (e == cudaSuccess) e = cudaGraphLaunch(static_cast<cudaGraphExec_t>(graph_exec), st);
    if (e == cudaSuccess) e = cudaEventRecord(static_cast<cudaEvent_t>(done_event), st);
    if (e == cudaSuccess) e = cudaEventSynchronize(static_cast<cudaEvent_t>(done_event));
    if (e != cudaSuccess) return cuda_fail(e, "pf_frame_run");
    *voxel_num = pinned_voxel_num[0];
    return PF_OK;
}

int pf_frame_copy_out(int device, void *dst0, const void *src0, size_t bytes0, void *dst1, const void *src1, size_t bytes1,
                      void *dst2, const void *src2, size_t bytes2, void *stream, void *record_event, int synchronize)
{
    DeviceGuard guard(device);
    if (int rc = pf_copy3(dst0, src0, bytes0, dst1, src1, bytes1, dst2, src2, bytes2, stream)) return rc;
    cudaError_t e = cudaSuccess;
    if (record_event) e = cudaEventRecord(static_cast<cudaEvent_t>(record_event), static_cast<cudaStream_t>(stream));
    if (e == cudaSuccess && synchronize) e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_frame_copy_out");
}

int pf_upload_frames(void *device_points, void *pinned_stage, const void *const *frames, const int64_t *rows,
                     const int32_t *kinds, int n_frames, int n_slots, int row_bytes, int threads, int32_t *pinned_offsets,
                     void *stream)
{
    if (n_frames < 0 || n_slots < n_frames || row_bytes < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_upload_frames: bad sizes");
    if (!device_points || !pinned_offsets || (n_frames > 0 && (!frames || !rows || !kinds)))
        return fail(PF_ERR_INVALID_ARGUMENT, "pf_upload_frames: null argument");
    Range nvtx_range("pf_upload_frames");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int64_t pos = 0;
    for (int b = 0; b < n_slots; ++b) {
        if (pos > 0x7fffffff) return fail(PF_ERR_CAPACITY, "pf_upload_frames: more than 2^31 points");
        pinned_offsets[b] = (int32_t)pos;
        if (b >= n_frames || rows[b] <= 0) continue;
        const size_t bytes = (size_t)rows[b] * row_bytes;
        char *dst = static_cast<char *>(device_points) + (size_t)pos * row_bytes;
        const void *src = frames[b];
        int kind = kinds[b];
        if (kind < 0) {                         // ask the driver what kind of memory this is
            cudaPointerAttributes at{};
            if (cudaPointerGetAttributes(&at, src) != cudaSuccess) cudaGetLastError();
            kind = at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged ? 2 : (at.type == cudaMemoryTypeHost ? 1 : 0);
        }
        if (kind == 0) {                        // pageable host memory: through the pinned staging buffer
            if (!pinned_stage) return fail(PF_ERR_INVALID_ARGUMENT, "pf_upload_frames: pageable frame without a staging buffer");
            char *stage = static_cast<char *>(pinned_stage) + (size_t)pos * row_bytes;
            pf_stage_copy(stage, src, bytes, threads);
            src = stage;
        }
        cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, st);
        if (e != cudaSuccess) return cuda_fail(e, "pf_upload_frames");
        pos += rows[b];
    }
    if (pos > 0x7fffffff) return fail(PF_ERR_CAPACITY, "pf_upload_frames: more than 2^31 points");
    pinned_offsets[n_slots] = (int32_t)pos;
    return PF_OK;
}

int pf_debug_set_fill_variant(int variant)
{
    if (variant < -1 || variant > 8) return fail(PF_ERR_INVALID_ARGUMENT, "pf_debug_set_fill_variant: variant %d outside -1..8", variant);
    pf::set_fill_variant(variant);
    return pf::fill_variant();
}

int pf_debug_set_pdl(int enable)
{
    pf::set_pdl(enable);
    return pf::pdl_state();
}

int pf_debug_stage_times(pf_handle *h, float *ms, int capacity)
{
    if (!h || !ms) return fail(PF_ERR_INVALID_ARGUMENT, "pf_debug_stage_times: null argument");
    if (!h->w.ev) return fail(PF_ERR_INVALID_ARGUMENT, "pf_debug_stage_times: timing is not enabled");
    if (capacity < pf::kStageEvents - 1) return fail(PF_ERR_CAPACITY, "pf_debug_stage_times: need room for %d values", pf::kStageEvents - 1);
    DeviceGuard guard(h->device);
    for (int k = 1; k < pf::kStageEvents; ++k) {
        ms[k - 1] = -1.f;
        if (cudaEventSynchronize(h->events[k]) != cudaSuccess || cudaEventElapsedTime(&ms[k - 1], h->events[k - 1], h->events[k]) != cudaSuccess) {
            cudaGetLastError();   // a stage that did not run in the last call
            ms[k - 1] = -1.f;
        }
    }
    return pf::kStageEvents - 1;
}

static int voxelize_common(pf_handle *h, const pf::VoxelizeArgs &a, void *stream, const char *who)
{
    if (!h) return fail(PF_ERR_INVALID_ARGUMENT, "%s: null handle", who);
    if (a.n_frames < 0) return fail(PF_ERR_INVALID_ARGUMENT, "%s: n_frames < 0", who);
    if (a.n_frames == 0) return PF_OK;
    if (a.n_frames > h->max_frames) return fail(PF_ERR_CAPACITY, "%s: n_frames %d > handle capacity %d", who, a.n_frames, h->max_frames);
    if (a.max_frame_points < 0 || a.max_frame_points > h->max_points_per_frame)
        return fail(PF_ERR_CAPACITY, "%s: max_frame_points %d outside [0, %d]", who, a.max_frame_points, h->max_points_per_frame);
    if (!a.frame_offsets || (!a.points && a.max_frame_points > 0)) return fail(PF_ERR_INVALID_ARGUMENT, "%s: null points/frame_offsets", who);
    if (a.point_stride < h->p.F) return fail(PF_ERR_INVALID_ARGUMENT, "%s: point_stride %d < num_features %d", who, a.point_stride, h->p.F);
    if (a.coors_stride != 3 && a.coors_stride != 4) return fail(PF_ERR_INVALID_ARGUMENT, "%s: coors_stride must be 3 or 4", who);
    if (!a.voxel_num) return fail(PF_ERR_INVALID_ARGUMENT, "%s: voxel_num must not be null", who);
    if (a.voxels && !aligned16(a.voxels)) return fail(PF_ERR_INVALID_ARGUMENT, "%s: voxels must be 16-byte aligned", who);
    if (a.features && !aligned16(a.features)) return fail(PF_ERR_INVALID_ARGUMENT, "%s: features must be 16-byte aligned", who);
    if (a.keep_cell_map && h->p.nz != 1) return fail(PF_ERR_UNSUPPORTED, "%s: keep_cell_map needs nz == 1", who);
    if (h->map_pending > 0)
        return fail(PF_ERR_STATE, "%s: the cell map kept by the previous call (%d frames) was not consumed: call "
                                  "pf_scatter_from_map or pf_discard_cell_map first", who, h->map_pending);
    DeviceGuard guard(h->device);
    Range nvtx_range(who);
    cudaError_t e = pf::launch_voxelize(h->p, h->w, a, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, who);
    h->map_pending = a.keep_cell_map ? a.n_frames : 0;
    return PF_OK;
}

int pf_voxelize(pf_handle *h, const float *points, const int32_t *frame_offsets, int n_frames, int max_frame_points,
                int point_stride, int compact, int coors_stride, float *voxels, int32_t *coors, int32_t *num_points,
                int32_t *voxel_num, void *stream)
{
    pf::VoxelizeArgs a{};
    a.points = points, a.frame_offsets = frame_offsets, a.n_frames = n_frames, a.max_frame_points = max_frame_points;
    a.point_stride = point_stride, a.compact = compact, a.coors_stride = coors_stride;
    a.voxels = voxels, a.coors = coors, a.num_points = num_points, a.voxel_num = voxel_num;
    return voxelize_common(h, a, stream, "pf_voxelize");
}

int pf_voxelize_decorate(pf_handle *h, const float *points, const int32_t *frame_offsets, int n_frames,
                         int max_frame_points, int point_stride, int compact, int coors_stride, float vx, float vy,
                         float x_off, float y_off, int keep_cell_map, float *voxels, int32_t *coors,
                         int32_t *num_points, int32_t *voxel_num, float *features, void *stream)
{
    pf::VoxelizeArgs a{};
    a.points = points, a.frame_offsets = frame_offsets, a.n_frames = n_frames, a.max_frame_points = max_frame_points;
    a.point_stride = point_stride, a.compact = compact, a.coors_stride = coors_stride;
    a.vx = vx, a.vy = vy, a.x_off = x_off, a.y_off = y_off, a.keep_cell_map = keep_cell_map;
    a.voxels = voxels, a.coors = coors, a.num_points = num_points, a.voxel_num = voxel_num, a.features = features;
    return voxelize_common(h, a, stream, "pf_voxelize_decorate");
}

int pf_voxelize_pfn(pf_handle *h, const float *points, const int32_t *frame_offsets, int n_frames, int max_frame_points,
                    int point_stride, int compact, int coors_stride, float vx, float vy, float x_off, float y_off,
                    int keep_cell_map, const float *pfn_weight, const float *pfn_scale, const float *pfn_shift, int n_ch,
                    float *voxels, int32_t *coors, int32_t *num_points, int32_t *voxel_num, float *features,
                    float *pillar_feats, void *stream)
{
    if (n_ch != 64) return fail(PF_ERR_UNSUPPORTED, "pf_voxelize_pfn: n_ch must be 64 (got %d)", n_ch);
    if (!pfn_weight || !pfn_scale || !pfn_shift || !pillar_feats) return fail(PF_ERR_INVALID_ARGUMENT, "pf_voxelize_pfn: null PFN argument");
    if (h && (h->p.F != 4 || point_stride != 4)) return fail(PF_ERR_UNSUPPORTED, "pf_voxelize_pfn: needs 4 floats per point (F = 4, stride 4)");
    if (!aligned16(pillar_feats) || !aligned16(points)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_voxelize_pfn: points / pillar_feats must be 16-byte aligned");
    pf::VoxelizeArgs a{};
    a.points = points, a.frame_offsets = frame_offsets, a.n_frames = n_frames, a.max_frame_points = max_frame_points;
    a.point_stride = point_stride, a.compact = compact, a.coors_stride = coors_stride;
    a.vx = vx, a.vy = vy, a.x_off = x_off, a.y_off = y_off, a.keep_cell_map = keep_cell_map;
    a.voxels = voxels, a.coors = coors, a.num_points = num_points, a.voxel_num = voxel_num, a.features = features;
    a.pfn_w = pfn_weight, a.pfn_scale = pfn_scale, a.pfn_shift = pfn_shift, a.pfn_out = pillar_feats;
    return voxelize_common(h, a, stream, "pf_voxelize_pfn");
}

int pf_pfn(const float *voxels, const int32_t *num_points, const int32_t *coors, int coors_stride, int64_t n_voxels,
           int max_points, int num_features, float vx, float vy, float x_off, float y_off, const float *pfn_weight,
           const float *pfn_scale, const float *pfn_shift, int n_ch, float *pillar_feats, void *stream)
{
    if (n_voxels < 0) return fail(PF_ERR_INVALID_ARGUMENT, "pf_pfn: n_voxels < 0");
    if (n_voxels == 0) return PF_OK;
    if (!voxels || !num_points || !coors || !pfn_weight || !pfn_scale || !pfn_shift || !pillar_feats)
        return fail(PF_ERR_INVALID_ARGUMENT, "pf_pfn: null argument");
    if (n_ch != 64 || num_features != 4) return fail(PF_ERR_UNSUPPORTED, "pf_pfn: needs n_ch == 64 and num_features == 4");
    if (coors_stride < 2 || max_points < 1 || max_points > 512) return fail(PF_ERR_UNSUPPORTED, "pf_pfn: coors_stride >= 2, 1 <= max_points <= 512");
    if (!aligned16(voxels) || !aligned16(pillar_feats)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_pfn: voxels / pillar_feats must be 16-byte aligned");
    Range nvtx_range("pf_pfn");
    cudaError_t e = pf::launch_decorate(voxels, num_points, coors, coors_stride, n_voxels, max_points, num_features, vx, vy,
                                        x_off, y_off, nullptr, static_cast<cudaStream_t>(stream), pfn_weight, pfn_scale,
                                        pfn_shift, pillar_feats);
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_pfn");
}

int pf_decorate(const float *voxels, const int32_t *num_points, const int32_t *coors, int coors_stride,
                int64_t n_voxels, int max_points, int num_features, float vx, float vy, float x_off, float y_off,
                float *features, void *stream)
{
    if (n_voxels < 0) return fail(PF_ERR_INVALID_ARGUMENT, "pf_decorate: n_voxels < 0");
    if (n_voxels == 0) return PF_OK;
    if (!voxels || !num_points || !coors || !features) return fail(PF_ERR_INVALID_ARGUMENT, "pf_decorate: null argument");
    if (coors_stride < 2) return fail(PF_ERR_INVALID_ARGUMENT, "pf_decorate: coors_stride must be >= 2");
    if (max_points < 1 || max_points > 4096) return fail(PF_ERR_UNSUPPORTED, "pf_decorate: max_points %d outside 1..4096", max_points);
    if (num_features < 3 || num_features > 16) return fail(PF_ERR_UNSUPPORTED, "pf_decorate: num_features %d outside 3..16", num_features);
    if (!aligned16(features)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_decorate: features must be 16-byte aligned");
    if (int rc = check_gather_shape(max_points, num_features, -1, "pf_decorate")) return rc;
    Range nvtx_range("pf_decorate");
    cudaError_t e = pf::launch_decorate(voxels, num_points, coors, coors_stride, n_voxels, max_points, num_features, vx, vy,
                                        x_off, y_off, features, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_decorate");
}

int pf_scatter(pf_handle *h, const float *feats, int64_t feat_stride_v, int64_t feat_stride_c, const int32_t *coors,
               int coors_stride, int64_t n_voxels, int batch, int n_ch, float *canvas, void *stream)
{
    if (!h || !canvas) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter: null handle/canvas");
    if (n_voxels < 0 || n_voxels >= (1ll << 31)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter: n_voxels out of range");
    if (n_voxels > 0 && (!feats || !coors)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter: null feats/coors");
    if (coors_stride < 2) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter: coors_stride must be >= 2");
    if (batch < 1 || batch > h->max_frames) return fail(PF_ERR_CAPACITY, "pf_scatter: batch %d outside [1, %d]", batch, h->max_frames);
    if (n_ch < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter: n_ch < 1");
    DeviceGuard guard(h->device);
    Range nvtx_range("pf_scatter");
    cudaError_t e = pf::launch_scatter(h->p, h->w, feats, feat_stride_v, feat_stride_c, coors, coors_stride, n_voxels,
                                       batch, n_ch, canvas, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_scatter");
}

// Allocation-free variant of pf_scatter: no handle, the cell map is the file-scope device array of pf_scatter.cu.
// Calls no cudaMalloc / cudaFree and no allocating library routine; runs on the CURRENT device.
int pf_scatter_static(const float *feats, int64_t feat_stride_v, int64_t feat_stride_c, const int32_t *coors,
                      int coors_stride, int64_t n_voxels, int batch, int n_ch, int nx, int ny, float *canvas, void *stream)
{
    if (!canvas) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_static: null canvas");
    if (n_voxels < 0 || n_voxels >= (1ll << 31)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_static: n_voxels out of range");
    if (n_voxels > 0 && (!feats || !coors)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_static: null feats/coors");
    if (coors_stride < 2) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_static: coors_stride must be >= 2");
    if (batch < 1 || n_ch < 1 || nx < 1 || ny < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_static: batch, n_ch, nx, ny must be >= 1");
    const long long max_cells = pf::static_scatter_max_cells();   // == PF_SCATTER_STATIC_MAX_CELLS of the header
    if ((long long)nx * ny > max_cells || (long long)nx * ny * batch > max_cells)
        return fail(PF_ERR_CAPACITY, "pf_scatter_static: batch %d x %d x %d cells exceed the static cell map of %lld cells "
                    "(use pf_scatter with a handle for larger canvases)", batch, nx, ny, max_cells);
    Range nvtx_range("pf_scatter_static");
    cudaError_t e = pf::launch_scatter_static(feats, feat_stride_v, feat_stride_c, coors, coors_stride, n_voxels, batch, n_ch,
                                              nx, ny, canvas, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_scatter_static");
}

int pf_scatter_from_map(pf_handle *h, const float *feats, int64_t feat_stride_v, int64_t feat_stride_c, int n_frames,
                        int compact, int n_ch, float *canvas, void *stream)
{
    if (!h || !canvas || !feats) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_from_map: null argument");
    if (h->p.nz != 1) return fail(PF_ERR_UNSUPPORTED, "pf_scatter_from_map: needs nz == 1");
    if (n_frames < 1 || n_frames > h->max_frames) return fail(PF_ERR_CAPACITY, "pf_scatter_from_map: n_frames %d outside [1, %d]", n_frames, h->max_frames);
    if (n_ch < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_from_map: n_ch < 1");
    if (h->map_pending == 0) return fail(PF_ERR_STATE, "pf_scatter_from_map: no cell map is pending (voxelise with keep_cell_map first)");
    if (n_frames != h->map_pending)
        return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_from_map: n_frames %d, but the kept map covers %d frames", n_frames, h->map_pending);
    DeviceGuard guard(h->device);
    Range nvtx_range("pf_scatter_from_map");
    cudaError_t e = pf::launch_scatter_from_map(h->p, h->w, feats, feat_stride_v, feat_stride_c, n_frames, compact, n_ch,
                                                canvas, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "pf_scatter_from_map");
    h->map_pending = 0;   // the fill puts every cell it reads back to empty
    return PF_OK;
}

int pf_discard_cell_map(pf_handle *h, void *stream)
{
    if (!h) return fail(PF_ERR_INVALID_ARGUMENT, "pf_discard_cell_map: null handle");
    if (h->map_pending == 0) return PF_OK;
    DeviceGuard guard(h->device);
    cudaError_t e = pf::launch_fill_i32(h->w.grid, pf::kEmpty, (size_t)h->map_pending * h->p.cells, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "pf_discard_cell_map");
    h->map_pending = 0;
    return PF_OK;
}

int pf_scatter_backward(const float *grad_canvas, const int32_t *coors, int coors_stride, int64_t n_voxels, int batch,
                        int n_ch, int nx, int ny, float *grad_feats, int64_t feat_stride_v, int64_t feat_stride_c,
                        void *stream)
{
    if (n_voxels < 0) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_backward: n_voxels < 0");
    if (n_voxels == 0) return PF_OK;
    if (!grad_canvas || !coors || !grad_feats) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_backward: null argument");
    if (coors_stride < 2 || batch < 1 || n_ch < 1 || nx < 1 || ny < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_scatter_backward: bad sizes");
    Range nvtx_range("pf_scatter_backward");
    cudaError_t e = pf::launch_scatter_backward(grad_canvas, coors, coors_stride, n_voxels, batch, n_ch, nx, ny, grad_feats,
                                                feat_stride_v, feat_stride_c, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_scatter_backward");
}

int pf_anchor_mask(const int32_t *coors, int coors_stride, const int32_t *voxel_num, int64_t n_rows, int rows_per_frame,
                   int n_frames, int nx, int ny, const int32_t *anchor_coors, int64_t n_anchors, int32_t *integral,
                   uint8_t *mask, void *stream)
{
    if (n_frames < 0 || n_rows < 0 || n_anchors < 0) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: negative size");
    if (n_frames == 0) return PF_OK;
    if (nx < 1 || ny < 1 || (long long)nx * ny >= (1ll << 31)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: bad grid %d x %d", nx, ny);
    if (!integral) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: null integral scratch");
    if (n_rows > 0 && !coors) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: null coors");
    if (n_anchors > 0 && (!anchor_coors || !mask)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: null anchor_coors/mask");
    if (!aligned16(anchor_coors)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: anchor_coors must be 16-byte aligned");
    if (coors_stride < 2) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: coors_stride must be >= 2");
    if (rows_per_frame < 0) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: rows_per_frame < 0");
    if (rows_per_frame > 0 && !voxel_num) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: the slab layout needs voxel_num");
    const bool frame_in_column = rows_per_frame == 0 && n_frames > 1;
    if (frame_in_column && coors_stride < 4) return fail(PF_ERR_INVALID_ARGUMENT, "pf_anchor_mask: frame id is column 3, coors_stride must be >= 4");
    Range nvtx_range("pf_anchor_mask");
    cudaError_t e = pf::launch_anchor_mask(coors, coors_stride, voxel_num, n_rows, rows_per_frame, n_frames, nx, ny,
                                           anchor_coors, n_anchors, integral, mask, nullptr, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_anchor_mask");
}

size_t pf_nms_scratch_bytes(int n_segments, int max_segment_boxes)
{
    if (n_segments < 1 || max_segment_boxes < 1 || max_segment_boxes > 4096) return 0;
    return pf::nms_scratch_bytes(n_segments, max_segment_boxes);
}

int pf_nms(const float *boxes, const float *scores, const int32_t *seg_offsets, int n_segments, int64_t n_total,
           int max_segment_boxes, float iou_threshold, int post_max, void *scratch, int32_t *keep, int32_t *keep_num,
           void *stream)
{
    if (n_segments < 0 || n_total < 0) return fail(PF_ERR_INVALID_ARGUMENT, "pf_nms: negative size");
    if (n_segments == 0) return PF_OK;
    if (max_segment_boxes < 1 || max_segment_boxes > 4096) return fail(PF_ERR_UNSUPPORTED, "pf_nms: max_segment_boxes %d outside 1..4096", max_segment_boxes);
    if (post_max < 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_nms: post_max < 1");
    if (!seg_offsets && n_segments != 1) return fail(PF_ERR_INVALID_ARGUMENT, "pf_nms: seg_offsets may be null only for one segment");
    if (!seg_offsets && n_total > max_segment_boxes) return fail(PF_ERR_CAPACITY, "pf_nms: %lld boxes > max_segment_boxes %d", (long long)n_total, max_segment_boxes);
    if (!scratch || !keep || !keep_num) return fail(PF_ERR_INVALID_ARGUMENT, "pf_nms: null scratch/keep/keep_num");
    if (n_total > 0 && (!boxes || !scores)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_nms: null boxes/scores");
    if (!aligned16(boxes) || (reinterpret_cast<uintptr_t>(scratch) & 255)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_nms: boxes must be 16-byte and scratch 256-byte aligned");
    Range nvtx_range("pf_nms");
    cudaError_t e = pf::launch_nms(boxes, scores, seg_offsets, n_segments, n_total, max_segment_boxes, iou_threshold, post_max,
                                   scratch, keep, keep_num, static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_nms");
}

}  // extern "C"

namespace {

// validates a pf_post_desc and the batch shape; 0 = fine
int post_check(const pf_post_desc *d, int n_frames, int64_t n_anchors, const char *who)
{
    if (!d) return fail(PF_ERR_INVALID_ARGUMENT, "%s: null desc", who);
    if (d->n_classes < 1 || d->n_classes > pf::post::kMaxClasses) return fail(PF_ERR_UNSUPPORTED, "%s: n_classes %d outside 1..%d", who, d->n_classes, pf::post::kMaxClasses);
    if (d->n_cls_channels < 1) return fail(PF_ERR_INVALID_ARGUMENT, "%s: n_cls_channels < 1", who);
    if (d->pre_max < 1 || d->pre_max > 4096) return fail(PF_ERR_UNSUPPORTED, "%s: pre_max %d outside 1..4096", who, d->pre_max);
    if (d->post_max < 1 || d->post_max > pf::post::kFinishThreads) return fail(PF_ERR_UNSUPPORTED, "%s: post_max %d outside 1..%d", who, d->post_max, pf::post::kFinishThreads);
    if (n_frames < 1 || n_frames > 65535) return fail(PF_ERR_INVALID_ARGUMENT, "%s: n_frames %d outside 1..65535", who, n_frames);
    if (n_anchors < 0 || n_anchors >= (1ll << 31)) return fail(PF_ERR_INVALID_ARGUMENT, "%s: n_anchors out of range", who);
    if ((long long)n_frames * d->n_classes > 65535) return fail(PF_ERR_CAPACITY, "%s: %lld segments > 65535", who, (long long)n_frames * d->n_classes);
    for (int c = 0; c < d->n_classes; ++c) {
        const long long lo = d->class_range[c][0], hi = d->class_range[c][1];
        if (lo < 0 || hi < lo || hi > n_anchors) return fail(PF_ERR_INVALID_ARGUMENT, "%s: class_range[%d] = [%lld, %lld) outside [0, %lld]", who, c, lo, hi, (long long)n_anchors);
        for (int e = 0; e < c; ++e)
            if (lo < d->class_range[e][1] && d->class_range[e][0] < hi) return fail(PF_ERR_INVALID_ARGUMENT, "%s: class ranges %d and %d overlap", who, e, c);
    }
    return 0;
}

}  // namespace

extern "C" {

size_t pf_post_scratch_layout(const pf_post_desc *d, int n_frames, int64_t n_anchors, size_t *offsets)
{
    if (post_check(d, n_frames, n_anchors, "pf_post_scratch_layout") != 0) return 0;
    return pf::post_scratch_layout(d->n_classes, d->pre_max, d->post_max, n_frames, n_anchors, offsets);
}

size_t pf_post_scratch_bytes(const pf_post_desc *d, int n_frames, int64_t n_anchors)
{
    return pf_post_scratch_layout(d, n_frames, n_anchors, nullptr);
}

int pf_postprocess(const pf_post_desc *d, const float *cls_preds, const float *box_preds, const float *dir_preds,
                   const uint8_t *anchors_mask, const float *anchors, int n_frames, int64_t n_anchors, void *scratch,
                   float *det_boxes, float *det_scores, int32_t *det_class, int32_t *det_anchor, int32_t *det_num,
                   void *stream)
{
    if (n_frames == 0) return PF_OK;
    const int rc = post_check(d, n_frames, n_anchors, "pf_postprocess");
    if (rc != 0) return rc;
    if (!scratch || (reinterpret_cast<uintptr_t>(scratch) & 255)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_postprocess: scratch must be non-null and 256-byte aligned");
    if (!det_boxes || !det_scores || !det_class || !det_num) return fail(PF_ERR_INVALID_ARGUMENT, "pf_postprocess: null output");
    if (n_anchors > 0 && (!cls_preds || !box_preds || !dir_preds || !anchors)) return fail(PF_ERR_INVALID_ARGUMENT, "pf_postprocess: null input");
    pf::post::Desc k{};
    k.n_classes = d->n_classes;
    for (int c = 0; c < d->n_classes; ++c) k.lo[c] = d->class_range[c][0], k.hi[c] = d->class_range[c][1];
    k.n_ch = d->n_cls_channels;
    k.pre_max = d->pre_max;
    k.post_max = d->post_max;
    k.score_thr = d->score_threshold;
    pf::post::set_prefilter(k);
    for (int i = 0; i < 6; ++i) k.limit[i] = d->center_limit[i];
    Range nvtx_range("pf_postprocess");
    cudaError_t e = pf::launch_postprocess(k, d->iou_threshold, cls_preds, box_preds, dir_preds, anchors_mask, anchors, n_frames,
                                           n_anchors, scratch, det_boxes, det_scores, det_class, det_anchor, det_num,
                                           static_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? PF_OK : cuda_fail(e, "pf_postprocess");
}

}  // extern "C"
