// Host-side launchers of the voxelisation chain (kernels: pf_voxel_kernels.cuh, pf_gather.cu) and the generic-F
// gather / decoration kernel.  Reference: framework/voxel_generator.py:28-40, 82-106.
#include <cstdlib>

#include "pf_voxel_kernels.cuh"

namespace pf {

// ---------------------------------------------------------------------------------------------------------
__global__ void k_fill_i32(int32_t *__restrict__ dst, int32_t value, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) dst[i] = value;
}

cudaError_t launch_fill_i32(int32_t *dst, int32_t value, size_t n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    size_t blocks = (n + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_fill_i32<<<(unsigned)blocks, threads, 0, stream>>>(dst, value, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------
// per-voxel gather, generic F: + pillar decoration (networks/pointpillars8_shared.py:32-54 of the reference).
// GROUP lanes (16 when P <= 16, else 32) own one voxel; a 256-thread CTA stages 256/GROUP voxels in shared
// memory and then streams the voxel rows and the decorated rows out with 16-byte stores, so every output
// byte -- padding zeros included -- is written exactly once and fully coalesced.
//   FROM_POINTS = true : rows come from the point cloud through slotidx (fused voxelise path)
//   FROM_POINTS = false: rows come from an existing voxels tensor (stand-alone pf_decorate)

__device__ __forceinline__ void copy_out(float *__restrict__ gdst, const float *ssrc, int count, int phase)
{
    // gdst and ssrc share the same 16-byte phase (ssrc = aligned base + phase floats)
    int head = (4 - phase) & 3;
    if (head > count) head = count;
    if ((int)threadIdx.x < head) gdst[threadIdx.x] = ssrc[threadIdx.x];
    const int nvec = (count - head) >> 2;
    float4 *gd = reinterpret_cast<float4 *>(gdst + head);
    const float4 *sd = reinterpret_cast<const float4 *>(ssrc + head);
    for (int j = threadIdx.x; j < nvec; j += blockDim.x) gd[j] = sd[j];
    const int done = head + (nvec << 2);
    if ((int)threadIdx.x < count - done) gdst[done + threadIdx.x] = ssrc[done + threadIdx.x];
}

template <int GROUP, bool FROM_POINTS, bool VEC4>
__global__ void __launch_bounds__(256) k_gather(Params p, Work w, GatherArgs a)
{
    constexpr int VPB = 256 / GROUP;
    extern __shared__ __align__(16) float s_mem[];
    const int P = p.P, F = p.F, FO = p.F + 5;
    float *s_vox = s_mem;                                 // [4 + VPB*P*F]
    float *s_feat = s_mem + ((VPB * P * F + 4 + 3) & ~3); // [4 + VPB*P*FO]

    const int f = FROM_POINTS ? blockIdx.y : 0;
    const long long n_vox = FROM_POINTS ? (long long)w.vnum[f] : a.n_voxels;
    const long long v0 = (long long)blockIdx.x * VPB;
    if (v0 >= n_vox) return;
    const int nv = (int)((n_vox - v0) < VPB ? (n_vox - v0) : VPB);

    long long row0 = v0;   // first output row of this CTA
    int pt_base = 0;
    if (FROM_POINTS) {
        pt_base = a.frame_offsets[f];
        if (a.compact) {
            long long s = 0;
            for (int k = 0; k < f; ++k) s += w.vnum[k];
            row0 += s;
        } else {
            row0 += (long long)f * p.max_voxels;
        }
    }
    const int phase_v = (int)((row0 * P * F) & 3);
    const int phase_f = (int)((row0 * P * FO) & 3);

    const int tid = threadIdx.x;
    const int g = tid / GROUP, gl = tid % GROUP;
    const bool active = g < nv;
    const long long v = v0 + g;

    int cnt = 0, cx = 0, cy = 0, cz = 0;
    if (active) {
        if (FROM_POINTS) {
            cnt = w.vcount[(size_t)f * p.max_voxels + v];
            const int cell = w.vcell[(size_t)f * p.max_voxels + v];
            cz = cell % p.nz;
            const int xy = cell / p.nz;
            cy = xy % p.ny;
            cx = xy / p.ny;
        } else {
            cnt = a.num_in[v];
            cx = a.coors_in[v * a.coors_stride + 0];
            cy = a.coors_in[v * a.coors_stride + 1];
        }
    }
    // stage the voxel rows, accumulate the xyz sums
    float sx = 0.f, sy = 0.f, sz = 0.f;
    float *vrow = s_vox + phase_v + (size_t)g * P * F;
    if (active) {
        for (int r = gl; r < P; r += GROUP) {
            float *dst = vrow + r * F;
            if (FROM_POINTS) {
                if (r < cnt) {
                    const int idx = w.slotidx[((size_t)f * p.max_voxels + v) * P + r];
                    if (VEC4) {
                        const float4 q = __ldg(reinterpret_cast<const float4 *>(a.points) + (size_t)(pt_base + idx));
                        dst[0] = q.x, dst[1] = q.y, dst[2] = q.z, dst[3] = q.w;
                        sx += q.x, sy += q.y, sz += q.z;
                    } else {
                        const float *q = a.points + (size_t)(pt_base + idx) * a.point_stride;
                        for (int k = 0; k < F; ++k) dst[k] = __ldg(q + k);
                        sx += dst[0], sy += dst[1], sz += dst[2];
                    }
                } else {
                    for (int k = 0; k < F; ++k) dst[k] = 0.f;
                }
            } else {
                const float *q = a.voxels_in + ((size_t)v * P + r) * F;
                for (int k = 0; k < F; ++k) dst[k] = __ldg(q + k);
                sx += dst[0], sy += dst[1], sz += dst[2];   // the reference sums every slot (:32)
            }
        }
    }
#pragma unroll
    for (int o = GROUP / 2; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o);
    }
    if (active && a.features) {
        const float fc = (float)cnt;
        const float mx = __fdiv_rn(sx, fc), my = __fdiv_rn(sy, fc), mz = __fdiv_rn(sz, fc);
        // pillar centre: float(c) * v rounded, then + offset rounded -- two fp32 ops, no FMA
        const float ccx = __fadd_rn(__fmul_rn((float)cx, a.vx), a.x_off);
        const float ccy = __fadd_rn(__fmul_rn((float)cy, a.vy), a.y_off);
        float *frow = s_feat + phase_f + (size_t)g * P * FO;
        for (int r = gl; r < P; r += GROUP) {
            const float *src = vrow + r * F;
            float *dst = frow + r * FO;
            if (r < cnt) {
                for (int k = 0; k < F; ++k) dst[k] = src[k];
                dst[F + 0] = __fsub_rn(src[0], mx);
                dst[F + 1] = __fsub_rn(src[1], my);
                dst[F + 2] = __fsub_rn(src[2], mz);
                dst[F + 3] = __fsub_rn(src[0], ccx);
                dst[F + 4] = __fsub_rn(src[1], ccy);
            } else {
                for (int k = 0; k < FO; ++k) dst[k] = 0.f;
            }
        }
    }
    if (FROM_POINTS && active && gl == 0) {
        const long long row = row0 + g;
        if (a.coors) {
            int32_t *c = a.coors + row * a.coors_stride;
            c[0] = cx, c[1] = cy, c[2] = cz;
            if (a.coors_stride >= 4) c[3] = f;
        }
        if (a.num_points) a.num_points[row] = cnt;
    }
    __syncthreads();
    if (FROM_POINTS && a.voxels) copy_out(a.voxels + (size_t)row0 * P * F, s_vox + phase_v, nv * P * F, phase_v);
    if (a.features) copy_out(a.features + (size_t)row0 * P * FO, s_feat + phase_f, nv * P * FO, phase_f);
}

// ---------------------------------------------------------------------------------------------------------
// host-side launchers
// Dynamic shared memory of the gather / decoration kernel a shape would be launched with (the F = 4 fast path of
// pf_gather.cu for P <= 512, else the generic kernel above): what the C ABI checks against the device's opt-in limit at
// validation time, so that an oversized (P, F) is refused as PF_ERR_UNSUPPORTED instead of failing at launch.
size_t gather_smem_needed(int P, int F)
{
    if (F == 4 && P <= 512) {
        const int vpw = (P == 15 || P == 10) ? 8 : (P == 100 ? 1 : (P <= 16 ? 2 : 1));
        return (size_t)8 * (((size_t)vpw * P * 9 + 4 + 3) & ~(size_t)3) * sizeof(float);
    }
    const int vpb = 256 / (P <= 16 ? 16 : 32);
    const size_t a = ((size_t)vpb * P * F + 4 + 3) & ~(size_t)3;
    const size_t b = (size_t)vpb * P * (F + 5) + 4;
    return (a + b) * sizeof(float);
}

static size_t gather_smem_bytes(int group, int P, int F)
{
    const int vpb = 256 / group;
    const size_t a = ((size_t)vpb * P * F + 4 + 3) & ~(size_t)3;
    const size_t b = (size_t)vpb * P * (F + 5) + 4;
    return (a + b) * sizeof(float);
}

template <int GROUP, bool FROM_POINTS, bool VEC4>
static cudaError_t run_gather(const Params &p, const Work &w, const GatherArgs &a, dim3 grid, cudaStream_t stream)
{
    const size_t smem = gather_smem_bytes(GROUP, p.P, p.F);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_gather<GROUP, FROM_POINTS, VEC4>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_gather<GROUP, FROM_POINTS, VEC4><<<grid, 256, smem, stream>>>(p, w, a);
    return cudaGetLastError();
}

template <bool FROM_POINTS>
static cudaError_t dispatch_gather(const Params &p, const Work &w, const GatherArgs &a, dim3 grid_g16, dim3 grid_g32,
                                   bool vec4, cudaStream_t stream)
{
    if (p.P <= 16) {
        return vec4 ? run_gather<16, FROM_POINTS, true>(p, w, a, grid_g16, stream)
                    : run_gather<16, FROM_POINTS, false>(p, w, a, grid_g16, stream);
    }
    return vec4 ? run_gather<32, FROM_POINTS, true>(p, w, a, grid_g32, stream)
                : run_gather<32, FROM_POINTS, false>(p, w, a, grid_g32, stream);
}

// Programmatic dependent launch of the chain kernels.  Modes: 0 = off (the default), 1 = on for the launches PF_PDL_MASK
// selects (PF_PDL=1), 2 = "auto" (PF_PDL=2): only the second quantize range of a SINGLE-frame call is a dependent.
// Measured (profiles/r2_chain_experiments.md E5): the second quantize range and the canvas fill as dependents save 4-5 us of a
// serial step, the five kernels between them lose 7-10 us (the early-resident dependents take issue slots and shared memory
// from their predecessor's tail), and with two or three front ends in flight every form of it is slower (0.290 -> 0.329 ms);
// a lone frame (the drop-in path: one front end) gains 2-5 us from the quantize boundary and nothing more -- inside generate()
// that is within the noise (call AN: 143.1 against 143.9 us), so even "auto" is not the default.
static int g_pdl = -1;
static int g_pdl_mask = 0x7f;   // experiments: which launches are dependents (bit 0 quantize range 2, 1 assign, 2 tile_rank,
                                // 3 column_scan, 4 slot, 5 gather, 6 canvas fill); PF_PDL_MASK
static int pdl_mode()
{
    if (g_pdl < 0) {
        const char *e = getenv("PF_PDL");
        g_pdl = !e ? 0 : (e[0] == '1' ? 1 : (e[0] == '2' ? 2 : 0));
        const char *m = getenv("PF_PDL_MASK");
        if (m) g_pdl_mask = (int)strtol(m, nullptr, 0);
    }
    return g_pdl;
}
bool pdl_enabled() { return pdl_mode() == 1; }
bool pdl_stage(int bit) { return pdl_enabled() && ((g_pdl_mask >> bit) & 1); }
void set_pdl(int on) { g_pdl = on < 0 ? 2 : (on ? 1 : 0); }
int pdl_state() { return pdl_mode(); }

cudaError_t launch_voxelize(const Params &p, const Work &w, const VoxelizeArgs &a, cudaStream_t stream)
{
    const int B = a.n_frames;
    if (B <= 0) return cudaSuccess;
    int npts = a.max_frame_points;
    if (npts < 1) npts = 1;
    const int tiles = (npts + kTile - 1) / kTile;
    const bool vec4 = (a.point_stride == 4) && (p.F == 4) && ((reinterpret_cast<uintptr_t>(a.points) & 15) == 0);
    constexpr int kPtsPerCta = 256 * kQPer;
    const dim3 g_pts((npts + kPtsPerCta - 1) / kPtsPerCta, B), g_tiles(tiles, B);
    // leading 1/16 of every frame first (see k_quantize); small frames go in one launch
    static const int split_den = [] { const char *e = getenv("PF_QUANTIZE_SPLIT"); const int d = e ? atoi(e) : 16; return d < 0 ? 16 : d; }();   // experiments: 0 = one launch
    const int split = (split_den > 1 && npts >= 32 * kPtsPerCta) ? ((npts / split_den + kPtsPerCta - 1) / kPtsPerCta) * kPtsPerCta : npts;
    // Every kernel after the first is launched as a programmatic dependent of the one in front of it (pf_internal.cuh):
    // its launch, index arithmetic and shared-memory clears run under the predecessor's tail.  Off while the stage events
    // are recorded (an event between two kernels is a full dependency anyway).
    const bool dep = pdl_enabled() && !w.ev;
    const bool dep_q = !w.ev && (pdl_mode() == 2 ? B == 1 : pdl_stage(0));   // second quantize range behind the first
    cudaError_t le = cudaSuccess;
    mark(w, kEvStart, stream);
    for (int pass = 0; pass < 2; ++pass) {
        const int lo = pass == 0 ? 0 : split, hi = pass == 0 ? split : npts;
        if (lo >= hi) continue;
        const dim3 g((hi - lo + kPtsPerCta - 1) / kPtsPerCta, B);
        if (vec4) le = launch_k(k_quantize<true>, g, dim3(256), 0, stream, dep_q && pass == 1, p, w, a.points, a.frame_offsets, a.point_stride, lo, hi);
        else      le = launch_k(k_quantize<false>, g, dim3(256), 0, stream, dep_q && pass == 1, p, w, a.points, a.frame_offsets, a.point_stride, lo, hi);
        if (le != cudaSuccess) return le;
    }
    mark(w, kEvQuantize, stream);
    le = launch_k(k_assign, dim3((npts + kATile - 1) / kATile, B), dim3(256), 0, stream, dep && pdl_stage(1), p, w, a.frame_offsets, a.voxel_num, npts);
    if (le != cudaSuccess) return le;
    mark(w, kEvAssign, stream);
    if ((size_t)p.vcols > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_tile_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, p.vcols);
        if (e != cudaSuccess) return e;
    }
    le = launch_k(k_tile_rank, g_tiles, dim3(kTileWarps * 32), (size_t)p.vcols, stream, dep && pdl_stage(2), p, w, a.frame_offsets, npts);
    if (le != cudaSuccess) return le;
    mark(w, kEvTileRank, stream);
    le = launch_k(k_column_scan, dim3((p.vcols / 4 + 127) / 128, B), dim3(128), 0, stream, dep && pdl_stage(3), p, w, a.frame_offsets, npts);
    if (le != cudaSuccess) return le;
    mark(w, kEvColumnScan, stream);
    le = launch_k(k_slot, g_pts, dim3(256), 0, stream, dep && pdl_stage(4), p, w, a.frame_offsets, npts, a.keep_cell_map ? 0 : 1);
    if (le != cudaSuccess) return le;
    mark(w, kEvSlot, stream);

    GatherArgs ga{};
    ga.points = a.points, ga.frame_offsets = a.frame_offsets, ga.point_stride = a.point_stride;
    ga.compact = a.compact, ga.coors_stride = a.coors_stride;
    ga.vx = a.vx, ga.vy = a.vy, ga.x_off = a.x_off, ga.y_off = a.y_off;
    ga.voxels = a.voxels, ga.coors = a.coors, ga.num_points = a.num_points, ga.features = a.features;
    ga.pfn_w = a.pfn_w, ga.pfn_scale = a.pfn_scale, ga.pfn_shift = a.pfn_shift, ga.pfn_out = a.pfn_out;
    cudaError_t e = vec4 ? launch_gather4(p, w, ga, B, true, stream) : cudaErrorNotSupported;
    if (e == cudaErrorNotSupported && a.pfn_w) return e;   // the fused PFN exists only on the F = 4 fast path
    if (e == cudaErrorNotSupported)   // generic F / stride / alignment
        e = dispatch_gather<true>(p, w, ga, dim3((p.max_voxels + 15) / 16, B), dim3((p.max_voxels + 7) / 8, B), vec4,
                                  stream);
    if (e != cudaSuccess) return e;
    mark(w, kEvGather, stream);
    return cudaGetLastError();
}

cudaError_t launch_decorate(const float *voxels, const int32_t *num, const int32_t *coors, int coors_stride,
                            int64_t n_voxels, int P, int F, float vx, float vy, float x_off, float y_off,
                            float *features, cudaStream_t stream, const float *pfn_w, const float *pfn_scale,
                            const float *pfn_shift, float *pfn_out)
{
    if (n_voxels <= 0) return cudaSuccess;
    Params p{};
    p.P = P, p.F = F, p.nx = p.ny = p.nz = 1;
    Work w{};
    GatherArgs ga{};
    ga.voxels_in = voxels, ga.num_in = num, ga.coors_in = coors, ga.coors_stride = coors_stride;
    ga.n_voxels = n_voxels;
    ga.vx = vx, ga.vy = vy, ga.x_off = x_off, ga.y_off = y_off;
    ga.features = features;
    ga.pfn_w = pfn_w, ga.pfn_scale = pfn_scale, ga.pfn_shift = pfn_shift, ga.pfn_out = pfn_out;
    const cudaError_t e4 = launch_gather4(p, w, ga, 1, false, stream);
    if (e4 != cudaErrorNotSupported || pfn_w) return e4;   // the fused PFN exists only on the F = 4 fast path
    const unsigned b16 = (unsigned)((n_voxels + 15) / 16), b32 = (unsigned)((n_voxels + 7) / 8);
    return dispatch_gather<false>(p, w, ga, dim3(b16, 1), dim3(b32, 1), false, stream);
}

}  // namespace pf
