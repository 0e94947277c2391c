// Per-voxel gather + pillar decoration, F = 4 fast path (reference: framework/voxel_generator.py:102-105 for
// the stored rows, networks/pointpillars8_shared.py:32-54 for the decoration).
//
// Warp-granular, no block barrier: GROUP lanes own one voxel and each lane owns the slots gl, gl+GROUP, ...
//   * voxel rows are 16 bytes, so a lane stores its float4 straight to global memory: the lanes of a group
//     write consecutive rows -> whole 32-byte sectors, padding zeros included;
//   * the 36-byte feature rows are staged per warp in shared memory (stride 9 floats) and streamed out with
//     16-byte stores; the staging buffer is shifted by the output's 16-byte phase so the vector body is aligned
//     on both sides whatever row the warp starts at (slab and compact layouts).
// HBM-write bound by design (16*P*V + 36*P*V bytes), but a pillar is only ~800 bytes, so the instruction count
// per pillar decides whether the bound is reached: the shipped P values (10, 15, 100) are compile-time
// instances (unrolled slot loops, folded index arithmetic), small P uses 4 lanes per voxel (8 voxels per warp,
// 2 shuffle rounds), the three mean divisions of a voxel run in three different lanes, and the cell ->
// (x, y, z) split uses a multiply-high instead of integer division.  (v2 of this kernel was a block-wide,
// runtime-P version: 38 M warp instructions per 128k pillars, issue bound -- profiles/r1_step_v2_ncu.md.)
#include "pf_internal.cuh"

namespace pf {

constexpr int kG4Warps = 8;

__host__ __device__ inline int g4_warp_floats(int vpw, int P) { return (vpw * P * 9 + 4 + 3) & ~3; }

// PT > 0: compile-time max_points; PT == 0: runtime p.P
template <int GROUP, int PT, bool FROM_POINTS, bool PFN>
__global__ void __launch_bounds__(kG4Warps * 32) k_gather4(Params p, Work w, GatherArgs a)
{
    constexpr int VPW = 32 / GROUP;   // voxels per warp
    extern __shared__ __align__(16) float s_g4[];
    const int P = PT ? PT : p.P;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane / GROUP, gl = lane % GROUP;
    const int f = FROM_POINTS ? blockIdx.y : 0;
    pdl_trigger();
    pdl_wait();   // last kernel of the voxelise chain: nothing to do before k_slot's lists exist (no-op when launched plainly)
    const int n_vox = FROM_POINTS ? w.vnum[f] : (int)a.n_voxels;
    const int v0 = (blockIdx.x * kG4Warps + warp) * VPW;   // first voxel of this warp
    if (v0 >= n_vox) return;
    const int nvw = (n_vox - v0) < VPW ? (n_vox - v0) : VPW;
    float *s_w = s_g4 + warp * g4_warp_floats(VPW, P);

    long long row0 = v0;   // first output row of this warp
    int pt_base = 0;
    if (FROM_POINTS) {
        pt_base = a.frame_offsets[f];
        if (a.compact) {
            int s = 0;
            for (int k = lane; k < f; k += 32) s += w.vnum[k];
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            row0 += s;
        } else {
            row0 += (long long)f * p.max_voxels;
        }
    }
    const bool active = g < nvw;
    const int v = v0 + g;   // voxel within the frame / input

    int cnt = 0, cx = 0, cy = 0, cz = 0;
    if (active) {
        if (FROM_POINTS) {
            cnt = w.vcount[(size_t)f * p.max_voxels + v];
            const int cell = w.vcell[(size_t)f * p.max_voxels + v];
            const int xy = div_small(cell, p.nz, p.nz_m);
            cz = cell - xy * p.nz;
            cx = div_small(xy, p.ny, p.ny_m);
            cy = xy - cx * p.ny;
        } else {
            cnt = a.num_in[v];
            cx = a.coors_in[(size_t)v * a.coors_stride + 0];
            cy = a.coors_in[(size_t)v * a.coors_stride + 1];
        }
    }
    const int phase = (int)((row0 * (P * 9)) & 3);
    float *frow = s_w + phase + g * (P * 9);                                       // this voxel's staged rows
    float4 *vout = FROM_POINTS && a.voxels ? reinterpret_cast<float4 *>(a.voxels) + (row0 + g) * P : nullptr;
    const float4 *src4 = FROM_POINTS ? reinterpret_cast<const float4 *>(a.points) + pt_base
                                     : reinterpret_cast<const float4 *>(a.voxels_in) + (size_t)v * P;
    const int32_t *slots = FROM_POINTS ? w.slotidx + ((size_t)f * p.max_voxels + v) * P : nullptr;
    const bool want_f = PFN || a.features != nullptr;   // stage the decorated rows

    // pass 1: fetch the rows, store the voxel rows, stage x,y,z,i, accumulate the xyz sums
    float sx = 0.f, sy = 0.f, sz = 0.f;
    if (active) {
        constexpr int ITERS = PT ? (PT + GROUP - 1) / GROUP : 1;
        if (PT) {
            float4 q[ITERS];
            int idx[ITERS];
#pragma unroll
            for (int k = 0; k < ITERS; ++k) {
                const int r = gl + k * GROUP;
                idx[k] = FROM_POINTS ? ((r < cnt) ? slots[r] : -1) : ((r < P) ? r : -1);
            }
#pragma unroll
            for (int k = 0; k < ITERS; ++k)
                q[k] = idx[k] >= 0 ? __ldg(src4 + idx[k]) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < ITERS; ++k) {
                const int r = gl + k * GROUP;
                if (r < P) {
                    if (FROM_POINTS && vout) vout[r] = q[k];
                    sx += q[k].x, sy += q[k].y, sz += q[k].z;
                    if (want_f) {
                        float *d = frow + r * 9;
                        d[0] = q[k].x, d[1] = q[k].y, d[2] = q[k].z, d[3] = q[k].w;
                    }
                }
            }
        } else {
            for (int r = gl; r < P; r += GROUP) {
                float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                if (FROM_POINTS) {
                    if (r < cnt) q = __ldg(src4 + slots[r]);
                    if (vout) vout[r] = q;
                } else {
                    q = __ldg(src4 + r);   // the reference sums every slot, padding included (:32)
                }
                sx += q.x, sy += q.y, sz += q.z;
                if (want_f) {
                    float *d = frow + r * 9;
                    d[0] = q.x, d[1] = q.y, d[2] = q.z, d[3] = q.w;
                }
            }
        }
    }
#pragma unroll
    for (int o = GROUP / 2; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o);
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
    if (!want_f) return;

    // the three mean divisions of a voxel run in lanes 0,1,2 of its group and are then broadcast
    const float mine = gl == 0 ? sx : (gl == 1 ? sy : sz);
    const float mean = __fdiv_rn(mine, (float)cnt);
    const int gbase = g * GROUP;
    const float mx = __shfl_sync(0xffffffffu, mean, gbase + 0);
    const float my = __shfl_sync(0xffffffffu, mean, gbase + 1);
    const float mz = __shfl_sync(0xffffffffu, mean, gbase + 2);

    // pass 2: decorate in shared memory (each lane revisits the slots it staged itself)
    if (active) {
        // pillar centre: float(c) * v rounded, then + offset rounded -- two fp32 ops, no FMA
        const float ccx = __fadd_rn(__fmul_rn((float)cx, a.vx), a.x_off);
        const float ccy = __fadd_rn(__fmul_rn((float)cy, a.vy), a.y_off);
        constexpr int ITERS = PT ? (PT + GROUP - 1) / GROUP : 1;
        const int iters = PT ? ITERS : (P + GROUP - 1) / GROUP;
#pragma unroll
        for (int k = 0; k < iters; ++k) {
            const int r = gl + k * GROUP;
            if (r < P) {
                float *d = frow + r * 9;
                if (r < cnt) {
                    const float x = d[0], y = d[1], z = d[2];
                    d[4] = __fsub_rn(x, mx);
                    d[5] = __fsub_rn(y, my);
                    d[6] = __fsub_rn(z, mz);
                    d[7] = __fsub_rn(x, ccx);
                    d[8] = __fsub_rn(y, ccy);
                } else {
#pragma unroll
                    for (int c = 0; c < 9; ++c) d[c] = 0.f;   // features *= mask (:47-54)
                }
            }
        }
    }
    __syncwarp();

    if (PFN) {
        // Fused pillar feature net (networks/pointpillars8_shared.py:24-28,57-60 in eval mode): Conv1d(9 -> 64, k = 1,
        // no bias) + folded BatchNorm1d + ReLU + max over the P slots.  A lane owns two channels and keeps their 18
        // weights in registers; the staged feature rows are read as shared-memory broadcasts.  Padded slots are zero
        // rows, i.e. the candidate relu(shift): taken once when the pillar is not full instead of P - cnt times.
        const int c0 = lane * 2;
        float w0[9], w1[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) w0[k] = __ldg(a.pfn_w + c0 * 9 + k), w1[k] = __ldg(a.pfn_w + (c0 + 1) * 9 + k);
        const float sc0 = __ldg(a.pfn_scale + c0), sc1 = __ldg(a.pfn_scale + c0 + 1);
        const float sh0 = __ldg(a.pfn_shift + c0), sh1 = __ldg(a.pfn_shift + c0 + 1);
        for (int q = 0; q < nvw; ++q) {
            const int cnt_q = __shfl_sync(0xffffffffu, cnt, q * GROUP);
            const float *fr = s_w + phase + q * (P * 9);
            float m0 = cnt_q < P ? sh0 : -3.402823466e38f, m1 = cnt_q < P ? sh1 : -3.402823466e38f;
#pragma unroll 2
            for (int r = 0; r < cnt_q; ++r) {
                float d0 = 0.f, d1 = 0.f;
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const float fv = fr[r * 9 + k];
                    d0 = fmaf(fv, w0[k], d0);
                    d1 = fmaf(fv, w1[k], d1);
                }
                m0 = fmaxf(m0, fmaf(d0, sc0, sh0));
                m1 = fmaxf(m1, fmaf(d1, sc1, sh1));
            }
            *reinterpret_cast<float2 *>(a.pfn_out + (row0 + q) * 64 + c0) = make_float2(fmaxf(m0, 0.f), fmaxf(m1, 0.f));
        }
        if (!a.features) return;
    }

    // copy-out: the warp's nvw*P*9 floats are contiguous in the output
    const int total = nvw * (P * 9);
    float *gdst = a.features + row0 * (P * 9);
    const float *ssrc = s_w + phase;
    const int head = min((4 - phase) & 3, total);
    if (lane < head) gdst[lane] = ssrc[lane];
    const int nvec = (total - head) >> 2;
    float4 *gd = reinterpret_cast<float4 *>(gdst + head);
    const float4 *sd = reinterpret_cast<const float4 *>(ssrc + head);
#pragma unroll 4
    for (int j = lane; j < nvec; j += 32) gd[j] = sd[j];
    const int done = head + (nvec << 2);
    if (lane < total - done) gdst[done + lane] = ssrc[done + lane];
}

template <int GROUP, int PT, bool FROM_POINTS, bool PFN>
static cudaError_t run_g4x(const Params &p, const Work &w, const GatherArgs &a, long long voxels, int n_frames,
                           cudaStream_t stream)
{
    const bool dependent = FROM_POINTS && pdl_stage(5) && !w.ev;   // behind k_slot on the same stream
    constexpr int VPW = 32 / GROUP;
    const size_t smem = (size_t)kG4Warps * g4_warp_floats(VPW, p.P) * sizeof(float);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_gather4<GROUP, PT, FROM_POINTS, PFN>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    const long long per_cta = (long long)kG4Warps * VPW;
    const dim3 grid((unsigned)((voxels + per_cta - 1) / per_cta), (unsigned)n_frames);
    return launch_k(k_gather4<GROUP, PT, FROM_POINTS, PFN>, grid, dim3(kG4Warps * 32), smem, stream, dependent, p, w, a);
}

template <int GROUP, int PT, bool FROM_POINTS>
static cudaError_t run_g4(const Params &p, const Work &w, const GatherArgs &a, long long voxels, int n_frames,
                          cudaStream_t stream)
{
    return a.pfn_w ? run_g4x<GROUP, PT, FROM_POINTS, true>(p, w, a, voxels, n_frames, stream)
                   : run_g4x<GROUP, PT, FROM_POINTS, false>(p, w, a, voxels, n_frames, stream);
}

template <bool FROM_POINTS>
static cudaError_t pick_g4(const Params &p, const Work &w, const GatherArgs &a, long long voxels, int n_frames,
                           cudaStream_t stream)
{
    switch (p.P) {   // max_num_points of the shipped configs: 15 (20 cm, 16 cm), 10 (10 cm), 100 (nuScenes)
    case 15: return run_g4<4, 15, FROM_POINTS>(p, w, a, voxels, n_frames, stream);
    case 10: return run_g4<4, 10, FROM_POINTS>(p, w, a, voxels, n_frames, stream);
    case 100: return run_g4<32, 100, FROM_POINTS>(p, w, a, voxels, n_frames, stream);
    default: break;
    }
    if (p.P <= 16) return run_g4<16, 0, FROM_POINTS>(p, w, a, voxels, n_frames, stream);
    return run_g4<32, 0, FROM_POINTS>(p, w, a, voxels, n_frames, stream);
}

cudaError_t launch_gather4(const Params &p, const Work &w, const GatherArgs &a, int n_frames, bool from_points,
                           cudaStream_t stream)
{
    auto al16 = [](const void *q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (p.F != 4 || p.P > 512) return cudaErrorNotSupported;
    if (from_points) {
        if (a.point_stride != 4 || !al16(a.points) || !al16(a.voxels)) return cudaErrorNotSupported;
    } else {
        if (!al16(a.voxels_in) || a.n_voxels >= (1ll << 31)) return cudaErrorNotSupported;
    }
    if (a.features && !al16(a.features)) return cudaErrorNotSupported;
    if (a.pfn_w && (!a.pfn_scale || !a.pfn_shift || !a.pfn_out || (reinterpret_cast<uintptr_t>(a.pfn_out) & 7))) return cudaErrorNotSupported;
    if (from_points) return pick_g4<true>(p, w, a, p.max_voxels, n_frames, stream);
    return pick_g4<false>(p, w, a, a.n_voxels, 1, stream);
}

}  // namespace pf
