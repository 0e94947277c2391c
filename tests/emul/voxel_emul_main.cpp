// TEST INFRASTRUCTURE ONLY.  Stand-alone driver of the emulated voxelisation kernels for the HOST sanitizers
// (compute-sanitizer is closed on this GPU pool): the very kernel bodies of csrc/pf_voxel_kernels.cuh run on host threads
// under AddressSanitizer + UBSan (out-of-bounds / misaligned accesses into the workspace arrays -- the memcheck role) and
// under ThreadSanitizer (accesses to shared memory or to the workspace that are not ordered by a barrier, an atomic or
// a flag -- the racecheck role), and every result is compared with the sequential C oracle.
//   g++ -fsanitize=address,undefined ... voxel_emul_main.cpp voxel_emul.cpp ../../oracle/pillar_oracle.c
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

extern "C" {
void *emul_vox_create(const float *offset, const float *voxel_size, const int *grid, int max_voxels, int max_points,
                      int num_features, int max_frames, int max_points_per_frame);
void emul_vox_destroy(void *h);
int emul_voxelize(void *h, const float *points, const int *frame_offsets, int n_frames, int max_frame_points,
                  int point_stride, int keep_cell_map, float *voxels, int *coors, int *num_points, int *voxel_num);
int32_t oracle_generate(const float *points, int64_t n_points, int32_t n_feat, float *voxels, int32_t *num_per_voxel,
                        int32_t *coors, const float *voxel_size, const float *offset, const int32_t *grid_size,
                        int32_t max_voxels, int32_t max_points);
}

static uint32_t rng_state = 12345u;
static float urand()
{
    rng_state = rng_state * 1664525u + 1013904223u;
    return (float)(rng_state >> 8) * (1.0f / 16777216.0f);
}

int main(int argc, char **argv)
{
    const int n_cases = argc > 1 ? atoi(argv[1]) : 3;
    int bad = 0;
    for (int c = 0; c < n_cases; ++c) {
        const int grid[3] = {20 + 7 * c, 16, c == 2 ? 2 : 1};
        const float vs[3] = {0.5f, 0.25f, 1.0f}, off[3] = {-3.0f, 1.0f, -2.0f};
        const int mv = c == 1 ? 40 : 400, P = c == 0 ? 3 : 200, F = 4, B = 2;
        const int n0 = 5000 + 1500 * c, n1 = c == 1 ? 0 : 2300;
        std::vector<float> pts((size_t)(n0 + n1 + 1) * F);
        for (int i = 0; i < n0 + n1; ++i) {
            // a quarter of the points pile into a few cells (saturating counters), the rest spread past the range
            const bool hot = (i % 4) == 0;
            pts[(size_t)i * F + 0] = off[0] + (hot ? 1.3f + 0.4f * (i % 3) : (urand() * 1.3f - 0.15f) * grid[0] * vs[0]);
            pts[(size_t)i * F + 1] = off[1] + (hot ? 0.6f : (urand() * 1.3f - 0.15f) * grid[1] * vs[1]);
            pts[(size_t)i * F + 2] = off[2] + urand() * grid[2] * vs[2];
            pts[(size_t)i * F + 3] = urand();
        }
        const int offs[3] = {0, n0, n0 + n1};
        void *h = emul_vox_create(off, vs, grid, mv, P, F, B, n0);
        std::vector<float> vox((size_t)B * mv * P * F), wv((size_t)mv * P * F);
        std::vector<int> coors((size_t)B * mv * 3), num((size_t)B * mv), vnum(B), wc((size_t)mv * 3), wn(mv);
        for (int rep = 0; rep < 2; ++rep) {
            const int rc = emul_voxelize(h, pts.data(), offs, B, n0, F, rep, vox.data(), coors.data(), num.data(), vnum.data());
            if (rc != 0 && !(rep == 1 && rc == 0)) { printf("case %d rep %d: workspace invariant %d\n", c, rep, rc); ++bad; }
            for (int f = 0; f < B; ++f) {
                const int n = offs[f + 1] - offs[f];
                const int V = oracle_generate(pts.data() + (size_t)offs[f] * F, n, F, wv.data(), wn.data(), wc.data(), vs, off, grid, mv, P);
                const size_t r0 = (size_t)f * mv;
                if (V != vnum[f] || memcmp(wn.data(), num.data() + r0, sizeof(int) * V) || memcmp(wc.data(), coors.data() + 3 * r0, sizeof(int) * 3 * V) ||
                    memcmp(wv.data(), vox.data() + r0 * P * F, sizeof(float) * (size_t)V * P * F)) {
                    printf("case %d rep %d frame %d: MISMATCH against the oracle (V %d vs %d)\n", c, rep, f, vnum[f], V);
                    ++bad;
                }
            }
            if (rep == 0) continue;
        }
        emul_vox_destroy(h);
        printf("case %d: grid %dx%dx%d max_voxels %d P %d frames %d+%d ok\n", c, grid[0], grid[1], grid[2], mv, P, n0, n1);
    }
    printf(bad ? "FAILED: %d problems\n" : "ALL CASES MATCH THE ORACLE (%d problems)\n", bad);
    return bad != 0;
}
