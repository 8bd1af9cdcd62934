/*
 * CPU oracle for the ImVoxelNet image->voxel path.  TEST INFRASTRUCTURE ONLY:
 * it is the checker for the CUDA kernels and the timed CPU baseline of
 * bench.py, never a product code path (see oracle/__init__.py).
 *
 * Plain C restatement of the arithmetic the reference performs with torch
 * ops in /root/reference/mmdet3d/models/detectors/imvoxelnet.py:
 *   get_points    :132-141   ivx_oracle_points
 *   backproject   :145-160   ivx_oracle_project / ivx_oracle_backproject
 *   view reduction :70-74    ivx_oracle_view_mean / ivx_oracle_backproject_mean
 *
 * Parity: PINNED.  tests/test_oracle_golden.py checks every function here
 * against outputs of the reference functions themselves (the .npz files under tests/golden,
 * produced by tests/golden/make_golden.py from the unmodified reference).
 *
 * Floating point contract (what makes pixel indices bit-exact):
 *   - voxel coordinate = rn(rn(i * voxel_size) + new_origin), two roundings
 *     (:139-140 computes it with separate fp32 mul and add);
 *   - each row of P.[x y z 1]^T is the K=4 dot product evaluated as
 *     fma(P3,1, fma(P2,z, fma(P1,y, rn(P0*x)))), which is what the CPU
 *     torch.bmm of :150 produces (checked against the goldens);
 *   - IEEE fp32 division, then round-half-to-even (torch.round, :151-152);
 *   - compile with -ffp-contract=off so the compiler fuses nothing itself.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

int ivx_oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ivx_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* get_points, reference :132-141.  points is [3][X*Y*Z], Z fastest. */
void ivx_oracle_new_origin(const int32_t n_voxels[3], const float voxel_size[3],
                           const float origin[3], float new_origin[3]) {
    for (int a = 0; a < 3; ++a) {
        volatile float half = (float)n_voxels[a] / 2.0f;  /* n_voxels / 2. */
        volatile float span = half * voxel_size[a];       /* ... * voxel_size */
        new_origin[a] = origin[a] - span;                 /* origin - ... */
    }
}

void ivx_oracle_points(const int32_t n_voxels[3], const float voxel_size[3],
                       const float origin[3], float *points) {
    float no[3];
    ivx_oracle_new_origin(n_voxels, voxel_size, origin, no);
    const int64_t X = n_voxels[0], Y = n_voxels[1], Z = n_voxels[2];
    const int64_t N = X * Y * Z;
    for (int64_t ix = 0; ix < X; ++ix)
        for (int64_t iy = 0; iy < Y; ++iy)
            for (int64_t iz = 0; iz < Z; ++iz) {
                const int64_t n = (ix * Y + iy) * Z + iz;
                volatile float px = (float)ix * voxel_size[0];
                volatile float py = (float)iy * voxel_size[1];
                volatile float pz = (float)iz * voxel_size[2];
                points[0 * N + n] = px + no[0];
                points[1 * N + n] = py + no[1];
                points[2 * N + n] = pz + no[2];
            }
}

static inline float dot4_row(const float *row, float x, float y, float z) {
    float acc = row[0] * x;
    acc = fmaf(row[1], y, acc);
    acc = fmaf(row[2], z, acc);
    acc = fmaf(row[3], 1.0f, acc);
    return acc;
}

/* float -> int64 the way x86 torch does (.long()): out of range / NaN -> INT64_MIN */
static inline int64_t to_long(float v) {
    if (!(v >= -9223372036854775808.0f && v < 9223372036854775808.0f)) return INT64_MIN;
    return (int64_t)v;
}

static inline int project_one(const float *P, float px, float py, float pz,
                              int h, int w, int64_t *xo, int64_t *yo) {
    const float p0 = dot4_row(P + 0, px, py, pz);
    const float p1 = dot4_row(P + 4, px, py, pz);
    const float p2 = dot4_row(P + 8, px, py, pz);
    const int64_t xi = to_long(rintf(p0 / p2));
    const int64_t yi = to_long(rintf(p1 / p2));
    *xo = xi;
    *yo = yi;
    return (xi >= 0) & (yi >= 0) & (xi < (int64_t)w) & (yi < (int64_t)h) & (p2 > 0.0f);
}

/* reference :148-154.  x, y: [V][N] int64, valid: [V][N] uint8. */
void ivx_oracle_project(const float *points, int64_t N, const float *proj, int V,
                        int h, int w, int64_t *x, int64_t *y, uint8_t *valid) {
#pragma omp parallel for collapse(2) schedule(static)
    for (int v = 0; v < V; ++v)
        for (int64_t n = 0; n < N; ++n) {
            int64_t xi, yi;
            const int ok = project_one(proj + 12 * v, points[n], points[N + n],
                                       points[2 * N + n], h, w, &xi, &yi);
            x[(int64_t)v * N + n] = xi;
            y[(int64_t)v * N + n] = yi;
            valid[(int64_t)v * N + n] = (uint8_t)ok;
        }
}

/*
 * reference :145-160.  feat is addressed with element strides (sv, sc, sh, sw)
 * so the strided [:, :, :h, :w] slice of :69 and channels-last maps both work.
 * volume: [V][C][N] fp32 (zeros where invalid), valid: [V][N] uint8.
 */
void ivx_oracle_backproject(const float *feat, int64_t sv, int64_t sc, int64_t sh,
                            int64_t sw, int V, int C, int h, int w,
                            const float *points, int64_t N, const float *proj,
                            float *volume, uint8_t *valid) {
#pragma omp parallel for schedule(static)
    for (int v = 0; v < V; ++v) {
        float *vol = volume + (int64_t)v * C * N;
        memset(vol, 0, sizeof(float) * (size_t)C * (size_t)N);
        for (int64_t n = 0; n < N; ++n) {
            int64_t xi, yi;
            const int ok = project_one(proj + 12 * v, points[n], points[N + n],
                                       points[2 * N + n], h, w, &xi, &yi);
            valid[(int64_t)v * N + n] = (uint8_t)ok;
            if (!ok) continue;
            const float *src = feat + v * sv + yi * sh + xi * sw;
            for (int c = 0; c < C; ++c) vol[(int64_t)c * N + n] = src[c * sc];
        }
    }
}

/* reference :70-74 on a materialised [V][C][N] volume. */
void ivx_oracle_view_mean(const float *volume, const uint8_t *valid, int V, int C,
                          int64_t N, float *out, uint8_t *valid_out, int32_t *count) {
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; ++n) {
        int32_t cnt = 0;
        for (int v = 0; v < V; ++v) cnt += valid[(int64_t)v * N + n];
        if (count) count[n] = cnt;
        valid_out[n] = cnt > 0;
        for (int c = 0; c < C; ++c) {
            float s = 0.0f;
            for (int v = 0; v < V; ++v) s += volume[((int64_t)v * C + c) * N + n];
            out[(int64_t)c * N + n] = cnt > 0 ? s / (float)cnt : 0.0f;
        }
    }
}

/*
 * backproject + view mean without the [V][C][N] temporary (same values as the
 * two calls above; used for the large configurations and as the CPU
 * baseline).  out: [C][N] fp32, valid_out: [N] uint8, count: [N] int32 or NULL.
 */
void ivx_oracle_backproject_mean(const float *feat, int64_t sv, int64_t sc,
                                 int64_t sh, int64_t sw, int V, int C, int h, int w,
                                 const float *points, int64_t N, const float *proj,
                                 float *out, uint8_t *valid_out, int32_t *count) {
#pragma omp parallel
    {
        float *acc = (float *)malloc(sizeof(float) * (size_t)C);
#pragma omp for schedule(static)
        for (int64_t n = 0; n < N; ++n) {
            int32_t cnt = 0;
            for (int c = 0; c < C; ++c) acc[c] = 0.0f;
            for (int v = 0; v < V; ++v) {
                int64_t xi, yi;
                if (!project_one(proj + 12 * v, points[n], points[N + n],
                                 points[2 * N + n], h, w, &xi, &yi))
                    continue;
                ++cnt;
                const float *src = feat + v * sv + yi * sh + xi * sw;
                for (int c = 0; c < C; ++c) acc[c] += src[c * sc];
            }
            if (count) count[n] = cnt;
            valid_out[n] = cnt > 0;
            for (int c = 0; c < C; ++c)
                out[(int64_t)c * N + n] = cnt > 0 ? acc[c] / (float)cnt : 0.0f;
        }
        free(acc);
    }
}
