// Projection set-up on the device: ImVoxelNet._compute_projection
// (detectors/imvoxelnet.py:114-129) for a whole batch in one launch instead of a Python loop of
// small CPU tensor ops per scene.  The result has to carry the exact bits of the reference's CPU
// computation (they decide which pixel a voxel rounds to), so the arithmetic is spelled out:
//   K'[i][:] = K[i][:] / float(ratio)   for rows 0, 1         (":119", IEEE fp32 division)
//   P[i][j]  = fma(K'[i][2], E[2][j], fma(K'[i][1], E[1][j], rn(K'[i][0] * E[0][j])))
// which is what `intrinsic @ extrinsic[:3]` (:128) evaluates to on the CPU (checked bit for bit
// against torch in tests/test_setup_gpu.py and against tests/golden/host_setup.npz).
#include "ivx_common.cuh"

namespace ivx {

__global__ void __launch_bounds__(256) compute_projection_kernel(const float *__restrict__ intrinsic,
                                                                 const float *__restrict__ ratio,
                                                                 const float *__restrict__ extrinsic,
                                                                 int n_views, long long total,
                                                                 float *__restrict__ proj) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const long long bv = idx / 12;
    const int e = (int)(idx % 12), i = e >> 2, j = e & 3;
    const long long b = bv / n_views;
    const float *K = intrinsic + b * 9 + i * 3;
    const float *E = extrinsic + bv * 16 + j;
    float k0 = K[0], k1 = K[1], k2 = K[2];
    if (i < 2) {
        const float r = ratio[b];
        k0 = __fdiv_rn(k0, r), k1 = __fdiv_rn(k1, r), k2 = __fdiv_rn(k2, r);
    }
    float acc = __fmul_rn(k0, E[0]);
    acc = __fmaf_rn(k1, E[4], acc);
    acc = __fmaf_rn(k2, E[8], acc);
    proj[idx] = acc;
}

}  // namespace ivx

using namespace ivx;

extern "C" int ivx_compute_projection(const float *intrinsic, const float *ratio, const float *extrinsic,
                                      int32_t n_scenes, int32_t n_views, float *proj, void *stream) {
    IVX_REQUIRE(intrinsic && ratio && extrinsic && proj, "NULL argument");
    IVX_REQUIRE(n_scenes > 0 && n_views > 0 && n_views <= IVX_MAX_VIEWS, "bad scene / view count");
    const long long total = (long long)n_scenes * n_views * 12;
    compute_projection_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        intrinsic, ratio, extrinsic, n_views, total, proj);
    IVX_CHECK_LAUNCH("projection set-up launch");
    return IVX_OK;
}
