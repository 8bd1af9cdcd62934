// Fused image->voxel back-projection for B200 (sm_100a).
//
// One launch replaces, for a whole batch of scenes, what the reference does
// with ~40 ATen launches and a Python loop per scene
// (mmdet3d/models/detectors/imvoxelnet.py:57-78 and 145-160): project every
// voxel through every view, round to a pixel, test validity, gather the
// C-channel feature of that pixel and reduce over the views.
//
// Work decomposition (both kernels):
//   * a warp owns a tile of T consecutive voxels of one scene;
//   * phase 1 -- all 32 lanes project (voxel, view) pairs of the tile.  The
//     3x4 matrices sit in shared memory, voxel coordinates are rebuilt in
//     registers from the grid description.  Hits are compacted on the fly
//     (ballot + popc) into a per-voxel list of pixel offsets in shared memory,
//     in ascending view order; single-view scenes keep the one offset in a
//     register instead;
//   * phase 2 -- G lanes own one voxel: they walk its hit list four pixels at
//     a time and stream the C contiguous channels of each pixel with 128-bit
//     loads (all four in flight), accumulating the sum in registers; the
//     epilogue divides by the hit count, zeroes unseen voxels and stores with
//     streaming 128-bit stores.  Single-view scenes instead keep four voxels
//     per group in flight, since a voxel has at most one pixel to fetch.
//
// Bit-exactness of the pixel index (checked against the reference's outputs):
// the K=4 dot product is evaluated as rn(P0*x) then three fma's -- what the
// reference's torch.bmm (:150) does on CPU -- followed by IEEE division and
// round-half-even.  Every float op on that path is an explicit _rn intrinsic
// so that nvcc can neither contract nor reassociate it.
#include <cuda_bf16.h>

#include <mutex>

#include "ivx_common.cuh"

namespace ivx {

// floor(n / d) for 0 <= n < 2^31 as one multiply-high and a shift
// (Granlund-Montgomery with p = 31 + ceil(log2 d)); d == 1 is flagged by m == 0.
struct FastDiv {
    unsigned m;
    int s;
    __device__ __forceinline__ int div(int n) const {
        return m ? (int)(__umulhi((unsigned)n, m) >> s) : n;
    }
};

static FastDiv make_fastdiv(int d) {
    FastDiv f = {0u, 0};
    if (d <= 1) return f;
    int lg = 0;
    while ((1ll << lg) < d) ++lg;
    const int p = 31 + lg;
    f.m = (unsigned)(((1ull << p) + (unsigned long long)d - 1) / (unsigned long long)d);
    f.s = p - 32;
    return f;
}

struct BpArgs {
    const float *feat;                       // (bf16 maps: same pointer, see feat_bf16)
    int feat_bf16;                           // 1: the maps are bf16 (vectorised kernel only)
    long long fs_b, fs_v, fs_c, fs_h, fs_w;  // element strides of [B][V][C][H][W]
    int q_v, q_h, q_w;                       // the same in float4 units (vectorised kernel)
    const float *proj;                       // [B][V][12]
    long long proj_stride_b;                 // elements between scenes' matrix blocks
    const float *points;                     // [3][N] or nullptr
    float vs[3];
    int V, C, X, Y, Z, N;
    FastDiv div_z, div_y;
    void *vol;
    long long os_b, os_c, os_n;
    uint8_t *valid;  // [B][N]
    int32_t *count;  // [B][N]
    int reduce;      // ivx_reduce
    int tile_log2;   // T = 1 << tile_log2, 2..5
    int table_stride;
    int hw[IVX_MAX_SCENES_PER_LAUNCH][2];
    float new_origin[IVX_MAX_SCENES_PER_LAUNCH][3];
};

__device__ __forceinline__ float dot4_rn(const float *P, float x, float y, float z) {
    float acc = __fmul_rn(P[0], x);
    acc = __fmaf_rn(P[1], y, acc);
    acc = __fmaf_rn(P[2], z, acc);
    return __fadd_rn(acc, P[3]);  // fma(P3, 1, acc) == rn(acc + P3)
}

// Pixel hit by one voxel in one view.  Returns validity; xr/yr are the rounded
// pixel coordinates as floats (exact integers when valid).
__device__ __forceinline__ bool project_rn(const float *P, float px, float py, float pz,
                                           int h, int w, float &xr, float &yr) {
    // depth first: a voxel behind the camera (or at depth 0: inf / NaN lanes) is invalid whatever
    // its pixel is (:154), so the two dot products and divisions are skipped for it
    const float p2 = dot4_rn(P + 8, px, py, pz);
    xr = yr = 0.f;
    if (!(p2 > 0.0f)) return false;
    const float p0 = dot4_rn(P, px, py, pz);
    // a pixel column far outside the image needs no division: x = p0 / p2 rounds to a column in [0, w)
    // only if -0.5 <= x < w - 0.5 (up to one rounding of the quotient, 1e-7 relative); the margins
    // below (0.01 pixel) keep the test strictly conservative, so the result is unchanged bit for bit
    if (p0 < -0.51f * p2 || p0 > ((float)w - 0.49f) * p2) return false;
    const float p1 = dot4_rn(P + 4, px, py, pz);
    if (p1 < -0.51f * p2 || p1 > ((float)h - 0.49f) * p2) return false;
    xr = rintf(__fdiv_rn(p0, p2));
    yr = rintf(__fdiv_rn(p1, p2));
    // -0.0f >= 0 holds (pixel 0); NaN / inf compare false, like the reference's
    // cast-then-compare on every platform.
    return (xr >= 0.0f) & (yr >= 0.0f) & (xr < (float)w) & (yr < (float)h);
}

__device__ __forceinline__ void voxel_coords(const BpArgs &a, int b, int n, float &px,
                                             float &py, float &pz) {
    if (a.points != nullptr) {
        px = __ldg(a.points + n);
        py = __ldg(a.points + a.N + n);
        pz = __ldg(a.points + 2ll * a.N + n);
    } else {
        const int t = a.div_z.div(n);
        const int iz = n - t * a.Z;
        const int ix = a.div_y.div(t);
        const int iy = t - ix * a.Y;
        px = __fadd_rn(__fmul_rn((float)ix, a.vs[0]), a.new_origin[b][0]);
        py = __fadd_rn(__fmul_rn((float)iy, a.vs[1]), a.new_origin[b][1]);
        pz = __fadd_rn(__fmul_rn((float)iz, a.vs[2]), a.new_origin[b][2]);
    }
}

// Phase 1.  For every voxel j of the tile, row j of `table` receives the
// offsets of the pixels it hits, in ascending view order (float4 units if VEC,
// else elements); the return value is the number of hits of the voxel this
// lane works for (lane & (T-1)), identical on all lanes of that voxel.  The
// per-voxel valid flag and view count are stored from here, 32 voxels per
// coalesced store.
template <bool VEC>
__device__ __forceinline__ int fill_table(const BpArgs &a, const float *sP, int *table, int b,
                                          int n0, int lane) {
    constexpr unsigned FULL = 0xffffffffu;
    const int T = 1 << a.tile_log2;
    const int j = lane & (T - 1);
    const int vq = lane >> a.tile_log2;
    const int vstep = 32 >> a.tile_log2;
    const int n = n0 + j;
    // lanes j, j+T, j+2T, ... work for voxel j on consecutive views
    unsigned pattern = 0;
    for (int q = 0; q < vstep; ++q) pattern |= 1u << (j + (q << a.tile_log2));
    const unsigned below = pattern & ((1u << lane) - 1u);
    float px = 0.f, py = 0.f, pz = 0.f;
    if (n < a.N) voxel_coords(a, b, n, px, py, pz);
    const int h = a.hw[b][0], w = a.hw[b][1];
    int *row = table + j * a.table_stride;
    int base = 0;
    for (int v0 = 0; v0 < a.V; v0 += vstep) {
        const int v = v0 + vq;
        bool ok = false;
        int off = 0;
        if (n < a.N && v < a.V) {
            float xr, yr;
            ok = project_rn(sP + 12 * v, px, py, pz, h, w, xr, yr);
            if (VEC)
                off = v * a.q_v + (int)yr * a.q_h + (int)xr * a.q_w;
            else
                off = (int)(v * a.fs_v + (long long)yr * a.fs_h + (long long)xr * a.fs_w);
        }
        const unsigned m = __ballot_sync(FULL, ok);
        if (ok) row[base + __popc(m & below)] = off;
        base += __popc(m & pattern);
    }
    if (lane < T && n < a.N) {
        if (a.valid) a.valid[(long long)b * a.N + n] = base > 0;
        if (a.count) a.count[(long long)b * a.N + n] = base;
    }
    __syncwarp();
    return base;
}

template <typename OutT>
__device__ __forceinline__ void store4(OutT *p, float4 v);

template <>
__device__ __forceinline__ void store4<float>(float *p, float4 v) {
    __stcs(reinterpret_cast<float4 *>(p), v);
}

template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16 *p, float4 v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<unsigned *>(&lo);
    u.y = *reinterpret_cast<unsigned *>(&hi);
    __stcs(reinterpret_cast<uint2 *>(p), u);
}

// ---------------------------------------------------------------------------
// Vectorised kernel: channels-last features (fs_c == 1), channels-last volume
// (os_c == 1).  G lanes per voxel, CPL float4 per lane; G * CPL * 4 >= C.
// Three schedules of phase 2, chosen by the number of views:
//   MODE_DEEP   one voxel per group at a time, DEPTH pixels of it in flight
//               (many views: the hit list of a voxel is long);
//   MODE_WIDE   WIDTH voxels per group at a time, one pixel of each in flight
//               (few views: most hit lists hold 0..2 pixels);
//   MODE_SINGLE one view: no table at all, the offset stays in a register and
//               the mean is the feature itself.
// ---------------------------------------------------------------------------
enum { MODE_DEEP = 0, MODE_SINGLE = 1, MODE_WIDE = 2 };

template <int CPL>
__device__ __forceinline__ void zero4(float4 (&acc)[CPL]) {
#pragma unroll
    for (int i = 0; i < CPL; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// One 16-byte unit per lane and step: 4 fp32 channels, or (FB) 8 bf16 channels widened exactly
// into two float4 -- the sums stay fp32 either way.
template <int G, int CPL, bool FB>
__device__ __forceinline__ void gather4(const void *feat, int off, int t, const bool (&ch_ok)[CPL],
                                        float4 (&r)[CPL * (FB ? 2 : 1)]) {
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        if constexpr (!FB) {
            r[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (off >= 0 && ch_ok[i]) r[i] = __ldg(reinterpret_cast<const float4 *>(feat) + off + t + i * G);
        } else {
            uint4 u = make_uint4(0u, 0u, 0u, 0u);
            if (off >= 0 && ch_ok[i]) u = __ldg(reinterpret_cast<const uint4 *>(feat) + off + t + i * G);
            // bf16 -> fp32 is a 16-bit shift
            r[2 * i] = make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u),
                                   __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
            r[2 * i + 1] = make_float4(__uint_as_float(u.z << 16), __uint_as_float(u.z & 0xffff0000u),
                                       __uint_as_float(u.w << 16), __uint_as_float(u.w & 0xffff0000u));
        }
    }
}

template <int CPL>
__device__ __forceinline__ void add4(float4 (&acc)[CPL], const float4 (&r)[CPL]) {
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
        acc[i].x += r[i].x;
        acc[i].y += r[i].y;
        acc[i].z += r[i].z;
        acc[i].w += r[i].w;
    }
}

// sum / count (count converted to float, :72; 0 where unseen, :74), then store.  Called by all lanes
// of the warp (`active` = this lane's voxel exists): one variant exchanges data between lanes.
template <int G, int CPL, bool DIVIDE, typename OutT, bool FB>
__device__ __forceinline__ void finish_voxel(const BpArgs &a, OutT *vol, int n, int t, int cnt,
                                             float4 (&acc)[CPL * (FB ? 2 : 1)], const bool (&ch_ok)[CPL],
                                             bool active) {
    constexpr int NF = FB ? 2 : 1;   // float4 per 16-byte feature unit
    if (DIVIDE && a.reduce == IVX_REDUCE_MEAN && cnt > 1) {
        const float c = (float)cnt;
#pragma unroll
        for (int i = 0; i < CPL * NF; ++i) {
            acc[i].x = __fdiv_rn(acc[i].x, c);
            acc[i].y = __fdiv_rn(acc[i].y, c);
            acc[i].z = __fdiv_rn(acc[i].z, c);
            acc[i].w = __fdiv_rn(acc[i].w, c);
        }
    }  // cnt == 1: x / 1.0f == x;  cnt == 0: the accumulator is still zero
    OutT *dst = vol + (long long)n * a.os_n;
    if constexpr (FB && sizeof(OutT) == 4) {
        // 8 bf16 channels in, 8 fp32 channels (32 bytes) out per lane: stored as they are, every store
        // instruction would write 16-byte pieces 32 bytes apart.  Neighbouring lanes swap halves instead,
        // so that each instruction writes whole 32-byte sectors: the even lane of a pair stores the first
        // half of both units, the odd lane the second half.
        const bool odd = t & 1;
#pragma unroll
        for (int i = 0; i < CPL; ++i) {
            const float4 send = odd ? acc[2 * i] : acc[2 * i + 1];
            float4 recv;
            recv.x = __shfl_xor_sync(0xffffffffu, send.x, 1), recv.y = __shfl_xor_sync(0xffffffffu, send.y, 1);
            recv.z = __shfl_xor_sync(0xffffffffu, send.z, 1), recv.w = __shfl_xor_sync(0xffffffffu, send.w, 1);
            const int unit_e = (t & ~1) + i * G;                 // unit of the even lane; the odd lane's is + 1
            float *p = reinterpret_cast<float *>(dst) + unit_e * 8 + (odd ? 4 : 0);
            if (active && unit_e * 8 < a.C) store4<float>(p, odd ? recv : acc[2 * i]);
            if (active && (unit_e + 1) * 8 < a.C) store4<float>(p + 8, odd ? acc[2 * i + 1] : recv);
        }
        return;
    }
    if (!active) return;
#pragma unroll
    for (int i = 0; i < CPL; ++i)
        if (ch_ok[i]) {
            if constexpr (FB && sizeof(OutT) == 2) {   // 8 bf16 channels: one 16-byte streaming store
                const float4 lo = acc[2 * i], hi = acc[2 * i + 1];
                __nv_bfloat162 p0 = __floats2bfloat162_rn(lo.x, lo.y), p1 = __floats2bfloat162_rn(lo.z, lo.w);
                __nv_bfloat162 p2 = __floats2bfloat162_rn(hi.x, hi.y), p3 = __floats2bfloat162_rn(hi.z, hi.w);
                uint4 u;
                u.x = *reinterpret_cast<unsigned *>(&p0), u.y = *reinterpret_cast<unsigned *>(&p1);
                u.z = *reinterpret_cast<unsigned *>(&p2), u.w = *reinterpret_cast<unsigned *>(&p3);
                __stcs(reinterpret_cast<uint4 *>(dst + (t + i * G) * 8), u);
            } else {
#pragma unroll
                for (int k = 0; k < NF; ++k) store4<OutT>(dst + ((t + i * G) * NF + k) * 4, acc[i * NF + k]);
            }
        }
}

template <int G, int CPL, int MODE, typename OutT, bool FB>
__global__ void __launch_bounds__(256, (CPL * (FB ? 2 : 1) <= 2 ? 4 : 2)) bp_vec_kernel(const __grid_constant__ BpArgs a) {
    constexpr int NF = FB ? 2 : 1;             // float4 accumulators per 16-byte feature unit
    constexpr int ACC = CPL * NF;
    extern __shared__ int smem[];
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int VPI = 32 / G;                // voxels a warp works on at once
    constexpr int DEPTH = (ACC >= 4) ? 2 : 4;  // MODE_DEEP: pixels in flight per voxel
    constexpr int WIDTH = (ACC >= 4) ? 1 : (ACC == 2 ? 2 : 4);  // voxels in flight per group
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    const int T = 1 << a.tile_log2;

    float *sP = reinterpret_cast<float *>(smem);
    for (int i = threadIdx.x; i < a.V * 12; i += blockDim.x)
        sP[i] = __ldg(a.proj + b * a.proj_stride_b + i);
    __syncthreads();
    int *table = smem + ((a.V * 12 + 3) & ~3) + warp * T * a.table_stride;

    const void *feat4 = FB ? static_cast<const void *>(reinterpret_cast<const __nv_bfloat16 *>(a.feat) + b * a.fs_b)
                           : static_cast<const void *>(a.feat + b * a.fs_b);
    OutT *vol = reinterpret_cast<OutT *>(a.vol) + b * a.os_b;
    const int g = lane / G;
    const int t = lane % G;
    const int ntiles = (a.N + T - 1) >> a.tile_log2;

    bool ch_ok[CPL];
#pragma unroll
    for (int i = 0; i < CPL; ++i) ch_ok[i] = (t + i * G) * 4 * NF < a.C;

    for (int tile = blockIdx.x * nwarps + warp; tile < ntiles; tile += gridDim.x * nwarps) {
        const int n0 = tile << a.tile_log2;

        if constexpr (MODE == MODE_SINGLE) {
            // lane j projects voxel j and keeps the offset of its pixel
            int my_off = -1;
            {
                const int n = n0 + lane;
                if (lane < T && n < a.N) {
                    float px, py, pz, xr, yr;
                    voxel_coords(a, b, n, px, py, pz);
                    if (project_rn(sP, px, py, pz, a.hw[b][0], a.hw[b][1], xr, yr))
                        my_off = (int)yr * a.q_h + (int)xr * a.q_w;
                    if (a.valid) a.valid[(long long)b * a.N + n] = my_off >= 0;
                    if (a.count) a.count[(long long)b * a.N + n] = my_off >= 0;
                }
            }
            for (int j0 = 0; j0 < T; j0 += VPI * WIDTH) {
                int o[WIDTH];
                float4 r[WIDTH][ACC];
#pragma unroll
                for (int u = 0; u < WIDTH; ++u) {
                    const int j = j0 + u * VPI + g;
                    o[u] = __shfl_sync(FULL, my_off, j & 31);
                    if (j >= T) o[u] = -1;
                }
#pragma unroll
                for (int u = 0; u < WIDTH; ++u) gather4<G, CPL, FB>(feat4, o[u], t, ch_ok, r[u]);
#pragma unroll
                for (int u = 0; u < WIDTH; ++u) {
                    const int j = j0 + u * VPI + g;
                    const int n = n0 + j;
                    finish_voxel<G, CPL, false, OutT, FB>(a, vol, n, t, o[u] >= 0, r[u], ch_ok, j < T && n < a.N);
                }
            }
        } else if constexpr (MODE == MODE_WIDE) {
            const int my_cnt = fill_table<true>(a, sP, table, b, n0, lane);
            for (int j0 = 0; j0 < T; j0 += VPI * WIDTH) {
                int cnt[WIDTH];
                float4 acc[WIDTH][ACC];
                int most = 0;
#pragma unroll
                for (int u = 0; u < WIDTH; ++u) {
                    const int j = j0 + u * VPI + g;
                    cnt[u] = __shfl_sync(FULL, my_cnt, j & 31);
                    if (j >= T || n0 + j >= a.N) cnt[u] = 0;
                    most = max(most, cnt[u]);
                    zero4<ACC>(acc[u]);
                }
                const int *rows = table + (j0 + g) * a.table_stride;
                for (int k = 0; k < most; ++k) {
                    float4 r[WIDTH][ACC];
#pragma unroll
                    for (int u = 0; u < WIDTH; ++u) {
                        const int o = (k < cnt[u]) ? rows[u * VPI * a.table_stride + k] : -1;
                        gather4<G, CPL, FB>(feat4, o, t, ch_ok, r[u]);
                    }
#pragma unroll
                    for (int u = 0; u < WIDTH; ++u) add4<ACC>(acc[u], r[u]);
                }
#pragma unroll
                for (int u = 0; u < WIDTH; ++u) {
                    const int j = j0 + u * VPI + g;
                    const int n = n0 + j;
                    finish_voxel<G, CPL, true, OutT, FB>(a, vol, n, t, cnt[u], acc[u], ch_ok, j < T && n < a.N);
                }
            }
            __syncwarp();  // the table is rewritten by the next tile
        } else {
            const int my_cnt = fill_table<true>(a, sP, table, b, n0, lane);
            for (int j0 = 0; j0 < T; j0 += VPI) {
                const int j = j0 + g;
                const int n = n0 + j;
                const bool active = (j < T) && (n < a.N);
                int cnt = __shfl_sync(FULL, my_cnt, j & 31);
                if (!active) cnt = 0;
                const int *row = table + j * a.table_stride;
                float4 acc[ACC];
                zero4<ACC>(acc);
                for (int k0 = 0; k0 < cnt; k0 += DEPTH) {
                    float4 r[DEPTH][ACC];
#pragma unroll
                    for (int u = 0; u < DEPTH; ++u) {
                        const int o = (k0 + u < cnt) ? row[k0 + u] : -1;
                        gather4<G, CPL, FB>(feat4, o, t, ch_ok, r[u]);
                    }
#pragma unroll
                    for (int u = 0; u < DEPTH; ++u) add4<ACC>(acc, r[u]);  // ascending view order
                }
                finish_voxel<G, CPL, true, OutT, FB>(a, vol, n, t, cnt, acc, ch_ok, active);
            }
            __syncwarp();  // the table is rewritten by the next tile
        }
    }
}

// ---------------------------------------------------------------------------
// Backward of the fused back-projection + view mean with respect to the
// features (what autograd does through :157 and :70-74 in the reference):
//   grad_feat[b, v, :, y, x] += grad_volume[b, :, n] / count[b, n]   for every hit (v, n)
// Same tiling and hit lists as the forward kernel; several voxels may hit one
// pixel, so the scatter uses 128-bit vector reductions (red.global.add.v4.f32).
// ---------------------------------------------------------------------------
template <int G, int CPL>
__global__ void __launch_bounds__(256, 4) bp_backward_kernel(const __grid_constant__ BpArgs a) {
    extern __shared__ int smem[];
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int VPI = 32 / G;
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    const int T = 1 << a.tile_log2;

    float *sP = reinterpret_cast<float *>(smem);
    for (int i = threadIdx.x; i < a.V * 12; i += blockDim.x)
        sP[i] = __ldg(a.proj + b * a.proj_stride_b + i);
    __syncthreads();
    int *table = smem + ((a.V * 12 + 3) & ~3) + warp * T * a.table_stride;

    // here a.feat is the GRADIENT of the features (written), a.vol the gradient of the volume (read)
    float4 *gfeat4 = reinterpret_cast<float4 *>(const_cast<float *>(a.feat) + b * a.fs_b);
    const float *gvol = reinterpret_cast<const float *>(a.vol) + b * a.os_b;
    const int g = lane / G;
    const int t = lane % G;
    const int ntiles = (a.N + T - 1) >> a.tile_log2;

    for (int tile = blockIdx.x * nwarps + warp; tile < ntiles; tile += gridDim.x * nwarps) {
        const int n0 = tile << a.tile_log2;
        const int my_cnt = fill_table<true>(a, sP, table, b, n0, lane);
        for (int j0 = 0; j0 < T; j0 += VPI) {
            const int j = j0 + g;
            const int n = n0 + j;
            int cnt = __shfl_sync(FULL, my_cnt, j & 31);
            if (j >= T || n >= a.N) cnt = 0;
            if (cnt == 0) continue;
            const int *row = table + j * a.table_stride;
            const float c = (float)cnt;
#pragma unroll
            for (int i = 0; i < CPL; ++i) {
                if ((t + i * G) * 4 >= a.C) continue;
                float4 gv = __ldcs(reinterpret_cast<const float4 *>(gvol + (long long)n * a.os_n) + t + i * G);
                if (a.reduce == IVX_REDUCE_MEAN) {
                    gv.x = __fdiv_rn(gv.x, c), gv.y = __fdiv_rn(gv.y, c);
                    gv.z = __fdiv_rn(gv.z, c), gv.w = __fdiv_rn(gv.w, c);
                }
                for (int k = 0; k < cnt; ++k) atomicAdd(gfeat4 + row[k] + t + i * G, gv);
            }
        }
        __syncwarp();
    }
}

static void *pick_backward_kernel(int C) {
    const int c4 = (C + 3) / 4;
    if (c4 <= 8) return (void *)bp_backward_kernel<8, 1>;
    if (c4 <= 16) return (void *)bp_backward_kernel<16, 1>;
    if (c4 <= 32) return (void *)bp_backward_kernel<32, 1>;
    if (c4 <= 64) return (void *)bp_backward_kernel<32, 2>;
    if (c4 <= 128) return (void *)bp_backward_kernel<32, 4>;
    if (c4 <= 256) return (void *)bp_backward_kernel<32, 8>;
    return nullptr;
}

// ---------------------------------------------------------------------------
// Strided kernel: any feature strides (the reference's NCHW slice, :69), any
// channel count, any fp32 volume strides.  Lanes own channels.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bp_strided_kernel(const __grid_constant__ BpArgs a) {
    extern __shared__ int smem[];
    constexpr unsigned FULL = 0xffffffffu;
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    const int T = 1 << a.tile_log2;

    float *sP = reinterpret_cast<float *>(smem);
    for (int i = threadIdx.x; i < a.V * 12; i += blockDim.x)
        sP[i] = __ldg(a.proj + b * a.proj_stride_b + i);
    __syncthreads();
    int *table = smem + ((a.V * 12 + 3) & ~3) + warp * T * a.table_stride;

    const float *feat = a.feat + b * a.fs_b;
    float *vol = reinterpret_cast<float *>(a.vol) + b * a.os_b;
    const int ntiles = (a.N + T - 1) >> a.tile_log2;

    for (int tile = blockIdx.x * nwarps + warp; tile < ntiles; tile += gridDim.x * nwarps) {
        const int n0 = tile << a.tile_log2;
        const int my_cnt = fill_table<false>(a, sP, table, b, n0, lane);
        for (int j = 0; j < T; ++j) {
            const int n = n0 + j;
            if (n >= a.N) break;
            const int *row = table + j * a.table_stride;
            const int cnt = __shfl_sync(FULL, my_cnt, j);
            const float fc = (float)cnt;
            for (int cb = 0; cb < a.C; cb += 32) {
                const int c = cb + lane;
                if (c < a.C) {
                    float acc = 0.f;
                    for (int k = 0; k < cnt; ++k) acc += __ldg(feat + row[k] + c * a.fs_c);
                    if (a.reduce == IVX_REDUCE_MEAN) acc = cnt ? __fdiv_rn(acc, fc) : 0.f;
                    vol[c * a.os_c + (long long)n * a.os_n] = acc;
                }
            }
        }
        __syncwarp();
    }
}

// x, y, valid of every voxel-view (the reference's locals at :151-154).
__global__ void __launch_bounds__(256) project_indices_kernel(
    const float *__restrict__ points, ivx_grid grid, float nox, float noy, float noz, int N,
    const float *__restrict__ proj, int V, int h, int w, int32_t *__restrict__ x,
    int32_t *__restrict__ y, uint8_t *__restrict__ valid) {
    const int v = blockIdx.y;
    __shared__ float P[12];
    if (threadIdx.x < 12) P[threadIdx.x] = proj[v * 12 + threadIdx.x];
    __syncthreads();
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
        float px, py, pz;
        if (points != nullptr) {
            px = points[n];
            py = points[N + n];
            pz = points[2ll * N + n];
        } else {
            const int iz = n % grid.n_voxels[2];
            const int t = n / grid.n_voxels[2];
            const int iy = t % grid.n_voxels[1];
            const int ix = t / grid.n_voxels[1];
            px = __fadd_rn(__fmul_rn((float)ix, grid.voxel_size[0]), nox);
            py = __fadd_rn(__fmul_rn((float)iy, grid.voxel_size[1]), noy);
            pz = __fadd_rn(__fmul_rn((float)iz, grid.voxel_size[2]), noz);
        }
        float xr, yr;
        const bool ok = project_rn(P, px, py, pz, h, w, xr, yr);
        const long long o = (long long)v * N + n;
        if (x) x[o] = ok ? (int)xr : INT32_MIN;
        if (y) y[o] = ok ? (int)yr : INT32_MIN;
        valid[o] = ok;
    }
}

__global__ void __launch_bounds__(256) finalize_mean_kernel(
    const float4 *__restrict__ sum, const int32_t *__restrict__ count, long long n_rows, int c4,
    int N, void *vol, long long os_b, long long os_c, long long os_n, int dtype,
    uint8_t *__restrict__ valid) {
    const long long total = n_rows * c4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / c4;
        const int q = (int)(i - row * c4);
        const int cnt = __ldg(count + row);
        float4 s = __ldcs(sum + i);
        const float c = (float)cnt;
        s.x = cnt ? __fdiv_rn(s.x, c) : 0.f;
        s.y = cnt ? __fdiv_rn(s.y, c) : 0.f;
        s.z = cnt ? __fdiv_rn(s.z, c) : 0.f;
        s.w = cnt ? __fdiv_rn(s.w, c) : 0.f;
        const long long b = row / N, n = row - b * N;
        const long long base = b * os_b + n * os_n;
        if (os_c == 1) {
            if (dtype == IVX_F32)
                store4<float>(reinterpret_cast<float *>(vol) + base + q * 4, s);
            else
                store4<__nv_bfloat16>(reinterpret_cast<__nv_bfloat16 *>(vol) + base + q * 4, s);
        } else {
            float *o = reinterpret_cast<float *>(vol) + base + (long long)q * 4 * os_c;
            o[0] = s.x;
            o[os_c] = s.y;
            o[2 * os_c] = s.z;
            o[3 * os_c] = s.w;
        }
        if (q == 0 && valid) valid[row] = cnt > 0;
    }
}

// NCHW -> NHWC through a 32x33 shared-memory tile, coalesced on both sides.
// Images are indexed (outer, inner) with their own strides so that a [B][V]
// batch of maps needs one launch; dst is dense [outer*inner][H][W][C].
__global__ void __launch_bounds__(256) repack_nchw_nhwc_kernel(
    const float *__restrict__ src, int n_inner, long long s_o, long long s_i, long long s_c,
    long long s_h, long long s_w, int C, int H, int W, float *__restrict__ dst) {
    __shared__ float tile[32][33];
    const int img = blockIdx.z;
    const int io = img / n_inner, ii = img - io * n_inner;
    const float *base = src + io * s_o + ii * s_i;
    const int HW = H * W;
    const int p0 = blockIdx.x * 32;  // pixel tile
    const int c0 = blockIdx.y * 32;  // channel tile
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int r = ty; r < 32; r += 8) {
        const int c = c0 + r, p = p0 + tx;
        float v = 0.f;
        if (c < C && p < HW) {
            const int yy = p / W, xx = p - yy * W;
            v = __ldcs(base + c * s_c + yy * s_h + xx * s_w);
        }
        tile[r][tx] = v;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int p = p0 + r, c = c0 + tx;
        if (c < C && p < HW) dst[((long long)img * HW + p) * C + c] = tile[tx][r];
    }
}

static int launch_repack(const float *src, int n_outer, int n_inner, long long s_o, long long s_i,
                         long long s_c, long long s_h, long long s_w, int C, int H, int W, float *dst,
                         cudaStream_t stream) {
    const long long hw = (long long)H * W;
    IVX_REQUIRE(hw < (1ll << 31) && (C + 31) / 32 <= 65535 && (long long)n_outer * n_inner <= 65535,
                "map or batch too large for the repack kernel");
    dim3 grid((unsigned)((hw + 31) / 32), (unsigned)((C + 31) / 32), (unsigned)(n_outer * n_inner));
    repack_nchw_nhwc_kernel<<<grid, 256, 0, stream>>>(src, n_inner, s_o, s_i, s_c, s_h, s_w, C, H, W, dst);
    IVX_CHECK_LAUNCH("repack launch");
    return IVX_OK;
}

// ------------------------------- host side --------------------------------

static int pick_tile_log2(long long n_points, int n_scenes, int table_stride) {
    // largest tile whose hit table stays small (<= 2 KB per warp keeps the
    // occupancy register-bound) and that still gives every SM ~32 warps of work
    int lg = 5;
    while (lg > 2 && (1 << lg) * table_stride > 512) --lg;
    for (; lg > 2; --lg) {
        const long long tiles = ((n_points + (1 << lg) - 1) >> lg) * n_scenes;
        if (tiles >= (long long)kNumSMs * 32) break;
    }
    return lg;
}

template <int MODE, typename OutT, bool FB>
static void *pick_vec_kernel(int C) {
    const int units = FB ? (C + 7) / 8 : (C + 3) / 4;   // 16-byte units per pixel
    if (units <= 8) return (void *)bp_vec_kernel<8, 1, MODE, OutT, FB>;
    if (units <= 16) {
        // several views, 64 channels (nuScenes): 8 lanes x 2 units per voxel instead of 16 x 1 -- four voxels
        // per warp step, half the table reads / shuffles / predicates per gathered byte (the kernel was
        // issue-bound there: 63 % issue slots busy at 53 % of the DRAM bandwidth)
        // (measured: nuScenes 200x200x4 0.775 -> 0.82 of HBM, 312x312x12 0.72 -> 0.815; 4 lanes x 4 units is
        // slower again, 0.63, and 16 x 4 for 256 channels changes nothing)
        if (MODE != MODE_SINGLE && units > 8) return (void *)bp_vec_kernel<8, 2, MODE, OutT, FB>;
        return (void *)bp_vec_kernel<16, 1, MODE, OutT, FB>;
    }
    if (units <= 32) return (void *)bp_vec_kernel<32, 1, MODE, OutT, FB>;
    if (units <= 64) return (void *)bp_vec_kernel<32, 2, MODE, OutT, FB>;
    if (units <= 128) return (void *)bp_vec_kernel<32, 4, MODE, OutT, FB>;
    if (!FB && units <= 256) return (void *)bp_vec_kernel<32, 8, MODE, OutT, false>;
    return nullptr;
}

constexpr int kWideMaxViews = 8;  // up to here most voxels are hit by 0..2 views

template <typename OutT, bool FB>
static void *pick_vec_kernel(int C, int V) {
    if (V == 1) return pick_vec_kernel<MODE_SINGLE, OutT, FB>(C);
    if (V <= kWideMaxViews) return pick_vec_kernel<MODE_WIDE, OutT, FB>(C);
    return pick_vec_kernel<MODE_DEEP, OutT, FB>(C);
}

static bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// One launch over at most IVX_MAX_SCENES_PER_LAUNCH scenes.
static int launch_scenes(BpArgs &a, int n_scenes, int vol_dtype, cudaStream_t stream, bool backward = false) {
    const int upe = a.feat_bf16 ? 8 : 4;   // elements per 16-byte unit of the feature maps
    const bool feat_vec = a.fs_c == 1 && a.C % upe == 0 && aligned16(a.feat) && a.fs_b % upe == 0 &&
                          a.fs_v % upe == 0 && a.fs_h % upe == 0 && a.fs_w % upe == 0;
    const size_t esz = vol_dtype == IVX_BF16 ? 2 : 4;
    const int vupe = (a.feat_bf16 && vol_dtype == IVX_BF16) ? 8 : 4;   // elements per vector store
    const bool vol_vec = a.os_c == 1 && a.os_b % vupe == 0 && a.os_n % vupe == 0 &&
                         (reinterpret_cast<uintptr_t>(a.vol) % (vupe * esz)) == 0;
    void *kernel = nullptr;
    if (backward) {
        if (!(feat_vec && vol_vec) || vol_dtype != IVX_F32 || (kernel = pick_backward_kernel(a.C)) == nullptr)
            return fail(IVX_ERR_UNSUPPORTED,
                        "backward needs channels-last fp32 gradients with C %% 4 == 0 and C <= 1024");
    } else if (feat_vec && vol_vec)
        kernel = a.feat_bf16 ? (vol_dtype == IVX_BF16 ? pick_vec_kernel<__nv_bfloat16, true>(a.C, a.V)
                                                      : pick_vec_kernel<float, true>(a.C, a.V))
                             : (vol_dtype == IVX_BF16 ? pick_vec_kernel<__nv_bfloat16, false>(a.C, a.V)
                                                      : pick_vec_kernel<float, false>(a.C, a.V));
    if (kernel != nullptr) {
        // table entries are int32 offsets in float4 units inside one scene's block
        long long hmax = 0, wmax = 0;
        for (int b = 0; b < n_scenes; ++b) {
            hmax = a.hw[b][0] > hmax ? a.hw[b][0] : hmax;
            wmax = a.hw[b][1] > wmax ? a.hw[b][1] : wmax;
        }
        if ((llabs(a.fs_v) * a.V + llabs(a.fs_h) * hmax + llabs(a.fs_w) * wmax) / upe >= (1ll << 31))
            return fail(IVX_ERR_UNSUPPORTED, "feature block of one scene exceeds 2^33 elements");
        const int q = backward ? 4 : upe;
        a.q_v = (int)(a.fs_v / q), a.q_h = (int)(a.fs_h / q), a.q_w = (int)(a.fs_w / q);
    }
    if (kernel == nullptr && a.feat_bf16)
        return fail(IVX_ERR_UNSUPPORTED,
                    "bf16 feature maps need the channels-last layout, C %% 8 == 0, C <= 1024 and 16-byte aligned rows");
    if (kernel == nullptr) {
        if (vol_dtype != IVX_F32)
            return fail(IVX_ERR_UNSUPPORTED,
                        "bf16 volumes need channels-last features and volume (C %% 4 == 0, C <= 1024)");
        kernel = (void *)bp_strided_kernel;
        // table entries are int32 element offsets inside one scene's block
        const long long span = llabs(a.fs_v) * a.V + llabs(a.fs_c) * a.C;
        long long hmax = 0, wmax = 0;
        for (int b = 0; b < n_scenes; ++b) {
            hmax = a.hw[b][0] > hmax ? a.hw[b][0] : hmax;
            wmax = a.hw[b][1] > wmax ? a.hw[b][1] : wmax;
        }
        if (span + llabs(a.fs_h) * hmax + llabs(a.fs_w) * wmax >= (1ll << 31))
            return fail(IVX_ERR_UNSUPPORTED, "feature block of one scene exceeds 2^31 elements");
    }

    a.table_stride = (a.V + 3) & ~3;
    a.tile_log2 = pick_tile_log2(a.N, n_scenes, a.table_stride);
    const int T = 1 << a.tile_log2;
    const size_t proj_words = (a.V * 12 + 3) & ~3;
    int warps = 8;
    size_t smem = 0;
    for (;; warps >>= 1) {
        smem = (proj_words + (size_t)warps * T * a.table_stride) * sizeof(int);
        if (smem <= 96 * 1024 || warps == 1) break;
    }
    if (smem > 200 * 1024) return fail(IVX_ERR_UNSUPPORTED, "too many views for one tile table");
    if (smem > 48 * 1024) {
        int rc = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)smem),
                            "cudaFuncSetAttribute");
        if (rc != IVX_OK) return rc;
    }
    const long long ntiles = ((long long)a.N + T - 1) >> a.tile_log2;
    long long gx = (ntiles + warps - 1) / warps;
    if (gx > 65535ll * 16) gx = 65535ll * 16;  // grid-stride loop covers the rest
    dim3 grid((unsigned)gx, (unsigned)n_scenes, 1), block(warps * 32, 1, 1);
    void *params[] = {&a};
    int rc = check_cuda(cudaLaunchKernel(kernel, grid, block, params, smem, stream), "backproject launch");
    return rc;
}

static int validate_grid(const float *points, const ivx_grid *grid, const float *new_origin,
                         int64_t n_points) {
    IVX_REQUIRE(n_points > 0 && n_points < (1ll << 31), "n_points must be in [1, 2^31)");
    if (points == nullptr) {
        IVX_REQUIRE(grid != nullptr && new_origin != nullptr,
                    "either points or grid + new_origin must be given");
        const long long n = (long long)grid->n_voxels[0] * grid->n_voxels[1] * grid->n_voxels[2];
        IVX_REQUIRE(grid->n_voxels[0] > 0 && grid->n_voxels[1] > 0 && grid->n_voxels[2] > 0 &&
                        n == n_points,
                    "grid %dx%dx%d does not hold n_points=%lld", grid->n_voxels[0],
                    grid->n_voxels[1], grid->n_voxels[2], (long long)n_points);
    }
    return IVX_OK;
}

static int backproject_batch(const float *feat, const long long fs[5], int B, int V, int C,
                             const int32_t *hw, bool hw_per_scene, const float *proj,
                             long long proj_stride_b, const float *points, const ivx_grid *grid,
                             const float *new_origin, bool origin_per_scene, int64_t N, void *volume,
                             const int64_t *vs, int vol_dtype, int reduce, uint8_t *valid,
                             int32_t *count, cudaStream_t stream, bool backward = false,
                             bool feat_bf16 = false) {
    const size_t esz = vol_dtype == IVX_BF16 ? 2 : 4;
    for (int b0 = 0; b0 < B; b0 += IVX_MAX_SCENES_PER_LAUNCH) {
        const int nb = (B - b0 < IVX_MAX_SCENES_PER_LAUNCH) ? B - b0 : IVX_MAX_SCENES_PER_LAUNCH;
        BpArgs a;
        a.feat = feat_bf16 ? reinterpret_cast<const float *>(reinterpret_cast<const __nv_bfloat16 *>(feat) + b0 * fs[0])
                           : feat + b0 * fs[0];
        a.feat_bf16 = feat_bf16 ? 1 : 0;
        a.fs_b = fs[0], a.fs_v = fs[1], a.fs_c = fs[2], a.fs_h = fs[3], a.fs_w = fs[4];
        a.proj = proj + b0 * proj_stride_b;
        a.proj_stride_b = proj_stride_b;
        a.points = points;
        a.V = V, a.C = C, a.N = (int)N;
        a.X = a.Y = a.Z = 1;
        a.div_z = a.div_y = make_fastdiv(1);
        a.q_v = a.q_h = a.q_w = 0;
        a.vs[0] = a.vs[1] = a.vs[2] = 0.f;
        if (points == nullptr) {
            a.X = grid->n_voxels[0], a.Y = grid->n_voxels[1], a.Z = grid->n_voxels[2];
            a.div_z = make_fastdiv(a.Z), a.div_y = make_fastdiv(a.Y);
            for (int k = 0; k < 3; ++k) a.vs[k] = grid->voxel_size[k];
        }
        a.vol = static_cast<char *>(volume) + (size_t)(b0 * vs[0]) * esz;
        a.os_b = vs[0], a.os_c = vs[1], a.os_n = vs[2];
        a.valid = valid ? valid + (long long)b0 * N : nullptr;
        a.count = count ? count + (long long)b0 * N : nullptr;
        a.reduce = reduce;
        for (int b = 0; b < nb; ++b) {
            const int32_t *p = hw + (hw_per_scene ? 2 * (b0 + b) : 0);
            a.hw[b][0] = p[0], a.hw[b][1] = p[1];
            for (int k = 0; k < 3; ++k)
                a.new_origin[b][k] =
                    new_origin ? new_origin[(origin_per_scene ? 3 * (b0 + b) : 0) + k] : 0.f;
        }
        int rc = launch_scenes(a, nb, vol_dtype, stream, backward);
        if (rc != IVX_OK) return rc;
    }
    return IVX_OK;
}

}  // namespace ivx

using namespace ivx;

extern "C" int ivx_project_indices(const float *points, const ivx_grid *grid, const float *new_origin,
                                   int64_t n_points, const float *proj, int n_views, int height,
                                   int width, int32_t *x, int32_t *y, uint8_t *valid, void *stream) {
    int rc = validate_grid(points, grid, new_origin, n_points);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(proj != nullptr && valid != nullptr, "proj and valid must not be NULL");
    IVX_REQUIRE(n_views > 0 && n_views <= 65535, "n_views out of range");
    IVX_REQUIRE(height > 0 && width > 0, "height/width must be positive");
    ivx_grid g = {{1, 1, 1}, {0.f, 0.f, 0.f}};
    float no[3] = {0.f, 0.f, 0.f};
    if (points == nullptr) {
        g = *grid;
        for (int k = 0; k < 3; ++k) no[k] = new_origin[k];
    }
    long long gx = (n_points + 255) / 256;
    if (gx > kNumSMs * 64) gx = kNumSMs * 64;
    project_indices_kernel<<<dim3((unsigned)gx, (unsigned)n_views), 256, 0, (cudaStream_t)stream>>>(
        points, g, no[0], no[1], no[2], (int)n_points, proj, n_views, height, width, x, y, valid);
    IVX_CHECK_LAUNCH("project_indices launch");
    return IVX_OK;
}

extern "C" int ivx_backproject_scenes(const float *feat, const int64_t *feat_strides, int n_scenes,
                                      int n_views, int n_channels, const int32_t *hw,
                                      const float *proj, const float *points, const ivx_grid *grid,
                                      const float *new_origin, int64_t n_points, void *volume,
                                      const int64_t *vol_strides, int vol_dtype, int reduce,
                                      uint8_t *valid, int32_t *count, void *stream) {
    int rc = validate_grid(points, grid, new_origin, n_points);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(feat && feat_strides && hw && proj && volume && vol_strides, "NULL argument");
    IVX_REQUIRE(n_scenes > 0 && n_channels > 0, "n_scenes and n_channels must be positive");
    IVX_REQUIRE(n_views > 0 && n_views <= IVX_MAX_VIEWS, "n_views must be in [1, %d]", IVX_MAX_VIEWS);
    IVX_REQUIRE(vol_dtype == IVX_F32 || vol_dtype == IVX_BF16, "unknown vol_dtype %d", vol_dtype);
    IVX_REQUIRE(reduce == IVX_REDUCE_MEAN || reduce == IVX_REDUCE_PARTIAL, "unknown reduce %d", reduce);
    if (reduce == IVX_REDUCE_PARTIAL)
        IVX_REQUIRE(count != nullptr && vol_dtype == IVX_F32, "partial sums need count and fp32 volume");
    for (int b = 0; b < n_scenes; ++b)
        IVX_REQUIRE(hw[2 * b] > 0 && hw[2 * b + 1] > 0, "scene %d: height/width must be positive", b);
    const long long fs[5] = {feat_strides[0], feat_strides[1], feat_strides[2], feat_strides[3],
                             feat_strides[4]};
    return backproject_batch(feat, fs, n_scenes, n_views, n_channels, hw, true, proj,
                             12ll * n_views, points, grid, new_origin, true, n_points, volume,
                             vol_strides, vol_dtype, reduce, valid, count, (cudaStream_t)stream);
}

extern "C" int ivx_backproject_scenes_bf16(const void *feat_bf16, const int64_t *feat_strides, int n_scenes,
                                           int n_views, int n_channels, const int32_t *hw, const float *proj,
                                           const float *points, const ivx_grid *grid, const float *new_origin,
                                           int64_t n_points, void *volume, const int64_t *vol_strides,
                                           int vol_dtype, int reduce, uint8_t *valid, int32_t *count,
                                           void *stream) {
    int rc = validate_grid(points, grid, new_origin, n_points);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(feat_bf16 && feat_strides && hw && proj && volume && vol_strides, "NULL argument");
    IVX_REQUIRE(n_scenes > 0 && n_channels > 0, "n_scenes and n_channels must be positive");
    IVX_REQUIRE(n_views > 0 && n_views <= IVX_MAX_VIEWS, "n_views must be in [1, %d]", IVX_MAX_VIEWS);
    IVX_REQUIRE(vol_dtype == IVX_F32 || vol_dtype == IVX_BF16, "unknown vol_dtype %d", vol_dtype);
    IVX_REQUIRE(reduce == IVX_REDUCE_MEAN || reduce == IVX_REDUCE_PARTIAL, "unknown reduce %d", reduce);
    if (reduce == IVX_REDUCE_PARTIAL)
        IVX_REQUIRE(count != nullptr && vol_dtype == IVX_F32, "partial sums need count and fp32 volume");
    for (int b = 0; b < n_scenes; ++b)
        IVX_REQUIRE(hw[2 * b] > 0 && hw[2 * b + 1] > 0, "scene %d: height/width must be positive", b);
    const long long fs[5] = {feat_strides[0], feat_strides[1], feat_strides[2], feat_strides[3],
                             feat_strides[4]};
    return backproject_batch(static_cast<const float *>(feat_bf16), fs, n_scenes, n_views, n_channels, hw, true,
                             proj, 12ll * n_views, points, grid, new_origin, true, n_points, volume,
                             vol_strides, vol_dtype, reduce, valid, count, (cudaStream_t)stream, false, true);
}

extern "C" int ivx_backproject_scenes_backward(const float *grad_volume, const int64_t *gvol_strides,
                                               float *grad_feat, const int64_t *feat_strides, int n_scenes,
                                               int n_views, int n_channels, const int32_t *hw,
                                               const float *proj, const float *points, const ivx_grid *grid,
                                               const float *new_origin, int64_t n_points, int reduce,
                                               void *stream) {
    int rc = validate_grid(points, grid, new_origin, n_points);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(grad_volume && gvol_strides && grad_feat && feat_strides && hw && proj, "NULL argument");
    IVX_REQUIRE(n_scenes > 0 && n_channels > 0, "n_scenes and n_channels must be positive");
    IVX_REQUIRE(n_views > 0 && n_views <= IVX_MAX_VIEWS, "n_views must be in [1, %d]", IVX_MAX_VIEWS);
    IVX_REQUIRE(reduce == IVX_REDUCE_MEAN || reduce == IVX_REDUCE_PARTIAL, "unknown reduce %d", reduce);
    for (int b = 0; b < n_scenes; ++b)
        IVX_REQUIRE(hw[2 * b] > 0 && hw[2 * b + 1] > 0, "scene %d: height/width must be positive", b);
    const long long fs[5] = {feat_strides[0], feat_strides[1], feat_strides[2], feat_strides[3],
                             feat_strides[4]};
    // the forward machinery with the roles of the buffers exchanged: the "features" are written
    return backproject_batch(grad_feat, fs, n_scenes, n_views, n_channels, hw, true, proj, 12ll * n_views,
                             points, grid, new_origin, true, n_points, const_cast<float *>(grad_volume),
                             gvol_strides, IVX_F32, reduce, nullptr, nullptr, (cudaStream_t)stream, true);
}

extern "C" int ivx_backproject_views(const float *feat, const int64_t *feat_strides, int n_views,
                                     int n_channels, int height, int width, const float *proj,
                                     const float *points, const ivx_grid *grid,
                                     const float *new_origin, int64_t n_points, float *volume,
                                     const int64_t *vol_strides, uint8_t *valid, void *stream) {
    int rc = validate_grid(points, grid, new_origin, n_points);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(feat && feat_strides && proj && volume && vol_strides && valid, "NULL argument");
    IVX_REQUIRE(n_views > 0 && n_channels > 0 && height > 0 && width > 0, "sizes must be positive");
    // every view is a one-view "scene": its sum is the gathered feature or 0
    const long long fs[5] = {feat_strides[0], 0, feat_strides[1], feat_strides[2], feat_strides[3]};
    const int32_t hw[2] = {height, width};
    return backproject_batch(feat, fs, n_views, 1, n_channels, hw, false, proj, 12, points, grid,
                             new_origin, false, n_points, volume, vol_strides, IVX_F32,
                             IVX_REDUCE_PARTIAL, valid, nullptr, (cudaStream_t)stream);
}

extern "C" int ivx_finalize_mean(const float *sum, const int32_t *count, int n_scenes, int n_channels,
                                 int64_t n_points, void *volume, const int64_t *vol_strides,
                                 int vol_dtype, uint8_t *valid, void *stream) {
    IVX_REQUIRE(sum && count && volume && vol_strides, "NULL argument");
    IVX_REQUIRE(n_scenes > 0 && n_points > 0 && n_points < (1ll << 31), "sizes out of range");
    IVX_REQUIRE(n_channels > 0 && n_channels % 4 == 0 && aligned16(sum),
                "partial sums must be channels-last with C %% 4 == 0 and 16-byte aligned");
    IVX_REQUIRE(vol_dtype == IVX_F32 || vol_dtype == IVX_BF16, "unknown vol_dtype %d", vol_dtype);
    const size_t esz = vol_dtype == IVX_BF16 ? 2 : 4;
    if (vol_strides[1] == 1)
        IVX_REQUIRE(vol_strides[0] % 4 == 0 && vol_strides[2] % 4 == 0 &&
                        reinterpret_cast<uintptr_t>(volume) % (4 * esz) == 0,
                    "channels-last volume must be aligned to 4 elements");
    else
        IVX_REQUIRE(vol_dtype == IVX_F32, "bf16 volumes must be channels-last");
    const long long rows = (long long)n_scenes * n_points;
    const int c4 = n_channels / 4;
    long long gx = (rows * c4 + 255) / 256;
    if (gx > kNumSMs * 32) gx = kNumSMs * 32;
    finalize_mean_kernel<<<(unsigned)gx, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(sum), count, rows, c4, (int)n_points, volume, vol_strides[0],
        vol_strides[1], vol_strides[2], vol_dtype, valid);
    IVX_CHECK_LAUNCH("finalize_mean launch");
    return IVX_OK;
}

extern "C" int ivx_repack_nchw_to_nhwc(const float *src, const int64_t *src_strides, int n_images,
                                       int n_channels, int height, int width, float *dst,
                                       void *stream) {
    IVX_REQUIRE(src && src_strides && dst, "NULL argument");
    IVX_REQUIRE(n_images > 0 && n_images <= 65535 && n_channels > 0 && height > 0 && width > 0,
                "sizes out of range");
    return launch_repack(src, 1, n_images, 0, src_strides[0], src_strides[1], src_strides[2],
                         src_strides[3], n_channels, height, width, dst, (cudaStream_t)stream);
}

static void max_hw(const int32_t *hw, int n_scenes, int &hmax, int &wmax) {
    hmax = wmax = 0;
    for (int b = 0; b < n_scenes; ++b) {
        hmax = hw[2 * b] > hmax ? hw[2 * b] : hmax;
        wmax = hw[2 * b + 1] > wmax ? hw[2 * b + 1] : wmax;
    }
}

extern "C" int64_t ivx_backproject_nchw_workspace_bytes(int n_scenes, int n_views, int n_channels,
                                                        const int32_t *hw) {
    if (n_scenes <= 0 || n_views <= 0 || n_channels <= 0 || hw == nullptr) return 0;
    int hmax, wmax;
    max_hw(hw, n_scenes, hmax, wmax);
    return (int64_t)n_scenes * n_views * hmax * wmax * n_channels * (int64_t)sizeof(float);
}

extern "C" int ivx_backproject_scenes_nchw(const float *feat, const int64_t *feat_strides,
                                           int n_scenes, int n_views, int n_channels,
                                           const int32_t *hw, const float *proj, const float *points,
                                           const ivx_grid *grid, const float *new_origin,
                                           int64_t n_points, void *volume, const int64_t *vol_strides,
                                           int vol_dtype, int reduce, uint8_t *valid, int32_t *count,
                                           void *workspace, int64_t workspace_bytes, void *stream) {
    IVX_REQUIRE(feat && feat_strides && hw && workspace, "NULL argument");
    IVX_REQUIRE(n_scenes > 0 && n_views > 0 && n_channels > 0, "sizes must be positive");
    if (n_channels % 4 != 0)
        return fail(IVX_ERR_UNSUPPORTED, "the repacking path needs C %% 4 == 0 (got %d)", n_channels);
    for (int b = 0; b < n_scenes; ++b)
        IVX_REQUIRE(hw[2 * b] > 0 && hw[2 * b + 1] > 0, "scene %d: height/width must be positive", b);
    const int64_t need = ivx_backproject_nchw_workspace_bytes(n_scenes, n_views, n_channels, hw);
    IVX_REQUIRE(workspace_bytes >= need, "workspace of %lld bytes is too small, %lld needed",
                (long long)workspace_bytes, (long long)need);
    IVX_REQUIRE(aligned16(workspace), "workspace must be 16-byte aligned");
    int hmax, wmax;
    max_hw(hw, n_scenes, hmax, wmax);
    float *ws = static_cast<float *>(workspace);
    int rc = launch_repack(feat, n_scenes, n_views, feat_strides[0], feat_strides[1], feat_strides[2],
                           feat_strides[3], feat_strides[4], n_channels, hmax, wmax, ws,
                           (cudaStream_t)stream);
    if (rc != IVX_OK) return rc;
    const int64_t row = (int64_t)wmax * n_channels, img = row * hmax;
    const int64_t ws_strides[5] = {img * n_views, img, 1, row, n_channels};
    return ivx_backproject_scenes(ws, ws_strides, n_scenes, n_views, n_channels, hw, proj, points,
                                  grid, new_origin, n_points, volume, vol_strides, vol_dtype, reduce,
                                  valid, count, stream);
}

// ---------------------------------------------------------------------------
// View-sharded finalize fused with its collective (multi-GPU).
//
// Every rank has written the fp32 partial sums / int32 counts of ITS views into
// a buffer that all ranks of the NVLink domain can address (symmetric memory:
// the same offset on every GPU).  Instead of an NCCL all-reduce followed by a
// separate finalize pass, one kernel reads the partials of all ranks and
// divides: through the NVSwitch multicast address with multimem.ld_reduce (the
// switch adds the G copies and returns one value: NVLS) when the buffer has
// one, else by loading each peer's copy over NVLink in rank order.  With the
// peer loads the result is bit-identical on every rank (same summation order
// everywhere); the switch's own order of the fp32 adds is not specified, so the
// multicast path only promises the 1e-5 bar.
// ---------------------------------------------------------------------------
namespace ivx {

constexpr int kMaxPeers = 16;

struct PeerArgs {
    const float4 *sum[kMaxPeers];
    const int32_t *count[kMaxPeers];
    const float4 *mc_sum;      // multicast alias of the sum buffers, or nullptr
    const int32_t *mc_count;   // multicast alias of the count buffers, or nullptr
    int world;
};

__device__ __forceinline__ float4 multimem_sum_f32x4(const float4 *p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}

__device__ __forceinline__ int multimem_sum_s32(const int32_t *p) {
    int v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Each thread keeps FU 16-byte units in flight.  On the multimem path the grid is at most two CTAs per SM
// (kPeerCtasPerSm): the kernel is then bound by NVLink, not by the SMs, and a small grid with more loads per
// thread moves the same bytes (measured at 8 GPUs: 0.50 ms per step either way).  It was meant to let a
// pipelined caller overlap this reduction with the neck kernels of the previous step; tools/overlap_probe.py
// shows that a neck layer and this kernel do not share SMs productively (layer 0.52 ms + reduction 2.85 ms
// = 4.18 ms together), so pipelining gains what the two streams' other gaps give, not the reduction time.  The view count of
// a voxel is fetched by one lane per row and warp and passed on with a shuffle (every lane of the row asked
// for it before: C / 4 identical requests per voxel through the switch).
constexpr int kPeerCtasPerSm = 2;

template <int FU>
__global__ void __launch_bounds__(256) finalize_mean_peers_kernel(
    const __grid_constant__ PeerArgs pa, long long n_rows, int c4, int N, void *vol, long long os_b,
    long long os_c, long long os_n, int dtype, uint8_t *__restrict__ valid, int32_t *__restrict__ count_out) {
    const long long total = n_rows * c4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 - threadIdx.x < total; i0 += stride * FU) {
        float4 s[FU];
        int cnt[FU];
        long long row[FU];
        int q[FU];
        bool ok[FU];
#pragma unroll
        for (int u = 0; u < FU; ++u) {
            const long long i = i0 + u * stride;
            ok[u] = i < total;
            row[u] = ok[u] ? i / c4 : 0;
            q[u] = ok[u] ? (int)(i - row[u] * c4) : 0;
            s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            cnt[u] = 0;
        }
        if (pa.mc_sum != nullptr) {
#pragma unroll
            for (int u = 0; u < FU; ++u)
                if (ok[u]) s[u] = multimem_sum_f32x4(pa.mc_sum + i0 + u * stride);
        } else {
            for (int r = 0; r < pa.world; ++r) {   // rank order: the same sum on every rank
                float4 t[FU];
#pragma unroll
                for (int u = 0; u < FU; ++u)
                    t[u] = ok[u] ? __ldcs(pa.sum[r] + i0 + u * stride) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < FU; ++u) s[u].x += t[u].x, s[u].y += t[u].y, s[u].z += t[u].z, s[u].w += t[u].w;
            }
        }
#pragma unroll
        for (int u = 0; u < FU; ++u) {
            // the first lane of this warp that works on the same row fetches the count for all of them
            const int src = max(lane - q[u], 0);
            int c = 0;
            if (ok[u] && src == lane) {
                if (pa.mc_sum != nullptr) c = multimem_sum_s32(pa.mc_count + row[u]);
                else
                    for (int r = 0; r < pa.world; ++r) c += __ldg(pa.count[r] + row[u]);
            }
            cnt[u] = __shfl_sync(0xffffffffu, c, src);
        }
#pragma unroll
        for (int u = 0; u < FU; ++u) {
            if (!ok[u]) continue;
            const float c = (float)cnt[u];
            float4 v = s[u];
            v.x = cnt[u] ? __fdiv_rn(v.x, c) : 0.f;
            v.y = cnt[u] ? __fdiv_rn(v.y, c) : 0.f;
            v.z = cnt[u] ? __fdiv_rn(v.z, c) : 0.f;
            v.w = cnt[u] ? __fdiv_rn(v.w, c) : 0.f;
            const long long b = row[u] / N, n = row[u] - b * N;
            const long long base = b * os_b + n * os_n;
            if (os_c == 1) {
                if (dtype == IVX_F32)
                    store4<float>(reinterpret_cast<float *>(vol) + base + q[u] * 4, v);
                else
                    store4<__nv_bfloat16>(reinterpret_cast<__nv_bfloat16 *>(vol) + base + q[u] * 4, v);
            } else {
                float *o = reinterpret_cast<float *>(vol) + base + (long long)q[u] * 4 * os_c;
                o[0] = v.x;
                o[os_c] = v.y;
                o[2 * os_c] = v.z;
                o[3 * os_c] = v.w;
            }
            if (q[u] == 0) {
                if (valid) valid[row[u]] = cnt[u] > 0;
                if (count_out) count_out[row[u]] = cnt[u];
            }
        }
    }
}

}  // namespace ivx

extern "C" int ivx_finalize_mean_peers(const float *const *peer_sums, const int32_t *const *peer_counts,
                                       int world, const float *mc_sum, const int32_t *mc_count,
                                       int n_scenes, int n_channels, int64_t n_points, void *volume,
                                       const int64_t *vol_strides, int vol_dtype, uint8_t *valid,
                                       int32_t *count_out, void *stream) {
    IVX_REQUIRE(volume && vol_strides, "NULL argument");
    IVX_REQUIRE(world >= 1 && world <= kMaxPeers, "world size must be in [1, %d]", kMaxPeers);
    IVX_REQUIRE((mc_sum != nullptr) == (mc_count != nullptr), "give both multicast pointers or neither");
    IVX_REQUIRE(mc_sum != nullptr || (peer_sums != nullptr && peer_counts != nullptr),
                "need the multicast pointers or the per-rank pointers");
    IVX_REQUIRE(n_scenes > 0 && n_points > 0 && n_points < (1ll << 31), "sizes out of range");
    IVX_REQUIRE(n_channels > 0 && n_channels % 4 == 0, "partial sums must be channels-last with C %% 4 == 0");
    IVX_REQUIRE(vol_dtype == IVX_F32 || vol_dtype == IVX_BF16, "unknown vol_dtype %d", vol_dtype);
    IVX_REQUIRE(mc_sum == nullptr || aligned16(mc_sum), "mc_sum must be 16-byte aligned (multimem.ld_reduce.v4.f32)");
    const size_t esz = vol_dtype == IVX_BF16 ? 2 : 4;
    if (vol_strides[1] == 1)
        IVX_REQUIRE(vol_strides[0] % 4 == 0 && vol_strides[2] % 4 == 0 &&
                        reinterpret_cast<uintptr_t>(volume) % (4 * esz) == 0,
                    "channels-last volume must be aligned to 4 elements");
    else
        IVX_REQUIRE(vol_dtype == IVX_F32, "bf16 volumes must be channels-last");
    PeerArgs pa;
    pa.world = world;
    pa.mc_sum = reinterpret_cast<const float4 *>(mc_sum);
    pa.mc_count = mc_count;
    for (int r = 0; r < kMaxPeers; ++r) {
        pa.sum[r] = (peer_sums && r < world) ? reinterpret_cast<const float4 *>(peer_sums[r]) : nullptr;
        pa.count[r] = (peer_counts && r < world) ? peer_counts[r] : nullptr;
        if (mc_sum == nullptr && r < world)
            IVX_REQUIRE(pa.sum[r] && pa.count[r] && aligned16(pa.sum[r]), "rank %d: bad partial pointers", r);
    }
    const long long rows = (long long)n_scenes * n_points;
    const int c4 = n_channels / 4;
    // through the switch (multimem) a rank takes in only its own slice, already reduced: two CTAs per SM with
    // four units per thread keep the links busy.  Peer loads read every rank's copy, the local one at HBM
    // speed: they keep one unit per thread and the big grid (measured at 2 GPUs: 0.53 ms against 0.61 ms with
    // four units per thread, 0.72 ms with two CTAs per SM)
    static const int force_small = [] {           // experiments: the small grid on the peer path too
        const char *e = getenv("IVX_FINALIZE_SMALL_GRID");
        return e ? atoi(e) : 0;
    }();
    if (mc_sum != nullptr) {
        long long gx = (rows * c4 + 256 * 4 - 1) / (256 * 4);
        if (gx > kNumSMs * kPeerCtasPerSm) gx = kNumSMs * kPeerCtasPerSm;
        finalize_mean_peers_kernel<4><<<(unsigned)gx, 256, 0, (cudaStream_t)stream>>>(
            pa, rows, c4, (int)n_points, volume, vol_strides[0], vol_strides[1], vol_strides[2], vol_dtype, valid,
            count_out);
    } else {
        long long gx = (rows * c4 + 255) / 256;
        if (gx > kNumSMs * 32) gx = kNumSMs * 32;
        finalize_mean_peers_kernel<1><<<(unsigned)gx, 256, 0, (cudaStream_t)stream>>>(
            pa, rows, c4, (int)n_points, volume, vol_strides[0], vol_strides[1], vol_strides[2], vol_dtype, valid,
            count_out);
    }
    IVX_CHECK_LAUNCH("finalize_mean_peers launch");
    return IVX_OK;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of ivx_backproject_scenes_nchw.
//
// Some users of this library run where device memory cannot be requested at
// call time: a plugin inside a host application that has already handed all
// GPU memory to its own pool (so cudaMalloc fails although the pool has room),
// or a serving / real-time thread that must not take the driver's allocation
// lock or suffer the device-wide synchronisation cudaMalloc and cudaFree imply.
// For them the channels-last copy of the maps lives in the fixed array below
// instead of a caller-provided workspace.  Every GPU that loads the library
// holds these IVX_NOALLOC_SCRATCH_BYTES until the process exits.
//
// `g_noalloc_scratch` is a device symbol: in host code its name is only a
// stand-in, so the device address is looked up with cudaGetSymbolAddress,
// once per device and with that device current, and cached.
// ---------------------------------------------------------------------------
static __device__ __align__(256) float g_noalloc_scratch[IVX_NOALLOC_SCRATCH_BYTES / sizeof(float)];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
float *g_noalloc_address[kMaxDevices] = {nullptr};

// Device address of the scratch array on the current device (set-up on first use).
int noalloc_scratch_address(float **out) {
    int device = -1;
    int rc = check_cuda(cudaGetDevice(&device), "cudaGetDevice");
    if (rc != IVX_OK) return rc;
    if (device < 0 || device >= kMaxDevices)
        return fail(IVX_ERR_UNSUPPORTED, "device ordinal %d is beyond the %d this build tracks", device,
                    kMaxDevices);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_address[device] == nullptr) {
        void *address = nullptr;
        rc = check_cuda(cudaGetSymbolAddress(&address, g_noalloc_scratch),
                        "cudaGetSymbolAddress(built-in scratch)");
        if (rc != IVX_OK) return rc;
        g_noalloc_address[device] = static_cast<float *>(address);
    }
    *out = g_noalloc_address[device];
    return IVX_OK;
}
}  // namespace

extern "C" int ivx_backproject_scenes_nchw_noalloc(
    const float *feat, const int64_t *feat_strides, int n_scenes, int n_views, int n_channels,
    const int32_t *hw, const float *proj, const float *points, const ivx_grid *grid,
    const float *new_origin, int64_t n_points, void *volume, const int64_t *vol_strides, int vol_dtype,
    int reduce, uint8_t *valid, int32_t *count, void *stream) {
    IVX_REQUIRE(hw != nullptr && n_scenes > 0 && n_views > 0 && n_channels > 0,
                "hw must be given and sizes must be positive");
    const int64_t need = ivx_backproject_nchw_workspace_bytes(n_scenes, n_views, n_channels, hw);
    if (need > (int64_t)IVX_NOALLOC_SCRATCH_BYTES)
        return fail(IVX_ERR_UNSUPPORTED,
                    "these maps need %lld bytes of scratch but the built-in scratch holds %lld: split "
                    "the batch, or call ivx_backproject_scenes_nchw with a workspace of your own",
                    (long long)need, (long long)IVX_NOALLOC_SCRATCH_BYTES);
    float *scratch = nullptr;
    int rc = noalloc_scratch_address(&scratch);
    if (rc != IVX_OK) return rc;
    // The transpose overwrites every element the gather will read, so nothing
    // in the shared array has to be cleared between calls.
    return ivx_backproject_scenes_nchw(feat, feat_strides, n_scenes, n_views, n_channels, hw, proj,
                                       points, grid, new_origin, n_points, volume, vol_strides,
                                       vol_dtype, reduce, valid, count, scratch,
                                       (int64_t)IVX_NOALLOC_SCRATCH_BYTES, stream);
}
