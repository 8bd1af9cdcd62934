// Weight gradient of a Conv3d layer on tcgen05 (tf32 math), training mode.
//
//   dW[tap][co][ci] = sum over output voxels o of  dz[o][co] * x[o * stride - pad + tap][ci]
//
// is, per tap, a GEMM whose reduction runs over VOXELS.  tcgen05 wants both operands K-major in
// shared memory, i.e. voxels contiguous, so the caller hands in "tiled channel-major" copies
// (ivx_channel_major_ndhwc writes them): with (D0, D1, D2) a permutation of the layer's (X, Y, Z),
//   dz_t [batch][D2][D1][D0 / 32][c_out][32]
//   x_t  [k0][batch][I2 + 2 pad2][I1][D0 / 32][c_in][32]
// i.e. runs of 32 voxels along D0, and all channels of a run next to each other: the operand tile of
// a k-step ([channel rows][32 voxels] = [rows][128 bytes], the K-major swizzled layout the forward
// kernel feeds to the MMA) is ONE contiguous block of 8..32 KB in memory.  (A plain [C][voxels]
// layout works too but reads one 128-byte row from each of c_in + c_out planes per k-step: 27 %
// tensor-pipe activity and 2.3 TB/s of scattered DRAM reads in ncu.)
// Two TMA rules found by measurement shape x_t: a swizzled box cannot start at an innermost coordinate
// that is not 16-byte aligned (an x box starting one voxel to the left is an illegal instruction) and
// the innermost dimension cannot be strided, so the tap shift and the layer stride along D0 are baked
// into one copy of x per tap position along that axis,
//   x_t[j][..][o] = x[..][o * stride0 - pad0 + j]      (0 outside the volume),
// indexed by OUTPUT coordinate.  The shift along D1 is a TMA coordinate (zero fill outside); the planes
// along D2 are merged with the batch into one TMA dimension, so x_t carries pad2 zero planes on
// either side of every volume instead.
// One CTA owns (a group of taps, 128 output channels, up to 256 input channels, a slice of the voxel
// range): warp 0 issues the loads, warp 1 the MMAs (M = 128 rows of c_out, N = c_in tile, K = 8 per
// instruction, one fp32 accumulator per tap in TMEM), warps 2..5 add the accumulators into dW with
// 128-bit vector reductions when the slice is done (dW must be zero-filled by the caller).
// The kernel is bound by the L2 -> shared-memory traffic (every tap re-reads both operands), so the
// taps along axis D1 share one grad_z tile per k-step when their accumulators fit the 512 TMEM columns.
#include "ivx_tc_ptx.cuh"

namespace ivx {

constexpr int WG_THREADS = 192;      // warp 0: TMA, warp 1: MMA, warps 2..5: epilogue
constexpr int WG_A_BYTES = 128 * 128;

struct WgArgs {
    int c_in, c_out, n_tile;         // n_tile: input channels per CTA tile
    int ci_tiles, co_tiles;
    int kernel[3], stride[3], pad[3];   // in the permuted axis order (index 0 = innermost)
    int orig_axis[3];                // permuted axis i is axis orig_axis[i] of (X, Y, Z)
    int orig_kernel[3];              // kernel sizes along X, Y, Z
    int box[3], chunks[3], batch;    // voxels per k-step along each permuted axis, k-steps per axis
    long long k_steps;               // batch * chunks[0] * chunks[1] * chunks[2]
    // 3xTF32 (fp32-grade gradients, see ivx_conv3d_tc_halo.cu): both operands carry 2 C channels (hi | lo) and
    // the k-steps run three passes over the voxels: (x_lo, dz_hi), (x_hi, dz_lo), (x_hi, dz_hi)
    int x3;
    long long k_total;               // k_steps, or 3 * k_steps (x3)
    int k_splits;
    long long steps_per_split;
    int stages;
    int a_rows;                      // rows of the grad_z box: min(c_out, 128)
    int dz_planes, x_planes;         // planes per volume along D2 in dz_t, and in x_t (with its zero planes)
    int group;                       // taps along D1 handled by one CTA (kernel[1], or 1)
    int debug;                       // IVX_WG_DEBUG bits: 1 no TMA loads, 2 no MMAs, 4 no epilogue reads
    float *dw;                       // [taps (kx,ky,kz order)][c_out][c_in]
    TileDiv div_tap, div_ci, div_co, div_c0, div_c1, div_c2;
};

__global__ void __launch_bounds__(WG_THREADS, 1)
conv3d_wgrad_tc_kernel(const __grid_constant__ CUtensorMap map_dz, const __grid_constant__ CUtensorMap map_x,
                       const __grid_constant__ WgArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int b_bytes = a.n_tile * 128;
    const int stage_bytes = WG_A_BYTES + a.group * b_bytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)a.stages * stage_bytes);
    uint64_t *full_bar = bars, *empty_bar = bars + a.stages, *acc_full = bars + 2 * a.stages;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int acc_cols = a.n_tile * a.group;
    const int tmem_cols = acc_cols <= 32 ? 32 : (acc_cols <= 64 ? 64 : (acc_cols <= 128 ? 128 : (acc_cols <= 256 ? 256 : 512)));

    if (threadIdx.x == 0) {
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(smem_u32(full_bar + s), 1);
            mbar_init(smem_u32(empty_bar + s), 1);
        }
        mbar_init(smem_u32(acc_full), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_dz) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), tmem_cols);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    // work item of this CTA: (tap group, c_in tile, c_out tile, k-slice), tap group fastest: the CTAs that
    // run at the same time work on the same voxels, so all but the first find grad_z and x in L2
    int item = blockIdx.x, tap_g, ci_t, co_t;               // tap group: (t0, t2) and, if group == 1, t1
    item = a.div_tap.divmod(item, tap_g);
    item = a.div_ci.divmod(item, ci_t);
    const int split = a.div_co.divmod(item, co_t);
    const int t0 = tap_g % a.kernel[0];
    const int t1_lo = a.group > 1 ? 0 : (tap_g / a.kernel[0]) % a.kernel[1];
    const int t2 = a.group > 1 ? tap_g / a.kernel[0] : tap_g / (a.kernel[0] * a.kernel[1]);
    const long long k_lo = (long long)split * a.steps_per_split;
    const long long k_hi = k_lo + a.steps_per_split < a.k_total ? k_lo + a.steps_per_split : a.k_total;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (long long ks0 = k_lo; ks0 < k_hi; ++ks0) {
                const int pass = a.x3 ? (int)(ks0 / a.k_steps) : 2;
                const long long ks = ks0 - (a.x3 ? pass * a.k_steps : 0);
                const int dz_ch = (a.x3 && pass == 1) ? a.c_out : 0;      // lo half of grad_z
                const int x_ch = (a.x3 && pass == 0) ? a.c_in : 0;        // lo half of x
                int r0, r1, r2;
                int t = a.div_c0.divmod((int)ks, r0);
                t = a.div_c1.divmod(t, r1);
                const int b = a.div_c2.divmod(t, r2);
                const int o1 = r1, o2 = r2;      // (r0 counts runs of 32 voxels along D0)
                mbar_wait(smem_u32(empty_bar + stage), phase ^ 1);
                const uint32_t bar = smem_u32(full_bar + stage);
                const uint32_t sa = smem_u32(smem + (size_t)stage * stage_bytes);
                if (a.debug & 1) {
                    mbar_arrive(bar);
                } else {
                    // a box always arrives in full, whatever part of it lies outside the tensor
                    mbar_expect_tx(bar, (uint32_t)(((a.debug & 8) ? 0 : a.a_rows * 128) +
                                                   ((a.debug & 16) ? 0 : a.group * b_bytes)));
                    if (!(a.debug & 8)) tma_load_5d(sa, &map_dz, bar, 0, dz_ch + co_t * 128, r0, o1, b * a.dz_planes + o2);
                    if (!(a.debug & 16))
                        for (int g = 0; g < a.group; ++g)
                            tma_load_5d(sa + WG_A_BYTES + g * b_bytes, &map_x, bar, 0, x_ch + ci_t * a.n_tile, r0,
                                        o1 * a.stride[1] - a.pad[1] + t1_lo + g,
                                        (t0 * a.batch + b) * a.x_planes + o2 * a.stride[2] + t2);
                }
                if (++stage == (uint32_t)a.stages) stage = 0, phase ^= 1;
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // the x tiles of a tap group lie one after the other in the stage: while they fit N <= 256 they
            // are ONE B operand (rows = tap x input channel), i.e. a third of the MMA instructions
            const bool fused = a.group * a.n_tile <= 256;
            const int n_mma = fused ? a.group * a.n_tile : a.n_tile, n_groups = fused ? 1 : a.group;
            const uint32_t idesc = umma_idesc_tf32(n_mma);
            uint32_t stage = 0, phase = 0;
            for (long long ks = k_lo; ks < k_hi; ++ks) {
                mbar_wait(smem_u32(full_bar + stage), phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = smem_u32(smem + (size_t)stage * stage_bytes);
                const uint64_t da = umma_desc_sw128(sa);
                for (int g = 0; g < n_groups; ++g) {
                    const uint64_t db = umma_desc_sw128(sa + WG_A_BYTES + g * b_bytes);
#pragma unroll
                    for (int k = 0; k < 4; ++k)   // 8 voxels (32 bytes) per MMA inside the 128-byte row
                        if (!(a.debug & 2))
                            umma_tf32(tmem_base + g * a.n_tile, da + 2 * k, db + 2 * k, idesc, (ks > k_lo) || k != 0);
                }
                umma_commit(smem_u32(empty_bar + stage));
                if (++stage == (uint32_t)a.stages) stage = 0, phase ^= 1;
            }
            umma_commit(smem_u32(acc_full));
        }
    } else if (k_hi > k_lo) {
        // epilogue: lane = output channel of the tile, columns = input channels
        const int q = warp & 3;
        const int co = co_t * 128 + q * 32 + lane;
        mbar_wait(smem_u32(acc_full), 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int g = 0; g < a.group; ++g) {
            // tap index in the (kx, ky, kz) order of the weight tensor
            int k3[3];
            k3[a.orig_axis[0]] = t0, k3[a.orig_axis[1]] = t1_lo + g, k3[a.orig_axis[2]] = t2;
            const int tap = (k3[0] * a.orig_kernel[1] + k3[1]) * a.orig_kernel[2] + k3[2];
            float *dst = a.dw + ((long long)tap * a.c_out + co) * a.c_in + ci_t * a.n_tile;
            for (int c = 0; c < ((a.debug & 4) ? 0 : a.n_tile); c += 32) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + g * a.n_tile + c, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (co < a.c_out) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        if (c + j < a.n_tile)
                            atomicAdd(reinterpret_cast<float4 *>(dst + c + j),
                                      make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                  __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
                }
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace ivx

using namespace ivx;

// Plan shared by the entry point and the layout query: which axis goes innermost.  A k-step is a run
// of 32 voxels along that axis (one full 128-byte swizzle row per channel: boxes with a shorter inner
// dimension are refused by the TMA in this mode -- measured -- so a short axis is padded by the
// zero fill instead).  The axis wasting the fewest voxels wins, ties go to the tensor's own fastest
// axis (Z).
static int plan_wgrad(const ivx_conv3d_params *p, int axis[3], int box[3]) {
    int inner = -1;
    double best = -1.0;
    for (int d = 2; d >= 0; --d) {
        const int n = p->out_dims[d];
        const double eff = (double)n / (double)(((n + 31) / 32) * 32);
        if (eff > best + 1e-9) best = eff, inner = d;
    }
    if (inner < 0) return -1;
    axis[0] = inner;
    int n = 1;
    for (int d = 2; d >= 0; --d)
        if (d != inner) axis[n++] = d;
    box[0] = 32, box[1] = 1, box[2] = 1;
    return 0;
}

extern "C" int ivx_conv3d_wgrad_tf32_axes(const ivx_conv3d_params *p, int32_t *axes /* host[3] */) {
    IVX_REQUIRE(p && axes, "NULL argument");
    if (p->transposed || p->c_in % 32 != 0 || p->c_out % 32 != 0)
        return fail(IVX_ERR_UNSUPPORTED, "tensor-core weight gradient: plain convolutions with c_in, c_out multiples of 32");
    int axis[3], box[3];
    if (plan_wgrad(p, axis, box) != 0)
        return fail(IVX_ERR_UNSUPPORTED, "tensor-core weight gradient: no usable axis");
    for (int i = 0; i < 3; ++i) axes[i] = axis[i];
    return IVX_OK;
}

static int launch_wgrad(const ivx_conv3d_params *p, const float *x_t, const float *dz_t, float *grad_w, void *stream,
                        bool x3);

extern "C" int ivx_conv3d_wgrad_tf32_tc(const ivx_conv3d_params *p, const float *x_t, const float *dz_t,
                                        float *grad_w, void *stream) {
    return launch_wgrad(p, x_t, dz_t, grad_w, stream, false);
}

extern "C" int ivx_conv3d_wgrad_f32x3_tc(const ivx_conv3d_params *p, const float *x_t_split, const float *dz_t_split,
                                         float *grad_w, void *stream) {
    return launch_wgrad(p, x_t_split, dz_t_split, grad_w, stream, true);
}

static int launch_wgrad(const ivx_conv3d_params *p, const float *x_t, const float *dz_t, float *grad_w, void *stream,
                        bool x3) {
    IVX_REQUIRE(p && x_t && dz_t && grad_w, "NULL argument");
    int32_t axes[3];
    int rc = ivx_conv3d_wgrad_tf32_axes(p, axes);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x_t) | reinterpret_cast<uintptr_t>(dz_t) |
                 reinterpret_cast<uintptr_t>(grad_w)) % 16 == 0,
                "tensors must be 16-byte aligned");
    WgArgs a;
    int axis[3];
    plan_wgrad(p, axis, a.box);
    a.c_in = p->c_in, a.c_out = p->c_out;
    a.n_tile = p->c_in % 256 == 0 ? 256 : (p->c_in % 128 == 0 ? 128 : (p->c_in % 64 == 0 ? 64 : 32));
    // taps along D1 share the grad_z tile when their accumulators fit TMEM (<= 128 input channels each)
    const int k1 = p->kernel[axis[1]];
    a.group = 1;
    if (k1 > 1 && k1 <= 4) {
        if (a.n_tile > 128) a.n_tile = 128;
        a.group = k1;
    }
    a.ci_tiles = p->c_in / a.n_tile;
    a.co_tiles = (p->c_out + 127) / 128;
    a.batch = p->batch;
    long long k_steps = p->batch;
    for (int i = 0; i < 3; ++i) {
        a.kernel[i] = p->kernel[axis[i]], a.stride[i] = p->stride[axis[i]], a.pad[i] = p->pad[axis[i]];
        a.orig_axis[i] = axis[i], a.orig_kernel[i] = p->kernel[i];
        a.chunks[i] = (p->out_dims[axis[i]] + a.box[i] - 1) / a.box[i];
        k_steps *= a.chunks[i];
    }
    IVX_REQUIRE(k_steps < (1ll << 31) / 3, "too many voxels");
    a.k_steps = k_steps;
    a.x3 = x3 ? 1 : 0;
    a.k_total = x3 ? 3 * k_steps : k_steps;
    const long long k_all = a.k_total;
    const int ch_mul = x3 ? 2 : 1;          // channels per voxel in the operand copies: (hi | lo)
    const int n_taps = p->kernel[0] * p->kernel[1] * p->kernel[2];
    const long long tiles = (long long)(n_taps / a.group) * a.co_tiles * a.ci_tiles;
    // enough k-slices to fill the GPU a few times over, at least 64 k-steps each
    long long splits = (4ll * kNumSMs + tiles - 1) / tiles;
    if (splits > k_all / 64) splits = k_all / 64;
    if (splits < 1) splits = 1;
    a.steps_per_split = (k_all + splits - 1) / splits;
    a.k_splits = (int)((k_all + a.steps_per_split - 1) / a.steps_per_split);
    IVX_REQUIRE(tiles * a.k_splits < (1ll << 31), "too many work items");
    a.dw = grad_w;
    a.a_rows = p->c_out < 128 ? p->c_out : 128;
    {
        const char *e = getenv("IVX_WG_DEBUG");
        a.debug = e ? atoi(e) : 0;
    }
    a.div_tap = make_tilediv(n_taps / a.group), a.div_ci = make_tilediv(a.ci_tiles), a.div_co = make_tilediv(a.co_tiles);
    a.div_c0 = make_tilediv(a.chunks[0]), a.div_c1 = make_tilediv(a.chunks[1]), a.div_c2 = make_tilediv(a.chunks[2]);
    const int stage_bytes = WG_A_BYTES + a.group * a.n_tile * 128;
    a.stages = (227 * 1024 - 2048) / stage_bytes;
    if (a.stages > 8) a.stages = 8;
    const size_t smem = (size_t)a.stages * stage_bytes + 2048;

    a.dz_planes = p->out_dims[axis[2]];
    a.x_planes = p->in_dims[axis[2]] + 2 * p->pad[axis[2]];
    const CUtensorMapDataType dt = (a.debug & 32) ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32;
    const cuuint64_t n_runs = (cuuint64_t)a.chunks[0];
    const cuuint32_t es[5] = {1, 1, 1, 1, 1};
    alignas(64) CUtensorMap map_dz, map_x;
    {   // (voxel in run, channel, run along D0, D1, plane x batch)
        const cuuint64_t dims[5] = {32, (cuuint64_t)p->c_out * ch_mul, n_runs, (cuuint64_t)p->out_dims[axis[1]],
                                    (cuuint64_t)a.dz_planes * p->batch};
        cuuint64_t st[4] = {128, (cuuint64_t)p->c_out * ch_mul * 128, 0, 0};
        st[2] = st[1] * n_runs, st[3] = st[2] * p->out_dims[axis[1]];
        const cuuint32_t box[5] = {32, (cuuint32_t)a.a_rows, 1, 1, 1};
        rc = encode_map_sw128(&map_dz, dt, 5, const_cast<float *>(dz_t), dims, st, box, es, "grad_z (tiled channel-major)");
        if (rc != IVX_OK) return rc;
    }
    {   // the same for x, with one copy per tap position along D0 and the zero planes along D2
        const cuuint64_t dims[5] = {32, (cuuint64_t)p->c_in * ch_mul, n_runs, (cuuint64_t)p->in_dims[axis[1]],
                                    (cuuint64_t)a.x_planes * p->batch * p->kernel[axis[0]]};
        cuuint64_t st[4] = {128, (cuuint64_t)p->c_in * ch_mul * 128, 0, 0};
        st[2] = st[1] * n_runs, st[3] = st[2] * p->in_dims[axis[1]];
        const cuuint32_t box[5] = {32, (cuuint32_t)a.n_tile, 1, 1, 1};
        rc = encode_map_sw128(&map_x, dt, 5, const_cast<float *>(x_t), dims, st, box, es, "x (tiled channel-major)");
        if (rc != IVX_OK) return rc;
    }
    rc = check_cuda(cudaFuncSetAttribute(conv3d_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem),
                    "cudaFuncSetAttribute");
    if (rc != IVX_OK) return rc;
    conv3d_wgrad_tc_kernel<<<(unsigned)(tiles * a.k_splits), WG_THREADS, smem, (cudaStream_t)stream>>>(map_dz, map_x, a);
    IVX_CHECK_LAUNCH("conv3d wgrad (tensor cores) launch");
    return IVX_OK;
}

// ---------------------------------------------------------------------------------------------
// Tiled channel-major copies for the weight-gradient kernel: x [batch][X][Y][Z][C] (channels-last) ->
//   out [n_shifts][batch][D2 + 2 halo][D1][runs][C][32]
//   out[j][b][i2 + halo][i1][o / 32][c][o % 32] = x[b][..][c] at position o * step + shift0 + j along D0
// (zero where that position is outside the volume, for o >= out_len, and in the halo planes).
// 32 x 32 tiles of (position along D0, channel) go through shared memory: reads are contiguous
// along channels, a tile is written as one contiguous 4 KB block; the source rows are staged once
// for all shifted copies.
// ---------------------------------------------------------------------------------------------
namespace ivx {

struct CmArgs {
    const float *x;
    float *out;
    int C, n_shifts, shift0, step;
    int d0, d1, d2;                  // input extents along the permuted axes
    long long s0, s1, s2, sb;        // element strides of those axes (and of the batch) in x
    int out_len, halo, batch;
    int o_tiles;                     // runs of 32 output positions
};

constexpr int CM_ROWS = 96;          // source positions one block stages: 31 * step + n_shifts of them

__global__ void __launch_bounds__(256) channel_major_kernel(const CmArgs a) {
    __shared__ float tile[CM_ROWS][33];
    const int planes = a.d2 + 2 * a.halo;
    const long long blk = blockIdx.x;
    const int ot = (int)(blk % a.o_tiles);
    const long long line = blk / a.o_tiles;
    const int i1 = (int)(line % a.d1);
    const long long t = line / a.d1;
    const int i2h = (int)(t % planes), b = (int)(t / planes);
    const int i2 = i2h - a.halo;
    const bool real_plane = i2 >= 0 && i2 < a.d2;
    const int c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8 threads
    if (real_plane) {
        // every source position the 32 outputs of this tile need, for all shifts, read once:
        // rows = positions, columns = channels (coalesced reads)
        const float *src = a.x + b * a.sb + i2 * a.s2 + i1 * a.s1;
        const int i_lo = ot * 32 * a.step + a.shift0;
        const int n_rows = 31 * a.step + a.n_shifts;
        for (int r = ty; r < n_rows; r += 8) {
            const int i = i_lo + r;
            float v = 0.f;
            if (i >= 0 && i < a.d0 && c0 + tx < a.C) v = __ldg(src + i * a.s0 + c0 + tx);
            tile[r][tx] = v;
        }
    }
    __syncthreads();
    // rows = channels, columns = output positions: 32 x 128 contiguous bytes per tile and copy
    const int o = ot * 32 + tx;
    for (int j = 0; j < a.n_shifts; ++j) {
        float *dst = a.out + ((((((long long)j * a.batch + b) * planes + i2h) * a.d1 + i1) * a.o_tiles + ot) * a.C) * 32;
#pragma unroll
        for (int r = ty; r < 32; r += 8) {
            const int c = c0 + r;
            if (c < a.C) dst[c * 32 + tx] = (real_plane && o < a.out_len) ? tile[tx * a.step + j][r] : 0.f;
        }
    }
}

}  // namespace ivx

extern "C" int ivx_channel_major_ndhwc(const float *x, int32_t batch, const int32_t *dims, int32_t n_channels,
                                       const int32_t *axes, int32_t n_shifts, int32_t shift0, int32_t step,
                                       int32_t out_len, int32_t halo, float *out, void *stream) {
    IVX_REQUIRE(x && dims && axes && out, "NULL argument");
    IVX_REQUIRE(batch > 0 && n_channels > 0 && n_shifts > 0 && step > 0 && out_len > 0 && halo >= 0, "bad sizes");
    for (int i = 0; i < 3; ++i) IVX_REQUIRE(dims[i] > 0 && axes[i] >= 0 && axes[i] < 3, "bad dims / axes");
    IVX_REQUIRE(axes[0] != axes[1] && axes[0] != axes[2] && axes[1] != axes[2], "axes must be a permutation");
    IVX_REQUIRE(31 * step + n_shifts <= ivx::CM_ROWS, "step / n_shifts too large (31 * step + n_shifts <= 96)");
    const long long st[3] = {(long long)dims[1] * dims[2] * n_channels, (long long)dims[2] * n_channels, n_channels};
    ivx::CmArgs a;
    a.x = x, a.out = out, a.C = n_channels, a.n_shifts = n_shifts, a.shift0 = shift0, a.step = step;
    a.d0 = dims[axes[0]], a.d1 = dims[axes[1]], a.d2 = dims[axes[2]];
    a.s0 = st[axes[0]], a.s1 = st[axes[1]], a.s2 = st[axes[2]], a.sb = st[0] * dims[0];
    a.out_len = out_len, a.halo = halo, a.batch = batch;
    a.o_tiles = (out_len + 31) / 32;
    const long long blocks = (long long)batch * (a.d2 + 2 * halo) * a.d1 * a.o_tiles;
    IVX_REQUIRE(blocks < (1ll << 31), "too many voxels");
    dim3 grid((unsigned)blocks, (unsigned)((n_channels + 31) / 32));
    ivx::channel_major_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    IVX_CHECK_LAUNCH("channel-major copy launch");
    return IVX_OK;
}
