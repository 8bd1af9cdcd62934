// Tensor-core Conv3d for the 3D necks: implicit GEMM on tcgen05 / TMEM fed by TMA.
//
// GEMM view of one layer (channels-last bf16 activations):
//   M = output voxels, N = c_out, K = kX*kY*kZ * c_in
//   A[m][(tap, ci)] = x[voxel(m) * stride - pad + tap][ci]   (0 outside the volume)
//   B[(tap, ci)][n] = w[n][tap][ci]
// One CTA owns a 128-row tile = a box of (bx, by, bz) output voxels and up to
// 256 output channels.  For every (tap, 64-channel chunk) k-step the producer
// thread issues
//   * one 5-D TMA box load of the activations with the box start shifted by
//     the tap (and the element strides of the tensor map equal to the conv
//     stride): zero padding is the TMA's out-of-bounds fill, and the box lands
//     in shared memory as 128 rows x 128 bytes, 128B-swizzled -- exactly the
//     K-major operand layout tcgen05.mma expects;
//   * one 2-D TMA load of the [n][64] weight slice, same layout.
// A single thread issues tcgen05.mma (M=128, N=c_out tile, K=16, bf16 in, fp32
// accumulate in TMEM); tcgen05.commit releases the smem stage and, after the
// last k-step, hands the accumulator to the eight epilogue warps (two per TMEM
// lane quarter, 32 of a slab's 64 channels each), which read it with
// tcgen05.ld, apply folded BN scale/shift, residual, ReLU, convert to bf16,
// stage 64-channel slabs in swizzled shared memory and TMA-store them (the
// store clips partial tiles).  Where shared memory allows, the residual tile is
// TMA-loaded into a slab n_slabs - 1 chunks ahead (3 to 8 slabs rotate), so
// adding it is a read-modify-write of shared memory instead of an exposed
// global load: layers with few k-steps (1x1x1, transposed) are bound by this
// epilogue and by the latency of that load.
// Two accumulators in TMEM (2 x N columns) let the epilogue of tile i overlap
// the main loop of tile i+1; CTAs are persistent over tiles.
#include <type_traits>

#include "ivx_tc_ptx.cuh"

namespace ivx {

constexpr int TC_THREADS = 320;      // warp 0: TMA, warp 1: MMA, warps 2..9: epilogue
constexpr int TC_EPI_WARPS = 8;
constexpr int TC_ROWS = 128;         // UMMA M
constexpr int TC_KCHUNK = 64;        // bf16 channels per k-step = one 128B swizzle row
constexpr int TC_A_BYTES = TC_ROWS * 128;

struct TcArgs {
    int c_in, c_out;
    int n_tile;                 // output channels per CTA tile (<= 256, multiple of 16)
    int n_splits;               // c_out / n_tile
    int taps[3], stride[3], pad[3];
    int out_dims[3], batch;
    int box[3];                 // output voxels per tile along x, y, z
    int tiles[3];               // tiles along x, y, z
    long long n_tiles;          // batch * tiles * n_splits * k_splits (< 2^31)
    TileDiv div_k, div_n, div_z, div_y, div_x;   // tile index -> (split, n-tile, z, y, x, batch)
    int stages;
    int relu, residual_after_act;
    float post_scale;
    const float *scale, *shift;
    const void *res;            // residual, in the layer's storage type
    int n_slabs;                // output staging slots (64 channels each): 2 (1 if short of shared memory),
                                // or 3..8 when the residual comes in by TMA
    int res_tma;                // residual tiles are TMA-loaded into the slabs (needs n_slabs >= 3)
    // transposed layers (ConvTranspose3d, kernel == stride): rows are INPUT voxels, the GEMM
    // columns are (tap, c_out); every n-tile lies inside one tap and is stored with that
    // tap's offset through an output map whose element strides are the layer stride
    int transposed;
    int up[3];                  // upsampling factor = kernel = stride of the transposed layer
    // split-K for layers with too few tiles to fill the GPU: the k-steps of one output tile are
    // shared by k_splits CTAs, which add their fp32 partial accumulators into `ws` with vector
    // reductions; the CTA that arrives last (per-tile counter) reads the sums back, runs the
    // fused epilogue and leaves workspace and counter zeroed for the next launch
    int k_splits;
    int steps_per_split;
    float *ws;                  // [out tiles][128 rows][n_tile] fp32, zero on entry
    unsigned *counters;         // [out tiles], zero on entry
    // 3xTF32 (fp32-grade results, see ivx_conv3d_tc_halo.cu): activations are (hi | lo) channels, weight
    // rows (hi | lo) x taps x c_in; the k-steps run three passes (x_lo w_hi, x_hi w_lo, x_hi w_hi)
    int x3;
    int drain_steps;            // x3: k-steps (of 4 MMAs) per accumulation period (x3_accumulate_periods)
};

struct TileCoord {
    int b, x0, y0, z0, n0;
};

__device__ __forceinline__ TileCoord decode_tile(const TcArgs &a, long long tile) {
    TileCoord c;
    int t = a.div_k.div((int)tile);   // the split index is the fastest: CTAs of one tile run together
    int r;
    t = a.div_n.divmod(t, r);
    c.n0 = r * a.n_tile;
    t = a.div_z.divmod(t, r);
    c.z0 = r * a.box[2];
    t = a.div_y.divmod(t, r);
    c.y0 = r * a.box[1];
    c.b = a.div_x.divmod(t, r);
    c.x0 = r * a.box[0];
    return c;
}

// One 8-channel group of the fused epilogue: y = post * (max(z * sc + sh [+ r], floor) [+ r]).
// RES: 0 no residual, 1 residual before the activation, 2 after it.
template <int RES, bool POST>
__device__ __forceinline__ uint4 epilogue8(const uint32_t *v, uint32_t ss_addr, uint4 res, float act_floor,
                                           float post_scale) {
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; e += 2) {   // (scale, shift) of two channels per 16-byte shared-memory read
        const uint4 ss = lds128(ss_addr + (uint32_t)e * 8);
        f[e] = fmaf(__uint_as_float(v[e]), __uint_as_float(ss.x), __uint_as_float(ss.y));
        f[e + 1] = fmaf(__uint_as_float(v[e + 1]), __uint_as_float(ss.z), __uint_as_float(ss.w));
    }
    float r[8];
    if (RES != 0) {
        const uint32_t w[4] = {res.x, res.y, res.z, res.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {   // bf16 -> fp32 is a 16-bit shift
            r[2 * e] = __uint_as_float(w[e] << 16);
            r[2 * e + 1] = __uint_as_float(w[e] & 0xffff0000u);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        if (RES == 1) f[e] += r[e];
        f[e] = fmaxf(f[e], act_floor);
        if (RES == 2) f[e] += r[e];
        if (POST) f[e] *= post_scale;
    }
    uint4 o;
    __nv_bfloat162 b0 = __floats2bfloat162_rn(f[0], f[1]), b1 = __floats2bfloat162_rn(f[2], f[3]);
    __nv_bfloat162 b2 = __floats2bfloat162_rn(f[4], f[5]), b3 = __floats2bfloat162_rn(f[6], f[7]);
    o.x = *reinterpret_cast<uint32_t *>(&b0), o.y = *reinterpret_cast<uint32_t *>(&b1);
    o.z = *reinterpret_cast<uint32_t *>(&b2), o.w = *reinterpret_cast<uint32_t *>(&b3);
    return o;
}

// The same for fp32 storage (tf32 layers): four channels per 16-byte cell.
template <int RES, bool POST>
__device__ __forceinline__ uint4 epilogue4(const uint32_t *v, uint32_t ss_addr, uint4 res, float act_floor,
                                           float post_scale) {
    float f[4];
#pragma unroll
    for (int e = 0; e < 4; e += 2) {
        const uint4 ss = lds128(ss_addr + (uint32_t)e * 8);
        f[e] = fmaf(__uint_as_float(v[e]), __uint_as_float(ss.x), __uint_as_float(ss.y));
        f[e + 1] = fmaf(__uint_as_float(v[e + 1]), __uint_as_float(ss.z), __uint_as_float(ss.w));
    }
    const float r[4] = {__uint_as_float(res.x), __uint_as_float(res.y), __uint_as_float(res.z), __uint_as_float(res.w)};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        if (RES == 1) f[e] += r[e];
        f[e] = fmaxf(f[e], act_floor);
        if (RES == 2) f[e] += r[e];
        if (POST) f[e] *= post_scale;
    }
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
}

// TF32 = false: bf16 storage, kind::f16 MMAs, 64 channels per 128-byte k-step and per output slab.
// TF32 = true : fp32 storage, kind::tf32 MMAs (the TMA rounds the operands to tf32), 32 channels per
//               k-step; a 64-channel output chunk is two 32-channel slabs, one per epilogue warp half.
template <bool TF32, bool X3 = false>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv3d_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_y, const __grid_constant__ CUtensorMap map_r,
                 const __grid_constant__ TcArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve-up (all tiles 1024-byte aligned for the 128B swizzle)
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int b_bytes = a.n_tile * 128;
    const int stage_bytes = TC_A_BYTES + b_bytes;
    uint8_t *stage_base = smem;
    constexpr int EK = TF32 ? 32 : 64;    // channels per k-step (one 128-byte swizzle row)
    constexpr int SPC = TF32 ? 2 : 1;     // 128-byte-wide slabs per 64-channel output chunk
    uint8_t *slab_base = smem + (size_t)a.stages * stage_bytes;  // n_slabs slots of SPC slabs of 128 x 128 B
    uint64_t *bars = reinterpret_cast<uint64_t *>(slab_base + (size_t)a.n_slabs * SPC * TC_A_BYTES);
    uint64_t *full_bar = bars, *empty_bar = bars + a.stages;
    uint64_t *acc_full = bars + 2 * a.stages, *acc_empty = acc_full + 2;
    uint64_t *res_bar = acc_empty + 2;               // residual tile landed in slab s (up to 8 slabs)
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(res_bar + 8);
    float2 *ss_buf = reinterpret_cast<float2 *>(reinterpret_cast<uint8_t *>(bars) + 512);   // 64 x (scale, shift)
    volatile uint32_t *split_flag = tmem_slot + 1;   // "this CTA finishes the tile" (split-K)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tmem_cols = a.n_tile <= 16 ? 32 : (a.n_tile <= 32 ? 64 : (a.n_tile <= 64 ? 128 : (a.n_tile <= 128 ? 256 : 512)));

    if (threadIdx.x == 0) {
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(smem_u32(full_bar + s), 1);
            mbar_init(smem_u32(empty_bar + s), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(acc_full + i), 1);
            mbar_init(smem_u32(acc_empty + i), TC_EPI_WARPS);  // one arrive per epilogue warp
        }
        for (int i = 0; i < 8; ++i) mbar_init(smem_u32(res_bar + i), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_y) : "memory");
        if (a.res_tma) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_r) : "memory");
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), tmem_cols);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    const int acc_stride = tmem_cols / 2;  // columns between the two accumulators

    const int k_chunks = a.c_in / EK;
    const int n_taps = a.taps[0] * a.taps[1] * a.taps[2];
    const int pass_steps = n_taps * k_chunks;            // k-steps of one pass over (tap, chunk)
    const int k_steps = a.x3 ? 3 * pass_steps : pass_steps;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
                const TileCoord tc = decode_tile(a, t);
                const int k_lo = (int)(t % a.k_splits) * a.steps_per_split;
                const int k_hi = min(k_steps, k_lo + a.steps_per_split);
                for (int ks0 = k_lo; ks0 < k_hi; ++ks0) {
                    const int pass = a.x3 ? ks0 / pass_steps : 0, ks = ks0 - pass * pass_steps;
                    const int tap = ks / k_chunks, kc = ks - tap * k_chunks;
                    const int kz = tap % a.taps[2], t2 = tap / a.taps[2];
                    const int ky = t2 % a.taps[1], kx = t2 / a.taps[1];
                    const int cx = tc.x0 * a.stride[0] - a.pad[0] + kx;
                    const int cy = tc.y0 * a.stride[1] - a.pad[1] + ky;
                    const int cz = tc.z0 * a.stride[2] - a.pad[2] + kz;
                    mbar_wait(smem_u32(empty_bar + stage), phase ^ 1);
                    const uint32_t bar = smem_u32(full_bar + stage);
                    const uint32_t sa = smem_u32(stage_base + (size_t)stage * stage_bytes);
                    mbar_expect_tx(bar, (uint32_t)(a.box[0] * a.box[1] * a.box[2] * 128 + b_bytes));
                    // x3: pass 0 = (x_lo, w_hi), pass 1 = (x_hi, w_lo), pass 2 = (x_hi, w_hi)
                    const int xch = (a.x3 && pass == 0 ? a.c_in : 0) + kc * EK;
                    const int wcol = (a.x3 && pass == 1 ? n_taps * a.c_in : 0) + tap * a.c_in + kc * EK;
                    tma_load_5d(sa, &map_x, bar, xch, cz, cy, cx, tc.b);
                    tma_load_2d(sa + TC_A_BYTES, &map_w, bar, wcol, tc.n0);
                    if (++stage == (uint32_t)a.stages) stage = 0, phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer ------------------------------
        if (lane == 0) {
            const uint32_t idesc = TF32 ? umma_idesc_tf32(a.n_tile) : umma_idesc(a.n_tile);
            uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
            for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
                if (!X3) {
                    mbar_wait(smem_u32(acc_empty + acc), acc_phase ^ 1);  // epilogue drained this accumulator
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                uint32_t d_tmem = tmem_base + acc * acc_stride;
                const int k_lo = (int)(t % a.k_splits) * a.steps_per_split;
                const int my_steps = min(k_steps, k_lo + a.steps_per_split) - k_lo;
                int in_period = 0;                 // x3: k-steps issued into the current accumulation period
                for (int ks = 0; ks < my_steps; ++ks) {
                    if (X3 && in_period == 0) {    // a new period starts in the other accumulator
                        mbar_wait(smem_u32(acc_empty + acc), acc_phase ^ 1);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        d_tmem = tmem_base + acc * acc_stride;
                    }
                    mbar_wait(smem_u32(full_bar + stage), phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = smem_u32(stage_base + (size_t)stage * stage_bytes);
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + TC_A_BYTES);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {   // +32 bytes per MMA (K = 16 bf16 / 8 tf32) inside the swizzle row
                        if (TF32) umma_tf32(d_tmem, da + 2 * k, db + 2 * k, idesc, ((X3 ? in_period : ks) | k) != 0);
                        else umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (ks | k) != 0);
                    }
                    umma_commit(smem_u32(empty_bar + stage));  // frees the smem stage when the MMAs retire
                    if (++stage == (uint32_t)a.stages) stage = 0, phase ^= 1;
                    if (X3 && (++in_period == a.drain_steps || ks + 1 == my_steps)) {
                        umma_commit(smem_u32(acc_full + acc));  // period complete
                        if (++acc == 2) acc = 0, acc_phase ^= 1;
                        in_period = 0;
                    }
                }
                if (!X3) {
                    umma_commit(smem_u32(acc_full + acc));  // accumulator complete
                    if (++acc == 2) acc = 0, acc_phase ^= 1;
                }
            }
        }
    } else {
        // ------------------------------ epilogue (8 warps) ------------------------------
        const int q = warp & 3;             // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;   // which 32 of a slab's 64 channels this warp handles
        const int row = q * 32 + lane;      // tile row = accumulator lane
        const int box_rows = a.box[0] * a.box[1] * a.box[2];
        const int iz = row % a.box[2], iy = (row / a.box[2]) % a.box[1], ix = row / (a.box[2] * a.box[1]);
        const bool epi_leader = (warp == 2 && lane == 0);
        const int epi_tid = threadIdx.x - 64;
        const uint32_t ss_addr = smem_u32(ss_buf);
        // y = post * (max(z [+ r], floor) [+ r]): residual before or after the activation
        const int epi_mode = (a.res ? (a.residual_after_act ? 2 : 1) : 0) + (a.post_scale != 1.f ? 3 : 0);
        const float act_floor = a.relu ? 0.f : -INFINITY;
        uint32_t acc = 0, acc_phase = 0, res_phase = 0;   // res_phase: bit s = parity of res_bar[s]
        int slab = 0;
        // residual by TMA: chunk i (64 channels of a tile) is staged in slab i % n_slabs; its residual
        // is requested n_slabs - 1 chunks ahead, as soon as the store that last read that slab has drained
        const int chunks_per_tile = a.n_tile / 64;
        long long chunk_i = 0;
        auto request_residual = [&](long long ci) {       // leader only
            const long long t2 = blockIdx.x + (ci / chunks_per_tile) * (long long)gridDim.x;
            if (t2 >= a.n_tiles) return;
            const TileCoord c2 = decode_tile(a, t2);
            const int into = (int)(ci % a.n_slabs);
            const uint32_t bar = smem_u32(res_bar + into);
            mbar_expect_tx(bar, (uint32_t)(box_rows * 128 * SPC));
            for (int sl = 0; sl < SPC; ++sl)
                tma_load_5d(smem_u32(slab_base + (size_t)(into * SPC + sl) * TC_A_BYTES), &map_r, bar,
                            c2.n0 + (int)(ci % chunks_per_tile) * 64 + sl * 32, c2.z0, c2.y0, c2.x0, c2.b);
        };
        if (a.res_tma && epi_leader)
            for (int ci = 0; ci < a.n_slabs - 1; ++ci) request_residual(ci);
        for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
            const TileCoord tc = decode_tile(a, t);
            const int tap_t = a.transposed ? tc.n0 / a.c_out : 0;   // which (kx,ky,kz) this n-tile is
            const int ch0 = a.transposed ? tc.n0 - tap_t * a.c_out : tc.n0;
            const int tkz = tap_t % a.up[2], tky = (tap_t / a.up[2]) % a.up[1], tkx = tap_t / (a.up[2] * a.up[1]);
            const int ox = tc.x0 + ix, oy = tc.y0 + iy, oz = tc.z0 + iz;
            const bool row_ok = row < box_rows && ox < a.out_dims[0] && oy < a.out_dims[1] && oz < a.out_dims[2];
            const long long ovox = (((long long)tc.b * a.out_dims[0] + ox) * a.out_dims[1] + oy) * a.out_dims[2] + oz;
            if (X3) {
                // sum the accumulation periods of this tile in registers and put the sum back into the last
                // accumulator: this warp's 32 lanes x its 32 columns of every 64-column chunk -- exactly the
                // cells it reads below, so no other warp has to wait for the write-back
                const int k_lo = (int)(t % a.k_splits) * a.steps_per_split;
                const int my_steps = min(k_steps, k_lo + a.steps_per_split) - k_lo;
                x3_accumulate_periods(tmem_base, (uint32_t)acc_stride, ((uint32_t)(q * 32) << 16) + half * 32,
                                      a.n_tile / 64, 64u, (my_steps + a.drain_steps - 1) / a.drain_steps, acc_full,
                                      acc_empty, acc, acc_phase, lane);
            } else {
                mbar_wait(smem_u32(acc_full + acc), acc_phase);
            }
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float *ws_row = nullptr;
            if (a.k_splits > 1) {
                // phase A (every CTA of the tile): add the partial accumulator into the workspace
                const long long out_tile = t / a.k_splits;
                ws_row = a.ws + (out_tile * TC_ROWS + row) * a.n_tile;
                for (int c = half * 32; c < a.n_tile; c += 64) {
                    uint32_t v[32];
                    tmem_ld32(tmem_base + acc * acc_stride + ((uint32_t)(q * 32) << 16) + c, v);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        atomicAdd(reinterpret_cast<float4 *>(ws_row + c + j),
                                  make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                              __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
                }
                // the accumulator is free again; publish the partials, then count this CTA in
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(acc_empty + acc));
                if (++acc == 2) acc = 0, acc_phase ^= 1;
                __threadfence();
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (epi_leader) *split_flag = atomicAdd(a.counters + out_tile, 1u) == (unsigned)(a.k_splits - 1);
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (!*split_flag) continue;   // another CTA finishes this tile
                __threadfence();              // phase B (last CTA): all partials are visible
            }
            for (int c0 = 0; c0 < a.n_tile; c0 += 64) {
                uint8_t *slot_ptr = slab_base + (size_t)slab * SPC * TC_A_BYTES;
                uint8_t *slab_ptr = slot_ptr + (TF32 ? half * TC_A_BYTES : 0);   // the slab this warp writes
                // (scale, shift) of this chunk's 64 channels -> shared memory (everyone is past the
                // previous chunk's barrier, so nobody still reads the old ones)
                if (epi_tid < 64) {
                    const int c = ch0 + c0 + epi_tid;
                    ss_buf[epi_tid] = make_float2(a.scale ? __ldg(a.scale + c) : 1.f, a.shift ? __ldg(a.shift + c) : 0.f);
                }
                // the TMA store that last read this slab must be done before we overwrite it (with the
                // residual by TMA that was checked before its load was issued)
                if (!a.res_tma && epi_leader) {
                    if (a.n_slabs > 1) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (a.res_tma) {   // the residual of this chunk has landed in the slab
                    mbar_wait(smem_u32(res_bar + slab), (res_phase >> slab) & 1u);
                    res_phase ^= 1u << slab;
                }
                {
                    const int h = half * 32;
                    uint32_t v[32];
                    if (ws_row == nullptr) {
                        tmem_ld32(tmem_base + acc * acc_stride + ((uint32_t)(q * 32) << 16) + c0 + h, v);
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    } else {   // sums of all splits, straight from L2; leave zeros behind
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float4 *p4 = reinterpret_cast<float4 *>(ws_row + c0 + h + j);
                            const float4 f4 = __ldcg(p4);
                            __stcg(p4, make_float4(0.f, 0.f, 0.f, 0.f));
                            v[j] = __float_as_uint(f4.x), v[j + 1] = __float_as_uint(f4.y);
                            v[j + 2] = __float_as_uint(f4.z), v[j + 3] = __float_as_uint(f4.w);
                        }
                    }
                    const int ch = ch0 + c0 + h;
                    const uint32_t row_addr = smem_u32(slab_ptr) + row * 128;
                    auto run = [&](auto res_c, auto post_c) {
                        constexpr int RES = decltype(res_c)::value;
                        constexpr bool POST = decltype(post_c)::value;
                        if constexpr (!TF32) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8) {
                                const int chunk = (h + j) >> 3;  // 16-byte chunk inside the 128-byte row
                                const uint32_t cell = row_addr + ((chunk ^ (row & 7)) << 4);
                                uint4 u = make_uint4(0u, 0u, 0u, 0u);
                                if (RES != 0) {
                                    if (a.res_tma) u = lds128(cell);
                                    else if (row_ok)
                                        u = __ldg(reinterpret_cast<const uint4 *>(
                                            static_cast<const __nv_bfloat16 *>(a.res) + ovox * a.c_out + ch + j));
                                }
                                sts128(cell, epilogue8<RES, POST>(v + j, ss_addr + (uint32_t)(h + j) * 8, u, act_floor,
                                                                  a.post_scale));
                            }
                        } else {   // this warp's 32 channels are one whole 128-byte row of its own slab
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                const uint32_t cell = row_addr + (((j >> 2) ^ (row & 7)) << 4);
                                uint4 u = make_uint4(0u, 0u, 0u, 0u);
                                if (RES != 0) {
                                    if (a.res_tma) u = lds128(cell);
                                    else if (row_ok)
                                        u = __ldg(reinterpret_cast<const uint4 *>(
                                            static_cast<const float *>(a.res) + ovox * a.c_out + ch + j));
                                }
                                sts128(cell, epilogue4<RES, POST>(v + j, ss_addr + (uint32_t)(h + j) * 8, u, act_floor,
                                                                  a.post_scale));
                            }
                        }
                    };
                    using std::integral_constant;
                    switch (epi_mode) {
                        case 0: run(integral_constant<int, 0>{}, integral_constant<bool, false>{}); break;
                        case 1: run(integral_constant<int, 1>{}, integral_constant<bool, false>{}); break;
                        case 2: run(integral_constant<int, 2>{}, integral_constant<bool, false>{}); break;
                        case 3: run(integral_constant<int, 0>{}, integral_constant<bool, true>{}); break;
                        case 4: run(integral_constant<int, 1>{}, integral_constant<bool, true>{}); break;
                        default: run(integral_constant<int, 2>{}, integral_constant<bool, true>{}); break;
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // st.shared -> visible to TMA
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (epi_leader) {
                    for (int sl = 0; sl < SPC; ++sl) {
                        const uint32_t src = smem_u32(slot_ptr + (size_t)sl * TC_A_BYTES);
                        if (a.transposed)
                            tma_store_5d(&map_y, src, ch0 + c0 + sl * 32, tc.z0 * a.up[2] + tkz,
                                         tc.y0 * a.up[1] + tky, tc.x0 * a.up[0] + tkx, tc.b);
                        else
                            tma_store_5d(&map_y, src, tc.n0 + c0 + sl * 32, tc.z0, tc.y0, tc.x0, tc.b);
                    }
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    if (a.res_tma) {   // once the store of the previous chunk has drained, its slab takes a new chunk
                        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                        request_residual(chunk_i + a.n_slabs - 1);
                    }
                }
                ++chunk_i;
                if (++slab == a.n_slabs) slab = 0;
            }
            if (ws_row != nullptr) {
                if (epi_leader) a.counters[t / a.k_splits] = 0;   // ready for the next launch
                continue;                                          // (accumulator already handed back)
            }
            // accumulator fully read: hand it back to the MMA warp
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(acc_empty + acc));
            if (++acc == 2) acc = 0, acc_phase ^= 1;
        }
        if (epi_leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ------------------------------- host side --------------------------------
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

int encode_map_bf16(CUtensorMap *map, int rank, void *base, const cuuint64_t *dims,
                      const cuuint64_t *strides_bytes, const cuuint32_t *box, const cuuint32_t *estrides,
                      const char *what) {
    return encode_map_sw128(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, base, dims, strides_bytes, box, estrides,
                            what);
}

int encode_map_sw128(CUtensorMap *map, CUtensorMapDataType dtype, int rank, void *base, const cuuint64_t *dims,
                     const cuuint64_t *strides_bytes, const cuuint32_t *box, const cuuint32_t *estrides,
                     const char *what) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (fn == nullptr) return fail(IVX_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    CUresult r = fn(map, dtype, (cuuint32_t)rank, base, dims, strides_bytes, box,
                    estrides, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(IVX_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed with %d", what, (int)r);
    return IVX_OK;
}

// box of output voxels with bx*by*bz <= 128 that wastes the fewest accumulator rows
static void pick_box(const int32_t out[3], const int32_t stride[3], int box[3], int tiles[3]) {
    double best = -1.0;
    for (int bz = 1; bz <= out[2] && bz <= 128; ++bz) {
        if (bz * stride[2] > 256) break;
        for (int by = 1; by <= out[1] && bz * by <= 128; ++by) {
            if (by * stride[1] > 256) break;
            int bx = 128 / (bz * by);
            if (bx > out[0]) bx = out[0];
            if (bx * stride[0] > 256) bx = 256 / stride[0];
            const long long t0 = (out[0] + bx - 1) / bx, t1 = (out[1] + by - 1) / by, t2 = (out[2] + bz - 1) / bz;
            const double eff = (double)out[0] * out[1] * out[2] / (double)(t0 * t1 * t2 * 128);
            // prefer long z runs on ties: they are contiguous in memory
            if (eff > best + 1e-9) {
                best = eff;
                box[0] = bx, box[1] = by, box[2] = bz;
                tiles[0] = (int)t0, tiles[1] = (int)t1, tiles[2] = (int)t2;
            }
        }
    }
}

// ivx_conv3d_tc_halo.cu
double plan_halo(const ivx_conv3d_params *p, int perm[3], int *mt_out, double *cost_out = nullptr);
int launch_conv3d_tc_halo(const ivx_conv3d_params *p, const int perm[3], int mt, const void *x,
                          const void *weight_bf16, const float *scale, const float *shift,
                          const void *residual, void *y, cudaStream_t stream);

}  // namespace ivx

using namespace ivx;

static int validate_tc(const ivx_conv3d_params *p, bool has_residual, bool free_output = false) {
    IVX_REQUIRE(p != nullptr, "NULL params");
    const bool bf16 = p->in_dtype == IVX_BF16 && p->out_dtype == IVX_BF16;
    const bool tf32 = (p->in_dtype == IVX_F32 || p->in_dtype == IVX_F32X2) && p->out_dtype == IVX_F32;
    if (!bf16 && !tf32)
        return fail(IVX_ERR_UNSUPPORTED,
                    "tensor-core path: bf16 in and out, or fp32 out with fp32 (tf32 math) or split fp32 (3xTF32) in");
    if (p->c_in % (bf16 ? 64 : 32) != 0 || p->c_out % 64 != 0)
        return fail(IVX_ERR_UNSUPPORTED,
                    "tensor-core path: c_in must be a multiple of 64 (bf16) or 32 (tf32), c_out of 64");
    if (p->transposed && has_residual)
        return fail(IVX_ERR_UNSUPPORTED, "tensor-core path: transposed layers take no residual");
    for (int d = 0; d < 3; ++d) {
        const int expect = p->transposed ? p->in_dims[d] * p->stride[d]
                                         : (p->in_dims[d] + 2 * p->pad[d] - p->kernel[d]) / p->stride[d] + 1;
        IVX_REQUIRE(p->kernel[d] >= 1 && p->stride[d] >= 1 && p->stride[d] <= 8 && p->pad[d] >= 0 &&
                        (expect == p->out_dims[d] || free_output) && p->out_dims[d] > 0,
                    "bad geometry on axis %d", d);
        if (p->transposed)
            IVX_REQUIRE(p->kernel[d] == p->stride[d] && p->pad[d] == 0,
                        "transposed layers need kernel == stride and no padding (axis %d)", d);
    }
    IVX_REQUIRE(p->batch > 0, "batch must be positive");
    return IVX_OK;
}

static void plan_tc(const ivx_conv3d_params *p, TcArgs &a);

// does the halo-reuse kernel take this layer?  (IVX_TC_HALO=0 keeps everything on the per-tap one)
static bool use_halo(const ivx_conv3d_params *p, int perm[3], int *mt) {
    static const bool halo_on = [] {
        const char *e = getenv("IVX_TC_HALO");
        return e == nullptr || e[0] != '0';
    }();
    static const int halo_max_cout = [] {
        const char *e = getenv("IVX_HALO_MAX_COUT");
        return e ? atoi(e) : 1024;
    }();
    static const bool by_score = [] {
        const char *e = getenv("IVX_HALO_PLAN");
        return e != nullptr && e[0] == 's';
    }();
    if (!halo_on || p->c_out > halo_max_cout) return false;
    double halo_cost = 0.0;
    const double eff = plan_halo(p, perm, mt, &halo_cost);
    if (eff <= 0.0) return false;
    if (by_score)
        // old rule: with one tile per CTA (mt == 1) an MMA covers only 128 voxels, i.e. half the tensor rate:
        // still a win for <= 128 channels (the per-tap kernel is at half rate there too), not above
        return eff >= 0.7 && (*mt == 2 || p->c_out <= 128);
    // the same unit (M128 MMA issue slots per CTA) for the per-tap kernel: it re-reads its activation box for
    // every tap through L2, which measured at about half the MMA rate of the halo kernel (x 2); its MMAs
    // have N = n_tile <= 256 columns, the halo kernel's 128 lanes hold min(c_out, 128) channels
    TcArgs a;
    plan_tc(p, a);
    // Measured (ncu launch lists of the four necks at batch 1 / 2 / 8 / 32, profiles/planner_r2.txt): with
    // 256-channel tiles the per-tap kernel runs within 1.3x of its MMA slots, with 128 / 64-channel tiles
    // (half / quarter of the columns of an MMA) it is also bound by the L2 -> shared-memory traffic;
    // a split-K item pays ~30 k-steps per 64 channels for the fp32 atomics and the read-back.
    const double pen = a.n_tile >= 256 ? 1.3 : 1.5;
    const double tap_cost = (double)((a.n_tiles + kNumSMs - 1) / kNumSMs) * a.steps_per_split * 4.0 * pen + 50.0 +
                            (a.k_splits > 1 ? 4.0 * 30.0 * a.n_tile / 64.0 : 0.0);
    // both costs count whole tiles of their own shape, so they compare directly
    return halo_cost <= tap_cost;
}

// tiling of the per-tap kernel, including the split-K decision
static void plan_tc(const ivx_conv3d_params *p, TcArgs &a) {
    a.c_in = p->c_in, a.c_out = p->c_out;
    a.transposed = p->transposed;
    const int up_taps = p->transposed ? p->kernel[0] * p->kernel[1] * p->kernel[2] : 1;
    const int n_total = p->c_out * up_taps;      // GEMM columns
    a.n_tile = p->c_out % 256 == 0 ? 256 : (p->c_out % 128 == 0 ? 128 : 64);
    a.n_splits = n_total / a.n_tile;
    const int32_t ones[3] = {1, 1, 1};
    const int32_t *row_dims = p->transposed ? p->in_dims : p->out_dims;   // what the tiles cover
    const int32_t *ld_stride = p->transposed ? ones : p->stride;          // element strides of the loads
    for (int d = 0; d < 3; ++d) {
        a.taps[d] = p->transposed ? 1 : p->kernel[d];
        a.stride[d] = ld_stride[d];
        a.pad[d] = p->transposed ? 0 : p->pad[d];
        a.out_dims[d] = row_dims[d];
        a.up[d] = p->transposed ? p->stride[d] : 1;
    }
    a.batch = p->batch;
    pick_box(row_dims, p->transposed ? p->stride : ld_stride, a.box, a.tiles);
    const long long out_tiles = (long long)p->batch * a.tiles[0] * a.tiles[1] * a.tiles[2] * a.n_splits;
    // split-K: the persistent grid runs ceil(work items / 148) waves, so a layer whose tile count is
    // small or just above a multiple of the SM count leaves SMs idle in the last wave (200 tiles =
    // 1.35 waves cost 2).  Sharing each tile's k-steps between k_splits CTAs changes the item count;
    // pick the split that minimises waves x k-steps per item, charging 3% per extra split for the
    // fp32 partials that go through the L2 workspace (needs a workspace; the caller may not have
    // one -> k_splits reset to 1 later).
    a.x3 = p->in_dtype == IVX_F32X2 ? 1 : 0;
    {
        const char *e = getenv("IVX_X3_DRAIN_STEPS");    // experiments: a huge value = one period per tile
        a.drain_steps = e ? atoi(e) : 8;
        if (a.drain_steps < 1) a.drain_steps = 1;
    }
    const int k_steps = a.taps[0] * a.taps[1] * a.taps[2] * (p->c_in / (p->in_dtype == IVX_BF16 ? 64 : 32)) *
                        (a.x3 ? 3 : 1);
    a.k_splits = 1;
    a.steps_per_split = k_steps;
    {
        auto cost = [&](int splits, int steps) {
            const long long waves = (out_tiles * splits + kNumSMs - 1) / kNumSMs;
            // + 2: per-item epilogue; a split item also adds its 128 x n_tile fp32 partials with vector
            // atomics and the last one reads them back: ~30 k-steps per 64 channels (measured, batch-1 necks)
            return (double)waves * (steps + 2.0 + (splits > 1 ? 30.0 * a.n_tile / 64.0 : 0.0));
        };
        double best = cost(1, k_steps);
        for (int want = 2; want <= 8 && want * 4 <= k_steps; ++want) {           // at least 4 k-steps per CTA
            const int steps = (k_steps + want - 1) / want;
            const int splits = (k_steps + steps - 1) / steps;
            const double c = cost(splits, steps);
            if (c < 0.93 * best) best = c, a.k_splits = splits, a.steps_per_split = steps;
        }
    }
    a.n_tiles = out_tiles * a.k_splits;
    a.div_k = make_tilediv(a.k_splits), a.div_n = make_tilediv(a.n_splits);
    a.div_z = make_tilediv(a.tiles[2]), a.div_y = make_tilediv(a.tiles[1]), a.div_x = make_tilediv(a.tiles[0]);
}

extern "C" int64_t ivx_conv3d_tc_workspace_bytes(const ivx_conv3d_params *p) {
    if (validate_tc(p, false, true) != IVX_OK) return 0;
    int perm[3] = {0, 1, 2}, mt = 1;
    if (use_halo(p, perm, &mt)) return 0;
    TcArgs a;
    plan_tc(p, a);
    if (a.k_splits <= 1) return 0;
    const long long out_tiles = a.n_tiles / a.k_splits;
    return out_tiles * TC_ROWS * a.n_tile * (int64_t)sizeof(float) + out_tiles * (int64_t)sizeof(unsigned);
}

// Launch of the per-tap kernel in either storage type (the bf16 entry point tries the halo kernel first).
template <bool TF32, bool X3 = false>
static int launch_tc(const ivx_conv3d_params *p, const void *x, const void *weight, const float *scale,
                     const float *shift, const void *residual, void *y, void *workspace, int64_t workspace_bytes,
                     cudaStream_t stream, const int64_t *y_strides = nullptr) {
    constexpr int EK = TF32 ? 32 : 64;           // channels per 128-byte k-step
    constexpr int ES = TF32 ? 4 : 2;             // bytes per element
    constexpr int SPC = TF32 ? 2 : 1;            // 128-byte-wide slabs per 64-channel output chunk
    // operands of the tf32 MMAs are rounded to tf32 by the TMA; outputs and residuals stay plain fp32
    const CUtensorMapDataType dt_in = TF32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const CUtensorMapDataType dt_out = TF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    TcArgs a;
    plan_tc(p, a);
    a.ws = nullptr, a.counters = nullptr;
    const int k_steps_total = a.taps[0] * a.taps[1] * a.taps[2] * (p->c_in / EK) * (a.x3 ? 3 : 1);
    const int cin_mem = a.x3 ? 2 * p->c_in : p->c_in;    // channels per voxel in memory (hi | lo)
    if (a.k_splits > 1) {
        const long long out_tiles = a.n_tiles / a.k_splits;
        const int64_t ws_floats = out_tiles * TC_ROWS * a.n_tile;
        const int64_t need = ws_floats * (int64_t)sizeof(float) + out_tiles * (int64_t)sizeof(unsigned);
        if (workspace != nullptr && workspace_bytes >= need &&
            reinterpret_cast<uintptr_t>(workspace) % 16 == 0) {
            a.ws = static_cast<float *>(workspace);
            a.counters = reinterpret_cast<unsigned *>(a.ws + ws_floats);
        } else {   // no (or too small a) workspace: one CTA per tile, still correct
            a.n_tiles = out_tiles;
            a.k_splits = 1;
            a.div_k = make_tilediv(1);
            a.steps_per_split = k_steps_total;
        }
    }
    a.relu = p->relu, a.residual_after_act = p->residual_after_act, a.post_scale = p->post_scale;
    a.scale = scale, a.shift = shift, a.res = residual;
    const int32_t ones[3] = {1, 1, 1};
    const int32_t *ld_stride = p->transposed ? ones : p->stride;
    const int up_taps = p->transposed ? p->kernel[0] * p->kernel[1] * p->kernel[2] : 1;
    const int n_total = p->c_out * up_taps;

    const int stage_bytes = TC_A_BYTES + a.n_tile * 128;
    const int slot_bytes = SPC * TC_A_BYTES;     // staging of one 64-channel output chunk
    const int budget = 227 * 1024 - 1024 /* barriers */ - 1024 /* alignment slack */;
    auto stages_with = [&](int n_slots) {
        const int st = (budget - n_slots * slot_bytes) / stage_bytes;
        return st > 8 ? 8 : st;
    };
    // Output staging: two slots, so the store of one chunk overlaps the next (one, if that buys the
    // fourth pipeline stage).  Residual through TMA needs extra slots: a layer with few k-steps is bound
    // by the epilogue and by the latency of the residual load, not by the main loop, so it gets a short
    // pipeline and up to 8 slots (residual requested 7 chunks ahead); otherwise one extra slot, if 4
    // stages remain.
    a.n_slabs = (stages_with(2) < 4 && stages_with(1) >= 4) ? 1 : 2;
    a.res_tma = 0;
    if (residual != nullptr && a.k_splits == 1 && !p->transposed) {
        if (k_steps_total <= 16) {
            for (int st = 4; st >= 2 && !a.res_tma; --st) {
                int slots = (budget - st * stage_bytes) / slot_bytes;
                if (slots > 8) slots = 8;
                if (slots >= 4) a.n_slabs = slots, a.res_tma = 1;
            }
        } else if (stages_with(3) >= 4) {
            a.n_slabs = 3, a.res_tma = 1;
        }
    }
    a.stages = stages_with(a.n_slabs);
    IVX_REQUIRE(a.stages >= 2, "not enough shared memory for a 2-stage pipeline");
    const size_t smem = (size_t)a.stages * stage_bytes + (size_t)a.n_slabs * slot_bytes + 2048;

    alignas(64) CUtensorMap map_x, map_w, map_y, map_r;
    {
        const cuuint64_t dims[5] = {(cuuint64_t)cin_mem, (cuuint64_t)p->in_dims[2], (cuuint64_t)p->in_dims[1],
                                    (cuuint64_t)p->in_dims[0], (cuuint64_t)p->batch};
        cuuint64_t st[4];
        st[0] = (cuuint64_t)cin_mem * ES;
        st[1] = st[0] * p->in_dims[2];
        st[2] = st[1] * p->in_dims[1];
        st[3] = st[2] * p->in_dims[0];
        const cuuint32_t box[5] = {EK, (cuuint32_t)(a.box[2] * ld_stride[2]),
                                   (cuuint32_t)(a.box[1] * ld_stride[1]), (cuuint32_t)(a.box[0] * ld_stride[0]), 1};
        const cuuint32_t es[5] = {1, (cuuint32_t)ld_stride[2], (cuuint32_t)ld_stride[1], (cuuint32_t)ld_stride[0], 1};
        int rc = encode_map_sw128(&map_x, dt_in, 5, const_cast<void *>(x), dims, st, box, es, "activations");
        if (rc != IVX_OK) return rc;
    }
    {
        // [c_out][taps * c_in] for convolutions, [taps * c_out][c_in] for transposed layers
        const int taps = p->transposed ? 1 : p->kernel[0] * p->kernel[1] * p->kernel[2];
        const cuuint64_t dims[2] = {(cuuint64_t)taps * cin_mem, (cuuint64_t)n_total};
        const cuuint64_t st[1] = {(cuuint64_t)taps * cin_mem * ES};
        const cuuint32_t box[2] = {EK, (cuuint32_t)a.n_tile};
        const cuuint32_t es[2] = {1, 1};
        int rc = encode_map_sw128(&map_w, dt_in, 2, const_cast<void *>(weight), dims, st, box, es, "weights");
        if (rc != IVX_OK) return rc;
    }
    {
        const cuuint64_t dims[5] = {(cuuint64_t)p->c_out, (cuuint64_t)p->out_dims[2], (cuuint64_t)p->out_dims[1],
                                    (cuuint64_t)p->out_dims[0], (cuuint64_t)p->batch};
        cuuint64_t st[4];
        st[0] = (cuuint64_t)p->c_out * ES;
        st[1] = st[0] * p->out_dims[2];
        st[2] = st[1] * p->out_dims[1];
        st[3] = st[2] * p->out_dims[0];
        if (y_strides != nullptr) {   // y is a strided view: element strides along X, Y, Z and between scenes
            st[0] = (cuuint64_t)y_strides[2] * ES, st[1] = (cuuint64_t)y_strides[1] * ES;
            st[2] = (cuuint64_t)y_strides[0] * ES, st[3] = (cuuint64_t)y_strides[3] * ES;
        }
        const cuuint32_t box[5] = {128 / ES, (cuuint32_t)(a.box[2] * a.up[2]), (cuuint32_t)(a.box[1] * a.up[1]),
                                   (cuuint32_t)(a.box[0] * a.up[0]), 1};
        const cuuint32_t es[5] = {1, (cuuint32_t)a.up[2], (cuuint32_t)a.up[1], (cuuint32_t)a.up[0], 1};
        int rc = encode_map_sw128(&map_y, dt_out, 5, y, dims, st, box, es, "output");
        if (rc != IVX_OK) return rc;
        map_r = map_y;      // the residual has the output's shape: same map over its own memory
        if (a.res_tma) {
            rc = encode_map_sw128(&map_r, dt_out, 5, const_cast<void *>(residual), dims, st, box, es, "residual");
            if (rc != IVX_OK) return rc;
        }
    }

    int rc = check_cuda(cudaFuncSetAttribute(conv3d_tc_kernel<TF32, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem),
                        "cudaFuncSetAttribute");
    if (rc != IVX_OK) return rc;
    const unsigned grid = (unsigned)(a.n_tiles < kNumSMs ? a.n_tiles : kNumSMs);
    conv3d_tc_kernel<TF32, X3><<<grid, TC_THREADS, smem, stream>>>(map_x, map_w, map_y, map_r, a);
    IVX_CHECK_LAUNCH("conv3d_tc launch");
    return IVX_OK;
}

extern "C" int ivx_conv3d_ndhwc_bf16_tc(const ivx_conv3d_params *p, const void *x, const void *weight_bf16,
                                        const float *scale, const float *shift, const void *residual,
                                        void *y, void *workspace, int64_t workspace_bytes, void *stream) {
    IVX_REQUIRE(p && x && weight_bf16 && y, "NULL argument");
    int rc0 = validate_tc(p, residual != nullptr);
    if (rc0 != IVX_OK) return rc0;
    IVX_REQUIRE(p->in_dtype == IVX_BF16, "ivx_conv3d_ndhwc_bf16_tc takes bf16 tensors");
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) |
                 reinterpret_cast<uintptr_t>(weight_bf16) | reinterpret_cast<uintptr_t>(residual)) % 16 == 0,
                "tensors must be 16-byte aligned");
    {
        int perm[3] = {0, 1, 2}, mt = 1;
        if (use_halo(p, perm, &mt))
            return launch_conv3d_tc_halo(p, perm, mt, x, weight_bf16, scale, shift, residual, y,
                                         (cudaStream_t)stream);
    }
    return launch_tc<false>(p, x, weight_bf16, scale, shift, residual, y, workspace, workspace_bytes,
                            (cudaStream_t)stream);
}

extern "C" int ivx_conv3d_ndhwc_tf32_tc(const ivx_conv3d_params *p, const float *x, const float *weight,
                                        const float *scale, const float *shift, const float *residual, float *y,
                                        void *workspace, int64_t workspace_bytes, void *stream) {
    IVX_REQUIRE(p && x && weight && y, "NULL argument");
    int rc0 = validate_tc(p, residual != nullptr);
    if (rc0 != IVX_OK) return rc0;
    IVX_REQUIRE(p->in_dtype == IVX_F32, "ivx_conv3d_ndhwc_tf32_tc takes fp32 tensors");
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(weight) |
                 reinterpret_cast<uintptr_t>(residual)) % 16 == 0,
                "tensors must be 16-byte aligned");
    {
        int perm[3] = {0, 1, 2}, mt = 1;
        if (use_halo(p, perm, &mt))
            return launch_conv3d_tc_halo(p, perm, mt, x, weight, scale, shift, residual, y, (cudaStream_t)stream);
    }
    return launch_tc<true>(p, x, weight, scale, shift, residual, y, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int ivx_conv3d_ndhwc_tf32_tc_view(const ivx_conv3d_params *p, const float *x, const float *weight,
                                             const float *scale, const float *shift, float *y,
                                             const int64_t *y_strides, void *workspace, int64_t workspace_bytes,
                                             void *stream) {
    IVX_REQUIRE(p && x && weight && y && y_strides, "NULL argument");
    int rc0 = validate_tc(p, false, true);
    if (rc0 != IVX_OK) return rc0;
    IVX_REQUIRE(p->in_dtype == IVX_F32 && !p->transposed, "plain convolutions of fp32 tensors only");
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(weight)) % 16 == 0,
                "tensors must be 16-byte aligned");
    for (int i = 0; i < 4; ++i)
        IVX_REQUIRE(y_strides[i] > 0 && y_strides[i] % 4 == 0, "output strides must be positive multiples of 4 elements");
    return launch_tc<true>(p, x, weight, scale, shift, nullptr, y, workspace, workspace_bytes, (cudaStream_t)stream,
                           y_strides);
}

extern "C" int ivx_conv3d_ndhwc_f32x3_tc_view(const ivx_conv3d_params *p, const float *x_split,
                                              const float *weight_split, const float *scale, const float *shift,
                                              float *y, const int64_t *y_strides, void *workspace,
                                              int64_t workspace_bytes, void *stream) {
    IVX_REQUIRE(p && x_split && weight_split && y && y_strides, "NULL argument");
    int rc0 = validate_tc(p, false, true);
    if (rc0 != IVX_OK) return rc0;
    IVX_REQUIRE(p->in_dtype == IVX_F32X2 && p->out_dtype == IVX_F32 && !p->transposed,
                "plain convolutions of split fp32 activations (IVX_F32X2) only");
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x_split) | reinterpret_cast<uintptr_t>(y) |
                 reinterpret_cast<uintptr_t>(weight_split)) % 16 == 0,
                "tensors must be 16-byte aligned");
    for (int i = 0; i < 4; ++i)
        IVX_REQUIRE(y_strides[i] > 0 && y_strides[i] % 4 == 0, "output strides must be positive multiples of 4 elements");
    return launch_tc<true, true>(p, x_split, weight_split, scale, shift, nullptr, y, workspace, workspace_bytes,
                                 (cudaStream_t)stream, y_strides);
}

extern "C" int ivx_conv3d_ndhwc_f32x3_tc(const ivx_conv3d_params *p, const float *x_split, const float *weight_split,
                                         const float *scale, const float *shift, const float *residual, float *y,
                                         void *workspace, int64_t workspace_bytes, void *stream) {
    IVX_REQUIRE(p && x_split && weight_split && y, "NULL argument");
    int rc0 = validate_tc(p, residual != nullptr);
    if (rc0 != IVX_OK) return rc0;
    IVX_REQUIRE(p->in_dtype == IVX_F32X2 && p->out_dtype == IVX_F32,
                "ivx_conv3d_ndhwc_f32x3_tc takes split fp32 activations (IVX_F32X2) and writes fp32");
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x_split) | reinterpret_cast<uintptr_t>(y) |
                 reinterpret_cast<uintptr_t>(weight_split) | reinterpret_cast<uintptr_t>(residual)) % 16 == 0,
                "tensors must be 16-byte aligned");
    {
        int perm[3] = {0, 1, 2}, mt = 1;
        if (use_halo(p, perm, &mt))
            return launch_conv3d_tc_halo(p, perm, mt, x_split, weight_split, scale, shift, residual, y,
                                         (cudaStream_t)stream);
    }
    return launch_tc<true, true>(p, x_split, weight_split, scale, shift, residual, y, workspace, workspace_bytes,
                                 (cudaStream_t)stream);
}
