// 3x3x3 stride-1 Conv3d on tcgen05 with activation-halo reuse.
//
// ivx_conv3d_tc.cu re-reads, for each of the 27 taps, a full 128-row activation
// box and the weight slice through L2; for layers with few channels that L2->SM
// traffic, not the tensor pipe, is the bound (ncu: ~9.5 TB/s, tensor pipe 18 %
// at c_in = c_out = 64).  Here the three spatial axes are given roles
//   P: 8 consecutive outputs  = one 8-row group of the MMA operand,
//   Q: 16 * MT outputs        = the 16 groups of MT accumulator tiles,
//   R: 1 output plane,
// and a CTA loads, per 64-channel chunk and per input plane r-1, r, r+1, ONE box
// of 10 (P: 8 + halo) x (16*MT + 2) (Q) voxels.  All nine (dP, dQ) taps of that
// plane are then MMA operands that merely START at a different row of the same
// box: the 8-row groups stay 10 rows (1280 B) apart;
// the swizzle is a function of the absolute shared-memory address for both the
// TMA and the MMA unit, so the shifted start needs nothing else.  One weight slice serves the MT tiles.  Activation traffic
// drops ~4x and weight traffic MT x compared with the per-tap kernel.
//
// Operand roles are SWAPPED relative to the per-tap kernel: the weight slice is the MMA's
// A operand (M = 128 accumulator lanes = output channels; a 64-channel layer leaves
// the upper lanes unused) and the voxels are the N dimension (N = 128 * MT).  Measured
// on B200, one tcgen05.mma with M = 128 costs >= ~128 cycles whatever N is (the A
// operand read), so N = 64/128-channel layers ran at 25 % / 50 % of the tensor peak
// with channels on N; with 256 voxels on N they run at the full M = 128 rate.  The
// accumulator is therefore D^T (lane = channel, column = voxel): the epilogue has the
// folded BN scale/shift as per-thread constants and transposes through shared memory.
//
// 64-channel layers would leave accumulator lanes 64..127 idle; instead they hold the NEXT
// output plane (pack2): an input plane p feeds output plane r0 with tap dR = p - r0 + pad and
// plane r0 + 1 with tap dR - 1, through the same shifted box, so the weight stage gets the two
// 64-row slices (an out-of-range tap is fetched from beyond the weight tensor: the TMA
// zero-fills it) and 4 input planes x 9 MMAs produce two output planes instead of 2 x 27.
//
// Warp roles (224 threads): 0 activation-plane producer, 1 MMA issuer, 2 weight
// producer, 3..6 epilogue.
//
// TF32 = true is the same kernel for fp32 tensors with kind::tf32 MMAs: a 128-byte row is 32 channels,
// so a chunk is 32 input channels and the output is staged as 32-channel fp32 slabs (four per
// 128-channel tile, which leaves room for one slab set and two plane stages).
//
// x3 (fp32-grade results from tf32 MMAs, "3xTF32"): every fp32 operand is the exact sum of a tf32 value
// `hi` and a remainder `lo`; x*w ~= x_lo*w_hi + x_hi*w_lo + x_hi*w_hi (the dropped x_lo*w_lo term and
// the rounding of `lo` to tf32 are both ~2^-22 of the product).  The activations come as
// [..][2*c_in] = (hi | lo) channels (ivx_split_tf32_ndhwc) and the weights as [c_out][2][27*c_in],
// so the three products are simply three times as many 32-channel chunks of the SAME loop: chunk
// kc < nc takes (x_lo, w_hi), nc <= kc < 2nc (x_hi, w_lo), the rest (x_hi, w_hi) -- small terms
// first.  The tensor core's truncating accumulate would still leave an error linear in K, so the x3
// instantiation hands an accumulator over after every `drain_groups` (chunk, plane) groups of 36 MMAs
// and eight epilogue warps (warps 7..10 help: each warp takes half of the columns of its TMEM lane
// quarter) sum the periods in registers (x3_accumulate_periods, ivx_tc_ptx.cuh); the sum is written
// back to TMEM and the usual epilogue (warps 3..6) runs on it.
#include <type_traits>

#include "ivx_tc_ptx.cuh"

namespace ivx {

constexpr int HL_THREADS = 224;
constexpr int HL_THREADS_X3 = 352;   // + 4 warps that help summing the accumulation periods
constexpr int HL_P = 8, HL_PPAD = 10, HL_QT = 16;  // outputs per group, box rows along P (8 + halo), groups per tile

struct HaloArgs {
    int c_in, c_out, n_tile, n_splits;
    int mt;                    // accumulator tiles per CTA along Q (1 or 2)
    int pack2;                 // 64-channel tiles: lanes 64..127 hold the NEXT output plane (r0 + 1)
    int dims[3];               // output extent along P, Q, R (== input extent + 2*pad - 2)
    int pad[3];                // padding along P, Q, R
    int stride_r;              // conv stride along R (P and Q are stride 1)
    int axis[3];               // tensor axis (0 = X, 1 = Y, 2 = Z) playing role P, Q, R
    int tiles[3];              // super-tiles along P, Q, R
    int batch;
    long long n_tiles;
    int a_stages, b_stages, n_slabs;
    long long ostride[3];      // output / residual element strides of one step along P, Q, R
    long long ostride_b;       // ... and of one batch item
    int relu, residual_after_act;
    float post_scale;
    const float *scale, *shift;
    const void *res;
    int x3;                    // 3xTF32: (hi | lo) activations and weights, three chunk passes
    int drain_groups;          // x3: (chunk, plane) groups of 36 MMAs per accumulation period
    int debug;                 // IVX_HALO_DEBUG: 1 = epilogue does no work, 2 = no MMA issued (timing decomposition only)
};

struct HaloTile {
    int b, p0, q0, r0, n0;
};

__device__ __forceinline__ HaloTile decode_halo_tile(const HaloArgs &a, long long t) {
    HaloTile c;
    c.n0 = (int)(t % a.n_splits) * a.n_tile;
    t /= a.n_splits;
    c.p0 = (int)(t % a.tiles[0]) * HL_P;
    t /= a.tiles[0];
    c.q0 = (int)(t % a.tiles[1]) * (HL_QT * a.mt);
    t /= a.tiles[1];
    c.r0 = (int)(t % a.tiles[2]) * (a.pack2 ? 2 : 1);
    c.b = (int)(t / a.tiles[2]);
    return c;
}

template <bool TF32, bool X3>
__global__ void __launch_bounds__(X3 ? HL_THREADS_X3 : HL_THREADS, 1)
conv3d_tc_halo_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                      const __grid_constant__ CUtensorMap map_y, const __grid_constant__ CUtensorMap map_r,
                      const __grid_constant__ HaloArgs a) {
    constexpr int EK = TF32 ? 32 : 64;      // input channels per chunk = one 128-byte row
    constexpr int SC = TF32 ? 32 : 64;      // output channels per 128-byte-wide slab
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int plane_rows = HL_PPAD * (HL_QT * a.mt + 2);
    const int a_bytes = (plane_rows * 128 + 1023) & ~1023;   // stage pitch; the box itself is plane_rows * 128
    const int b_bytes = 128 * 128;                  // weight stage: 128 channel rows (n_tile of them loaded)
    const int w_bytes = a.n_tile * 128;             // bytes one weight load actually writes
    uint8_t *a_base = smem;
    uint8_t *b_base = a_base + (size_t)a.a_stages * a_bytes;
    uint8_t *slab_base = b_base + (size_t)a.b_stages * b_bytes;   // n_slabs output slabs of 128 x 128 B
    uint64_t *bars = reinterpret_cast<uint64_t *>(slab_base + (size_t)a.n_slabs * 128 * 128);
    uint64_t *a_full = bars, *a_empty = a_full + a.a_stages;
    uint64_t *b_full = a_empty + a.a_stages, *b_empty = b_full + a.b_stages;
    uint64_t *acc_full = b_empty + a.b_stages, *acc_empty = acc_full + 2;
    uint64_t *res_bar = acc_empty + 2;              // residual tile landed in slab set 0 / 1
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(res_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int acc_cols = a.mt * 128;                // columns (voxels) of one accumulator
    const int tmem_cols = 2 * acc_cols <= 32 ? 32 : (2 * acc_cols <= 64 ? 64 : (2 * acc_cols <= 128 ? 128 : (2 * acc_cols <= 256 ? 256 : 512)));

    if (threadIdx.x == 0) {
        for (int s = 0; s < a.a_stages; ++s) mbar_init(smem_u32(a_full + s), 1), mbar_init(smem_u32(a_empty + s), 1);
        for (int s = 0; s < a.b_stages; ++s) mbar_init(smem_u32(b_full + s), 1), mbar_init(smem_u32(b_empty + s), 1);
        for (int i = 0; i < 2; ++i) mbar_init(smem_u32(acc_full + i), 1), mbar_init(smem_u32(acc_empty + i), X3 ? 8 : 4);
        for (int i = 0; i < 2; ++i) mbar_init(smem_u32(res_bar + i), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_y) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_r) : "memory");
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), tmem_cols);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    const int nc = a.c_in / EK;                    // 128-byte rows per voxel (of each half, x3)
    const int k_chunks = a.x3 ? 3 * nc : nc;
    const int n_planes = a.pack2 ? 4 : 3;          // input planes feeding one super-tile
    const int n_groups = k_chunks * n_planes;      // (chunk, plane) groups of 36 MMAs per tile
    const int n_periods = X3 ? (n_groups + a.drain_groups - 1) / a.drain_groups : 1;

    if (warp == 0) {
        // -------- activation planes: (chunk, dR) -> one box of 10 x (16*MT + 2) voxels --------
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
                const HaloTile tc = decode_halo_tile(a, t);
                for (int kc = 0; kc < k_chunks; ++kc)
                    for (int dr = 0; dr < n_planes; ++dr) {
                        mbar_wait(smem_u32(a_empty + stage), phase ^ 1);
                        const uint32_t bar = smem_u32(a_full + stage);
                        mbar_expect_tx(bar, (uint32_t)(plane_rows * 128));
                        // x3: pass 0 reads the lo half of the channels, passes 1 and 2 the hi half
                        const int xch = a.x3 ? (kc < nc ? a.c_in + kc * EK : (kc % nc) * EK) : kc * EK;
                        tma_load_5d(smem_u32(a_base + (size_t)stage * a_bytes), &map_x, bar, xch,
                                    tc.p0 - a.pad[0], tc.q0 - a.pad[1], tc.r0 * a.stride_r - a.pad[2] + dr, tc.b);
                        if (++stage == (uint32_t)a.a_stages) stage = 0, phase ^= 1;
                    }
            }
        }
    } else if (warp == 2) {
        // -------- weight slices: (chunk, dR, dP, dQ) -> [n_tile][64] --------
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
                const HaloTile tc = decode_halo_tile(a, t);
                for (int kc = 0; kc < k_chunks; ++kc)
                    for (int tap = 0; tap < 9 * n_planes; ++tap) {   // (plane, dQ, dP) in MMA order
                        const int dp = tap % 3, dq = (tap / 3) % 3, ip = tap / 9;
                        mbar_wait(smem_u32(b_empty + stage), phase ^ 1);
                        const uint32_t bar = smem_u32(b_full + stage);
                        const uint32_t dst = smem_u32(b_base + (size_t)stage * b_bytes);
                        // weights stay packed [c_out][kX][kY][kZ][c_in]; column of tap (dp, dq, dr):
                        // x3: the row is (hi | lo) x 27 taps x c_in; pass 1 takes the lo half
                        const int wbase = (a.x3 && kc >= nc && kc < 2 * nc) ? 27 * a.c_in : 0;
                        const int wch = (a.x3 ? kc % nc : kc) * EK;
                        auto wcol = [&](int dr) {
                            if (dr < 0 || dr > 2) return (a.x3 ? 54 : 27) * a.c_in;   // beyond the tensor: TMA zero fill
                            int k3[3];                                   // tap along the tensor axes X, Y, Z
                            k3[a.axis[0]] = dp, k3[a.axis[1]] = dq, k3[a.axis[2]] = dr;
                            return wbase + ((k3[0] * 3 + k3[1]) * 3 + k3[2]) * a.c_in + wch;
                        };
                        if (a.pack2) {
                            mbar_expect_tx(bar, 2u * 64u * 128u);
                            tma_load_2d(dst, &map_w, bar, wcol(ip), tc.n0);                 // plane r0:     dR = ip
                            tma_load_2d(dst + 64 * 128, &map_w, bar, wcol(ip - 1), tc.n0);  // plane r0 + 1: dR = ip - 1
                        } else {
                            mbar_expect_tx(bar, (uint32_t)w_bytes);
                            tma_load_2d(dst, &map_w, bar, wcol(ip), tc.n0);
                        }
                        if (++stage == (uint32_t)a.b_stages) stage = 0, phase ^= 1;
                    }
            }
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer ------------------------------
        if (lane == 0) {
            const uint32_t idesc = TF32 ? umma_idesc_tf32(128 * a.mt) : umma_idesc(128 * a.mt);   // M = 128 channel lanes, N = voxels
            uint32_t sa = 0, pa = 0, sb = 0, pb = 0, acc = 0, acc_phase = 0;
            for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
                if (!X3) {
                    mbar_wait(smem_u32(acc_empty + acc), acc_phase ^ 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                uint32_t d_tmem = tmem_base + acc * acc_cols;
                bool first = true;
                int gi = 0;
                for (int kc = 0; kc < k_chunks; ++kc)
                    for (int dr = 0; dr < n_planes; ++dr) {
                        if (X3 && gi % a.drain_groups == 0) {   // a new accumulation period, in the other accumulator
                            mbar_wait(smem_u32(acc_empty + acc), acc_phase ^ 1);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            d_tmem = tmem_base + acc * acc_cols;
                            first = true;
                        }
                        mbar_wait(smem_u32(a_full + sa), pa);
                        const uint32_t plane = smem_u32(a_base + (size_t)sa * a_bytes);
                        for (int dq = 0; dq < 3; ++dq)
                            for (int dp = 0; dp < 3; ++dp) {
                                mbar_wait(smem_u32(b_full + sb), pb);
                                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                                // A operand: the weight slice [channel][64]
                                const uint64_t dw = umma_desc_sw128(smem_u32(b_base + (size_t)sb * b_bytes));
                                // B operand: 16*MT groups of 8 voxels starting at (q = dq, p = dp) of the plane
                                const uint32_t row0 = (uint32_t)(dq * HL_PPAD + dp);
                                const uint64_t dx = umma_desc_sw128(plane + row0 * 128, HL_PPAD * 128);
                                if (a.debug != 2) {
#pragma unroll
                                    for (int k = 0; k < 4; ++k) {
                                        if (TF32) umma_tf32(d_tmem, dw + 2 * k, dx + 2 * k, idesc, !(first && k == 0));
                                        else umma_bf16(d_tmem, dw + 2 * k, dx + 2 * k, idesc, !(first && k == 0));
                                    }
                                }
                                first = false;
                                umma_commit(smem_u32(b_empty + sb));
                                if (++sb == (uint32_t)a.b_stages) sb = 0, pb ^= 1;
                            }
                        umma_commit(smem_u32(a_empty + sa));
                        if (++sa == (uint32_t)a.a_stages) sa = 0, pa ^= 1;
                        ++gi;
                        if (X3 && (gi % a.drain_groups == 0 || gi == n_groups)) {
                            umma_commit(smem_u32(acc_full + acc));
                            if (++acc == 2) acc = 0, acc_phase ^= 1;
                        }
                    }
                if (!X3) {
                    umma_commit(smem_u32(acc_full + acc));
                    if (++acc == 2) acc = 0, acc_phase ^= 1;
                }
            }
        }
    } else if (X3 && warp >= 7) {
        // ---------------- x3 helpers (warps 7..10): upper half of the columns of each period ----------------
        const int q4 = warp & 3;
        const uint32_t half_cols = (uint32_t)acc_cols >> 1;
        uint32_t acc = 0, acc_phase = 0;
        for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
            x3_accumulate_periods(tmem_base, (uint32_t)acc_cols, ((uint32_t)(q4 * 32) << 16) + half_cols,
                                  (int)(half_cols >> 5), 32u, n_periods, acc_full, acc_empty, acc, acc_phase, lane);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            asm volatile("bar.sync 2, 256;" ::: "memory");     // the sum is in TMEM for warps 3..6
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(acc_empty + acc));
            if (++acc == 2) acc = 0, acc_phase ^= 1;
        }
    } else {
        // ------------------------------ epilogue (warps 3..6) ------------------------------
        // Accumulator lane = output channel, column = voxel row of the tile (group g = col >> 3,
        // row in group i = col & 7).  A voxel tile is staged as 64-channel slabs of 128 rows x 128 B
        // (TMA swizzle) and stored by TMA; the residual tile is TMA-loaded into the same slabs
        // first, so adding it is a read-modify-write of shared memory.
        const int q4 = warp & 3;                       // TMEM lane quarter of this warp
        const int ch_local = q4 * 32 + lane;           // channel inside the n-tile
        const bool ch_ok = a.pack2 || ch_local < a.n_tile;
        const bool epi_leader = (warp == 3 && lane == 0);
        const bool has_res = a.res != nullptr;
        // a voxel tile is stored as SC-channel slabs (128 bytes per voxel row): the channel groups of a
        // 128-channel tile, or (pack2) those of the 64 channels of output plane r0 and then of plane r0 + 1
        constexpr int ES = TF32 ? 4 : 2;               // bytes per output element
        const int slabs_per_tile = a.pack2 ? 128 / SC : a.n_tile / SC;
        const int slabs_per_plane = 64 / SC;           // pack2: slabs of one output plane
        const int slab_sets = a.n_slabs / slabs_per_tile;
        const int my_slab = ch_local / SC, my_chunk = ((ch_local % SC) * ES) >> 4, my_byte = ((ch_local % SC) * ES) & 15;
        uint32_t xoff[8];                              // byte offset inside a slab row, by (row & 7)
#pragma unroll
        for (int c = 0; c < 8; ++c) xoff[c] = (uint32_t)(((my_chunk ^ c) << 4) + my_byte);
        uint32_t acc = 0, acc_phase = 0, res_phase[2] = {0, 0};
        int set = 0;

        auto slab_free = [&]() {   // leader: the TMA stores that last read slab set `set` are done with it
            if (slab_sets > 1) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        };
        // leader only.  `pending` = TMA stores that may still be reading OTHER slab sets
        auto load_residual = [&](const HaloTile &tc, int mt, int into, int pending) {
            if (pending >= 1) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            const uint32_t bar = smem_u32(res_bar + into);
            mbar_expect_tx(bar, (uint32_t)(slabs_per_tile * 128 * 128));
            for (int sl = 0; sl < slabs_per_tile; ++sl)
                tma_load_5d(smem_u32(slab_base + (size_t)(into * slabs_per_tile + sl) * (128 * 128)), &map_r, bar,
                            tc.n0 + (a.pack2 ? (sl % slabs_per_plane) * SC : sl * SC), tc.p0, tc.q0 + mt * HL_QT,
                            tc.r0 + (a.pack2 ? sl / slabs_per_plane : 0), tc.b);
        };

        for (long long t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
            const HaloTile tc = decode_halo_tile(a, t);
            const int ch = tc.n0 + (a.pack2 ? (ch_local & 63) : ch_local);   // (pack2: my_slab already selects the plane)
            const float sc = (a.scale && ch_ok) ? __ldg(a.scale + ch) : 1.f;
            const float sh = (a.shift && ch_ok) ? __ldg(a.shift + ch) : 0.f;
            // residual tiles are fetched while the accumulator is still being produced: all of the
            // super-tile's when every tile has a slab set of its own, else only the first
            const bool res_ahead = slab_sets >= a.mt;
            if (has_res && epi_leader) {
                load_residual(tc, 0, set, slab_sets - 1);
                if (res_ahead)
                    for (int mt = 1; mt < a.mt; ++mt) load_residual(tc, mt, (set + mt) % slab_sets, slab_sets - 1 - mt);
            }
            if (X3) {   // sum the accumulation periods (lower half of the columns; warps 7..10 take the upper)
                x3_accumulate_periods(tmem_base, (uint32_t)acc_cols, (uint32_t)(q4 * 32) << 16, acc_cols >> 6, 32u,
                                      n_periods, acc_full, acc_empty, acc, acc_phase, lane);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                asm volatile("bar.sync 2, 256;" ::: "memory");
            } else {
                mbar_wait(smem_u32(acc_full + acc), acc_phase);
            }
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int mt = 0; mt < (a.debug == 1 ? 0 : a.mt); ++mt) {
                if (has_res) {
                    if (mt > 0 && !res_ahead && epi_leader) load_residual(tc, mt, set, slab_sets - 1);
                    mbar_wait(smem_u32(res_bar + set), res_phase[set]);
                    res_phase[set] ^= 1;
                } else {
                    if (epi_leader) slab_free();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
                uint8_t *my_rows = slab_base + (size_t)(set * slabs_per_tile + my_slab) * (128 * 128);
                for (int v0 = 0; v0 < 128; v0 += 32) {
                    uint32_t v[32];
                    tmem_ld32(tmem_base + acc * acc_cols + mt * 128 + ((uint32_t)(q4 * 32) << 16) + v0, v);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (ch_ok) {
                        uint8_t *rows = my_rows + v0 * 128;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            using Cell = typename std::conditional<TF32, float, __nv_bfloat16>::type;
                            Cell *cell = reinterpret_cast<Cell *>(rows + j * 128 + xoff[j & 7]);
                            float f = fmaf(__uint_as_float(v[j]), sc, sh);
                            if (has_res) {
                                const float r = (float)*cell;
                                if (!a.residual_after_act) f += r;
                                if (a.relu) f = fmaxf(f, 0.f);
                                if (a.residual_after_act) f += r;
                            } else if (a.relu) {
                                f = fmaxf(f, 0.f);
                            }
                            *cell = (Cell)(f * a.post_scale);
                        }
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (epi_leader) {
                    for (int sl = 0; sl < slabs_per_tile; ++sl)
                        tma_store_5d(&map_y, smem_u32(slab_base + (size_t)(set * slabs_per_tile + sl) * (128 * 128)),
                                     tc.n0 + (a.pack2 ? (sl % slabs_per_plane) * SC : sl * SC), tc.p0,
                                     tc.q0 + mt * HL_QT, tc.r0 + (a.pack2 ? sl / slabs_per_plane : 0), tc.b);
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
                if (++set == slab_sets) set = 0;
            }
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

// Decide whether (and how) the halo kernel takes a layer.  perm[role] = tensor axis (0=X,1=Y,2=Z)
// playing role P, Q, R.  Returns the fraction of accumulator rows that hold real outputs, or 0.
double plan_halo(const ivx_conv3d_params *p, int perm[3], int *mt_out, double *cost_out) {
    const int ek = p->in_dtype == IVX_BF16 ? 64 : 32;    // channels per 128-byte row: bf16 / tf32
    if (cost_out) *cost_out = 0.0;
    if (p->transposed || p->c_in % ek != 0 || p->c_out % 64 != 0) return 0.0;
    int strided_axis = -1;
    for (int d = 0; d < 3; ++d) {
        if (p->kernel[d] != 3 || p->pad[d] > 1) return 0.0;
        if (p->stride[d] != 1) {
            if (strided_axis >= 0) return 0.0;   // a stride is only supported along R (one axis)
            strided_axis = d;
        }
    }
    // Cost of a plan in MMA issue slots: every CTA tile (8 x 16*mt voxels of one plane -- of two planes for
    // 64-channel layers -- x 128 accumulator lanes) takes the same number of M128 MMAs whatever mt is (one
    // M = 128 MMA costs ~128 cycles for N = 128 and for N = 256), the persistent grid runs
    // ceil(tiles / 148) of them per CTA, plus a fixed pipeline fill + last epilogue.  Wave quantisation
    // matters for single scenes (200 tiles = 2 waves).  Ties: fewer wasted rows, then P along Z.
    static const bool by_score = [] {
        const char *e = getenv("IVX_HALO_PLAN");
        return e != nullptr && e[0] == 's';
    }();
    const int n_tile = p->c_out % 128 == 0 ? 128 : 64;
    const int n_splits = p->c_out / n_tile;
    const int chunks = (p->c_in / ek) * (p->in_dtype == IVX_F32X2 ? 3 : 1);
    double best = 0.0, best_cost = 0.0, best_eff = 0.0, min_cost = 1e300;
    // pass 0 finds the cheapest plan, pass 1 picks among the plans within 3 % of it by the rule the large
    // grids were tuned with (few wasted rows, two tiles per CTA, P along the contiguous Z axis)
    for (int pass = by_score ? 1 : 0; pass < 2; ++pass)
    for (int P = 0; P < 3; ++P)
        for (int Q = 0; Q < 3; ++Q) {
            if (Q == P) continue;
            const int R = 3 - P - Q;
            if (strided_axis >= 0 && R != strided_axis) continue;
            for (int mt = 1; mt <= 2; ++mt) {
                const long long tp = (p->out_dims[P] + HL_P - 1) / HL_P;
                const long long tq = (p->out_dims[Q] + HL_QT * mt - 1) / (HL_QT * mt);
                double eff = (double)p->out_dims[P] * p->out_dims[Q] / (double)(tp * HL_P * tq * HL_QT * mt);
                const bool pack2 = n_tile == 64 && p->stride[R] == 1;
                const long long tr = pack2 ? (p->out_dims[R] + 1) / 2 : p->out_dims[R];
                const long long tiles = (long long)p->batch * tp * tq * tr * n_splits;
                const double per_tile = (pack2 ? 36.0 : 27.0) * chunks * 4.0;      // MMAs of one CTA tile
                const double cost = (double)((tiles + kNumSMs - 1) / kNumSMs) * per_tile + 50.0;
                // old rule (IVX_HALO_PLAN=score): sharing a weight slice between two tiles is worth a few
                // wasted rows; P along the contiguous Z axis keeps the TMA rows contiguous in memory
                if (pass == 0) {
                    if (cost < min_cost) min_cost = cost;
                    continue;
                }
                if (!by_score && cost > 1.03 * min_cost) continue;
                const double score = eff * (mt == 2 ? 1.12 : 1.0) * (P == 2 ? 1.02 : 1.0);
                if (score > best) {
                    best = score, best_cost = cost, best_eff = eff;
                    perm[0] = P, perm[1] = Q, perm[2] = R;
                    *mt_out = mt;
                }
            }
        }
    if (best == 0.0) return 0.0;
    if (cost_out) *cost_out = best_cost;
    return best_eff;   // fraction of accumulator rows that hold real outputs
}

template <bool TF32, bool X3 = false>
static int launch_halo(const ivx_conv3d_params *p, const int perm[3], int mt, const void *x,
                       const void *weight_bf16, const float *scale, const float *shift,
                       const void *residual, void *y, cudaStream_t stream) {
    constexpr int EK = TF32 ? 32 : 64, SC = TF32 ? 32 : 64, ES = TF32 ? 4 : 2;
    const CUtensorMapDataType dt_in = TF32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const CUtensorMapDataType dt_out = TF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    HaloArgs a;
    a.c_in = p->c_in, a.c_out = p->c_out;
    a.x3 = X3 ? 1 : 0;
    {
        const char *e = getenv("IVX_X3_DRAIN_GROUPS");   // experiments: a huge value = one period per tile
        a.drain_groups = e ? atoi(e) : 2;
        if (a.drain_groups < 1) a.drain_groups = 1;
    }
    const int cin_mem = a.x3 ? 2 * p->c_in : p->c_in;    // channels per voxel in memory
    const int w_halves = a.x3 ? 2 : 1;
    a.n_tile = p->c_out % 128 == 0 ? 128 : 64;
    a.n_splits = p->c_out / a.n_tile;
    a.mt = mt;
    a.batch = p->batch;
    // element strides of the channels-last tensors along X, Y, Z
    const long long in_s[3] = {(long long)p->in_dims[1] * p->in_dims[2] * cin_mem,
                               (long long)p->in_dims[2] * cin_mem, (long long)cin_mem};
    const long long out_s[3] = {(long long)p->out_dims[1] * p->out_dims[2] * p->c_out,
                                (long long)p->out_dims[2] * p->c_out, (long long)p->c_out};
    for (int r = 0; r < 3; ++r) {
        a.dims[r] = p->out_dims[perm[r]];
        a.pad[r] = p->pad[perm[r]];
        a.axis[r] = perm[r];
        a.ostride[r] = out_s[perm[r]];
    }
    a.ostride_b = out_s[0] * p->out_dims[0];
    a.tiles[0] = (a.dims[0] + HL_P - 1) / HL_P;
    a.tiles[1] = (a.dims[1] + HL_QT * mt - 1) / (HL_QT * mt);
    a.stride_r = p->stride[perm[2]];
    a.pack2 = (a.n_tile == 64 && a.stride_r == 1) ? 1 : 0;
    a.tiles[2] = a.pack2 ? (a.dims[2] + 1) / 2 : a.dims[2];
    a.n_tiles = (long long)p->batch * a.tiles[0] * a.tiles[1] * a.tiles[2] * a.n_splits;
    a.relu = p->relu, a.residual_after_act = p->residual_after_act, a.post_scale = p->post_scale;
    a.scale = scale, a.shift = shift, a.res = residual;
    {
        const char *e = getenv("IVX_HALO_DEBUG");
        a.debug = e ? atoi(e) : 0;
    }
    const int plane_rows = HL_PPAD * (HL_QT * mt + 2);
    const int a_bytes = (plane_rows * 128 + 1023) & ~1023, b_bytes = 128 * 128;
    const int slabs_per_tile = a.pack2 ? 128 / SC : a.n_tile / SC;
    // shared memory: weight ring (16 KB stages), two sets of output slabs if they fit, the
    // rest goes to activation planes (each one feeds 36 MMAs)
    a.b_stages = 4;
    a.n_slabs = 2 * slabs_per_tile;
    const int budget = 227 * 1024 - 2048;
    a.a_stages = (budget - a.b_stages * b_bytes - a.n_slabs * 16384) / a_bytes;
    if (a.a_stages < 3 && slabs_per_tile >= 2) {   // wide tiles: one slab set buys plane stages
        a.n_slabs = slabs_per_tile;
        a.a_stages = (budget - a.b_stages * b_bytes - a.n_slabs * 16384) / a_bytes;
    }
    if (a.a_stages < 2) {                          // fp32 slabs are large: a shorter weight ring
        a.b_stages = 3;
        a.a_stages = (budget - a.b_stages * b_bytes - a.n_slabs * 16384) / a_bytes;
    }
    if (a.a_stages > 4) a.a_stages = 4;
    if (a.a_stages < 2) return fail(IVX_ERR_UNSUPPORTED, "halo kernel: not enough shared memory");
    const size_t smem = (size_t)a.a_stages * a_bytes + (size_t)a.b_stages * b_bytes +
                       (size_t)a.n_slabs * 16384 + 2048;

    alignas(64) CUtensorMap map_x, map_w, map_y, map_r;
    {
        const cuuint64_t dims[5] = {(cuuint64_t)cin_mem, (cuuint64_t)p->in_dims[perm[0]],
                                    (cuuint64_t)p->in_dims[perm[1]], (cuuint64_t)p->in_dims[perm[2]],
                                    (cuuint64_t)p->batch};
        const cuuint64_t st[4] = {(cuuint64_t)in_s[perm[0]] * ES, (cuuint64_t)in_s[perm[1]] * ES,
                                  (cuuint64_t)in_s[perm[2]] * ES, (cuuint64_t)in_s[0] * p->in_dims[0] * ES};
        const cuuint32_t box[5] = {EK, HL_PPAD, (cuuint32_t)(HL_QT * mt + 2), 1, 1};
        const cuuint32_t es[5] = {1, 1, 1, 1, 1};
        int rc = encode_map_sw128(&map_x, dt_in, 5, const_cast<void *>(x), dims, st, box, es, "activations (halo)");
        if (rc != IVX_OK) return rc;
    }
    {
        const cuuint64_t dims[2] = {(cuuint64_t)27 * w_halves * p->c_in, (cuuint64_t)p->c_out};
        const cuuint64_t st[1] = {(cuuint64_t)27 * w_halves * p->c_in * ES};
        const cuuint32_t box[2] = {EK, (cuuint32_t)a.n_tile};   // pack2: two such loads fill the 128 rows
        const cuuint32_t es[2] = {1, 1};
        int rc = encode_map_sw128(&map_w, dt_in, 2, const_cast<void *>(weight_bf16), dims, st, box, es, "weights (halo)");
        if (rc != IVX_OK) return rc;
    }
    {
        const cuuint64_t dims[5] = {(cuuint64_t)p->c_out, (cuuint64_t)a.dims[0], (cuuint64_t)a.dims[1],
                                    (cuuint64_t)a.dims[2], (cuuint64_t)p->batch};
        const cuuint64_t st[4] = {(cuuint64_t)a.ostride[0] * ES, (cuuint64_t)a.ostride[1] * ES,
                                  (cuuint64_t)a.ostride[2] * ES, (cuuint64_t)a.ostride_b * ES};
        const cuuint32_t box[5] = {SC, HL_P, HL_QT, 1, 1};
        const cuuint32_t es[5] = {1, 1, 1, 1, 1};
        int rc = encode_map_sw128(&map_y, dt_out, 5, y, dims, st, box, es, "output (halo)");
        if (rc != IVX_OK) return rc;
        map_r = map_y;
        if (residual != nullptr) {
            rc = encode_map_sw128(&map_r, dt_out, 5, const_cast<void *>(residual), dims, st, box, es, "residual (halo)");
            if (rc != IVX_OK) return rc;
        }
    }
    int rc = check_cuda(cudaFuncSetAttribute(conv3d_tc_halo_kernel<TF32, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem),
                        "cudaFuncSetAttribute");
    if (rc != IVX_OK) return rc;
    const unsigned grid = (unsigned)(a.n_tiles < kNumSMs ? a.n_tiles : kNumSMs);
    conv3d_tc_halo_kernel<TF32, X3><<<grid, X3 ? HL_THREADS_X3 : HL_THREADS, smem, stream>>>(map_x, map_w, map_y, map_r, a);
    IVX_CHECK_LAUNCH("conv3d_tc_halo launch");
    return IVX_OK;
}

int launch_conv3d_tc_halo(const ivx_conv3d_params *p, const int perm[3], int mt, const void *x,
                          const void *weight, const float *scale, const float *shift,
                          const void *residual, void *y, cudaStream_t stream) {
    if (p->in_dtype == IVX_F32X2)
        return launch_halo<true, true>(p, perm, mt, x, weight, scale, shift, residual, y, stream);
    if (p->in_dtype != IVX_BF16)
        return launch_halo<true>(p, perm, mt, x, weight, scale, shift, residual, y, stream);
    return launch_halo<false>(p, perm, mt, x, weight, scale, shift, residual, y, stream);
}

}  // namespace ivx
