// Inline-PTX wrappers shared by the tcgen05 kernels: mbarrier, TMA (cp.async.bulk.tensor),
// TMEM allocation, tcgen05.mma / commit / ld, and the tensor-map encoder entry point.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>

#include "ivx_common.cuh"

namespace ivx {

// ------------------------------ PTX wrappers ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}

__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0,
                                            int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0,
                                            int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

__device__ __forceinline__ void tma_store_5d(const CUtensorMap *map, uint32_t src, int c0, int c1, int c2,
                                             int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];"
        ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// fp32 operands read as tf32 (the low 13 mantissa bits are ignored), K = 8 per instruction
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          bool accumulate) {
    const uint32_t acc = accumulate ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
}

// arrives on the mbarrier once every tcgen05.mma issued so far has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                 : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
          "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
          "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
          "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

// registers -> TMEM: 32 lanes x 32 consecutive columns (the inverse of tmem_ld32)
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
          "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
          "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
          "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}

// 3xTF32 accumulation (fp32-grade layers).  The tensor core adds each MMA's products into the TMEM
// accumulator with truncation, so thousands of MMAs into ONE accumulator leave an error that grows
// linearly with K (measured: 1.6e-6 of max|out| at K = 1728, 1.1e-5 at K = 13 824).  The x3 kernels
// therefore give an accumulator only a short run of MMAs (a "period"), alternate between the two
// TMEM accumulators, and the epilogue warps add every finished period into fp32 registers
// (round-to-nearest) -- `master`, 128 values per thread, this warp's 32 TMEM lanes x up to 128 of the
// columns in 32-column chunks `stride` apart.  After the last period the sum is written back over the
// accumulator it came from and the ordinary fused epilogue runs on it.
//   n_chunks        32-column chunks of this warp (<= 4); chunk i starts at column col0 + i * stride
//   acc, acc_phase  accumulator / mbarrier parity bookkeeping, shared with the rest of the epilogue:
//                   on return `acc` is the buffer that holds the sum and is NOT yet released
__device__ __forceinline__ void x3_accumulate_periods(uint32_t tmem_base, uint32_t acc_cols_stride, uint32_t lane_col0,
                                                      int n_chunks, uint32_t stride, int n_periods,
                                                      uint64_t *acc_full, uint64_t *acc_empty, uint32_t &acc,
                                                      uint32_t &acc_phase, int lane) {
    float master[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) master[i] = 0.f;
    for (int per = 0; per < n_periods; ++per) {
        mbar_wait(smem_u32(acc_full + acc), acc_phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t base = tmem_base + acc * acc_cols_stride + lane_col0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i < n_chunks) {
                uint32_t v[32];
                tmem_ld32(base + (uint32_t)i * stride, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j) master[i * 32 + j] += __uint_as_float(v[j]);
            }
        }
        if (per + 1 < n_periods) {   // hand the accumulator back; the last one is kept for the epilogue
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(acc_empty + acc));
            if (++acc == 2) acc = 0, acc_phase ^= 1;
        }
    }
    const uint32_t base = tmem_base + acc * acc_cols_stride + lane_col0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (i < n_chunks) {
            uint32_t v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(master[i * 32 + j]);
            tmem_st32(base + (uint32_t)i * stride, v);
        }
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// bf16 x bf16 -> fp32, both operands K-major, M = 128
__device__ __forceinline__ uint32_t umma_idesc(int n) {
    return (1u << 4)                      // accumulator format F32
           | (1u << 7) | (1u << 10)       // A, B format BF16
           | ((uint32_t)(n >> 3) << 17)   // N
           | ((uint32_t)(128 >> 4) << 24);
}

// tf32 x tf32 -> fp32, both operands K-major, M = 128
__device__ __forceinline__ uint32_t umma_idesc_tf32(int n) {
    return (1u << 4)                      // accumulator format F32
           | (2u << 7) | (2u << 10)       // A, B format TF32
           | ((uint32_t)(n >> 3) << 17)   // N
           | ((uint32_t)(128 >> 4) << 24);
}

// K-major operand of 128-byte rows with the 128B swizzle.  `sbo_bytes` is the distance between
// consecutive 8-row groups.  The start address may be any 128-byte row of a box the TMA wrote
// with the same swizzle: measured on B200, the MMA unit applies the XOR pattern to the absolute
// shared-memory address (as the TMA does), so a row-shifted start needs no base offset
// (setting the base-offset field to (addr >> 7) & 7 gives wrong data; tools/halo_debug.py).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr, uint32_t sbo_bytes = 1024) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);       // start address, 16-byte units
    d |= (uint64_t)1 << 16;                           // leading byte offset (unused: swizzled K-major)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32; // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                           // SWIZZLE_128B
    return d;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

// bf16 tensor map with the 128B swizzle (cuTensorMapEncodeTiled, resolved through the runtime)
int encode_map_bf16(CUtensorMap *map, int rank, void *base, const cuuint64_t *dims,
                    const cuuint64_t *strides_bytes, const cuuint32_t *box, const cuuint32_t *estrides,
                    const char *what);
// the same for any element type (fp32 tensors: FLOAT32, or TFLOAT32 for MMA operands)
int encode_map_sw128(CUtensorMap *map, CUtensorMapDataType dtype, int rank, void *base, const cuuint64_t *dims,
                     const cuuint64_t *strides_bytes, const cuuint32_t *box, const cuuint32_t *estrides,
                     const char *what);

}  // namespace ivx
