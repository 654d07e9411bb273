// b2c_ptx.cuh — thin inline-PTX wrappers for the sm_100a features the conv kernels use:
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld) and the
// shared-memory matrix / instruction descriptors that tcgen05.mma consumes.
//
// Nothing here allocates or launches; every function is __device__ __forceinline__.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace b2c {

// ----------------------------------------------------------------------------------------
// error reporting from device code: a bounded mbarrier wait records why it gave up and traps,
// so a wrong transaction count or phase shows up as a launch failure instead of a hung GPU.
// ----------------------------------------------------------------------------------------
static __device__ unsigned int g_b2c_timeout_info[4];   // {tag, blockIdx.x, threadIdx.x, parity}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// One lane of the (fully active, converged) warp gets `true`. Issue tcgen05.mma / TMA / commit under this
// predicate with the WHOLE warp running the surrounding loop: ptxas then emits the uniform-datapath
// instruction directly. Guarding with `lane == 0` instead makes the code divergent and every UTCHMMA /
// UTMALDG gets wrapped in an ELECT/branch loop (~70-90 cycles per issue, measured: tools/tma_bw_probe.cu).
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
        "elect.sync rx|px, 0xffffffff;\n\t"
        "selp.b32 %0, 1, 0, px;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---------------------------------- mbarrier ---------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok;
}
// Bounded wait: ~2 s of SM clock, then record (tag, block, thread, parity) and trap.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, uint32_t tag) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (((++spins) & 0x3ffu) == 0 && (clock64() - t0) > 4000000000ll) {
            g_b2c_timeout_info[0] = tag;
            g_b2c_timeout_info[1] = blockIdx.x;
            g_b2c_timeout_info[2] = threadIdx.x;
            g_b2c_timeout_info[3] = parity;
            __threadfence_system();
            __trap();
        }
    }
}

// ---------------------------------- TMA --------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                            int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

__device__ __forceinline__ void tma_load_5d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                            int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

// ---------------------------------- programmatic dependent launch (b2c_launch.h) ----------
// pdl_launch_dependents(): the next kernel in the stream may be scheduled once every CTA of this grid has got here.
// pdl_wait(): blocks until the previous kernel in the stream has completed and its global writes are visible;
// nothing before it may touch global memory that another kernel writes. Both are no-ops without the launch attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Which kernel classes release their successor early (compile-time policy): tensor-core kernels (BIG), full-tensor
// element-wise passes (EW), latency-bound kernels of a few CTAs (TINY). Measured on one box, ms per step at batch
// 1024 / 128 per GPU: no PDL 5.389 / 1.375; all classes 5.506 / 1.354; TINY only 5.370 / 1.354; BIG+TINY 5.449 / 1.364;
// EW+TINY 5.387 / 1.336. Early successors of the tensor-core kernels take SM resources from CTAs that still have
// work, so those keep the implicit trigger at exit.
#ifndef B2C_PDL_BIG
#define B2C_PDL_BIG 0
#endif
#ifndef B2C_PDL_EW
#define B2C_PDL_EW 1
#endif
#ifndef B2C_PDL_TINY
#define B2C_PDL_TINY 1
#endif
__device__ __forceinline__ void pdl_trigger_big() { if (B2C_PDL_BIG) pdl_launch_dependents(); }
__device__ __forceinline__ void pdl_trigger_ew() { if (B2C_PDL_EW) pdl_launch_dependents(); }
__device__ __forceinline__ void pdl_trigger_tiny() { if (B2C_PDL_TINY) pdl_launch_dependents(); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---------------------------------- tcgen05 ----------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Whole warp executes. Writes the TMEM base address (lane 0, column c) to *dst_smem.
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem) {
    static_assert(kCols == 32 || kCols == 64 || kCols == 128 || kCols == 256 || kCols == 512, "pow2 >= 32");
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "n"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc]; one thread issues for the CTA.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets lane (lane_base + t).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
          "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
          "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// ---------------------------------- descriptors ------------------------------------------
// Shared-memory matrix descriptor (sm_100 "version 1"), SWIZZLE_128B.
//   K-major  operand: rows of 64 bf16 (128 B); 8-row groups `sbo` bytes apart (1024 when dense).
//   MN-major operand: each K index is one 128 B row of 64 MN elements; 8-row K groups `sbo`
//                     bytes apart; successive 64-element MN blocks `lbo` bytes apart.
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes,
                                                         uint32_t sbo_bytes, uint32_t base_offset = 0) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3ffffu) >> 4);            // [0,14)  start address
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;       // [16,30) leading byte offset
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;       // [32,46) stride byte offset
    d |= static_cast<uint64_t>(1) << 46;                                // [46,48) version = 1
    d |= static_cast<uint64_t>(base_offset & 7u) << 49;                 // [49,52) base offset
    d |= static_cast<uint64_t>(2) << 61;                                // [61,64) SWIZZLE_128B
    return d;
}

// Instruction descriptor for kind::f16, bf16 x bf16 -> fp32, dense.
//   a_mn / b_mn: 1 = operand is MN-major ("transposed"), 0 = K-major.
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, int a_mn, int b_mn) {
    return (1u << 4)                                   // c_format = F32
           | (1u << 7)                                 // a_format = BF16
           | (1u << 10)                                // b_format = BF16
           | (static_cast<uint32_t>(a_mn) << 15)       // a_major
           | (static_cast<uint32_t>(b_mn) << 16)       // b_major
           | (static_cast<uint32_t>(N >> 3) << 17)     // n_dim
           | (static_cast<uint32_t>(M >> 4) << 24);    // m_dim
}

// ---------------------------------- small helpers ----------------------------------------
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Column sums across the 32 lanes of a warp for 32 per-lane values, 31 shuffles:
// on return lane l holds sum over lanes of v[l] (the "transpose-reduce" butterfly).
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const bool up = lane & 16;
        const float send = up ? v[i] : v[i + 16];
        const float keep = up ? v[i + 16] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool up = lane & 8;
        const float send = up ? v[i] : v[i + 8];
        const float keep = up ? v[i + 8] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const bool up = lane & 4;
        const float send = up ? v[i] : v[i + 4];
        const float keep = up ? v[i + 4] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const bool up = lane & 2;
        const float send = up ? v[i] : v[i + 2];
        const float keep = up ? v[i + 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    {
        const bool up = lane & 1;
        const float send = up ? v[0] : v[1];
        const float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
    }
    return v[0];
}

// Epilogue global accesses of one warp's 32 (rows = lanes) x 32 (channels) bf16 chunk. A direct access has every
// lane touching 16 bytes of its own output row: 32 lines per instruction, and the load/store queue backs up into
// the register file (ncu: the top epilogue stall was the WAR wait on the STG source registers). Instead the chunk
// goes through a 2 KB per-warp shared-memory buffer: global instructions cover 8 rows x 64 contiguous bytes each,
// lanes read / write their own row in shared memory. The 16-byte pieces are XOR-swizzled by (row >> 1) so both
// the row-wise and the 8-row accesses are bank-conflict free.
//   rowoff[it] = element offset of row (it*8 + lane/4) in the tensor, ~0 for rows past the end (chunk_rowoffs)
//   col        = first channel of the chunk
// All 32 lanes must call.
__device__ __forceinline__ void chunk_rowoffs(size_t off, bool valid, int lane, unsigned long long (&rowoff)[4]) {
    const unsigned long long mine = valid ? static_cast<unsigned long long>(off) : ~0ull;
#pragma unroll
    for (int it = 0; it < 4; ++it) rowoff[it] = __shfl_sync(0xffffffffu, mine, it * 8 + (lane >> 2));
}

__device__ __forceinline__ void store_chunk_staged(uint8_t* stg, __nv_bfloat16* out,
                                                   const unsigned long long (&rowoff)[4], int col,
                                                   const uint4 (&u)[4], int lane) {
    __syncwarp();                                   // earlier reads of the buffer are done
    uint4* row = reinterpret_cast<uint4*>(stg + lane * 64);
    const int sw = (lane >> 1) & 3;
#pragma unroll
    for (int g = 0; g < 4; ++g) row[g ^ sw] = u[g];
    __syncwarp();
    const int piece = lane & 3;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const int r = it * 8 + (lane >> 2);
        const uint4 val = *reinterpret_cast<const uint4*>(stg + r * 64 + ((piece ^ ((r >> 1) & 3)) << 4));
        if (rowoff[it] != ~0ull) *reinterpret_cast<uint4*>(out + rowoff[it] + col + piece * 8) = val;
    }
}

// The reverse: u[4] = this lane's row of a bf16 tensor chunk (zeros for rows past the end).
__device__ __forceinline__ void load_chunk_staged(uint8_t* stg, const __nv_bfloat16* src,
                                                  const unsigned long long (&rowoff)[4], int col, uint4 (&u)[4],
                                                  int lane) {
    const int piece = lane & 3;
    uint4 val[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        val[it] = make_uint4(0u, 0u, 0u, 0u);
        if (rowoff[it] != ~0ull) val[it] = __ldg(reinterpret_cast<const uint4*>(src + rowoff[it] + col + piece * 8));
    }
    __syncwarp();                                   // earlier reads of the buffer are done
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const int r = it * 8 + (lane >> 2);
        *reinterpret_cast<uint4*>(stg + r * 64 + ((piece ^ ((r >> 1) & 3)) << 4)) = val[it];
    }
    __syncwarp();
    const uint4* row = reinterpret_cast<const uint4*>(stg + lane * 64);
    const int sw = (lane >> 1) & 3;
#pragma unroll
    for (int g = 0; g < 4; ++g) u[g] = row[g ^ sw];
}

}  // namespace b2c
