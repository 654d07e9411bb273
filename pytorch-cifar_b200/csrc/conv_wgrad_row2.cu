// conv_wgrad_row2.cu — conv_wgrad_row.cu on a CTA PAIR (tcgen05.mma.cta_group::2, M = 256).
//
// Same GEMM, boxes and epilogue as conv_wgrad_row.cu (reference site: autograd of nn.Conv2d, main.py:104; conv
// sites models/resnet.py:23-27). The two CTAs of a cluster walk the SAME pixel range and cout slab (K and N are
// shared), and each owns one "unit" of D rows = (kernel row a, 128-channel cin pair): three 128x128 blocks per CTA.
// Each CTA loads its own two haloed X boxes and ONE 64-channel slab of the dY tile (its half of N = 128); the TMA
// loads (cp.async.bulk.tensor...cta_group::2) signal the leader's barrier, the leader issues M = 256 MMAs and frees
// stages with a multicast commit. With an odd number of units the last pair's peer repeats the leader's unit for its
// A rows (its half of dY is still needed) and skips the epilogue.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kXSlab = 10240;                 // one cin slab's box: <= 80 pixels x 128 B
constexpr int kDYSlab = 8192;                 // 64 pixels x 64 channels (this CTA's half of the dY tile)
constexpr int kStageBytes = 2 * kXSlab + kDYSlab;   // 28 KB
constexpr int kStages = 7;
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
constexpr int kTmemCols = 512;

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_addr(const void* p, uint32_t rank) {
    uint32_t raddr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(p)), "r"(rank));
    return raddr;
}
__device__ __forceinline__ void tma2_load_4d(uint32_t dst_cluster, const CUtensorMap* m, uint32_t bar_cluster, int c0,
                                             int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst_cluster),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(static_cast<uint16_t>(3))
        : "memory");
}

// W16: true = 16x16 images (k-step = 4 rows x 16), false = 8x8 images (k-step = 8 rows x 8)
template <bool W16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
wgrad_row2_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapDY,
                  const __grid_constant__ WgradRowParams p) {
    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* aux = smem + kStages * kStageBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty = full + kStages;
    uint64_t* tfull = empty + kStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull + 1);

    constexpr int kRows = W16 ? 4 : 8;
    constexpr int kW = W16 ? 16 : 8;
    constexpr uint32_t kXBytes = static_cast<uint32_t>(kRows * (kW + 2) * 128);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();

    // pair index = (unit pair, cout slab, split); unit = (kernel row a, cin pair)
    const int units = 3 * p.cin_pairs;
    const int pi = blockIdx.x >> 1;
    const int up = pi / (p.cout_slabs * p.splits);
    const int rem_p = pi - up * p.cout_slabs * p.splits;
    const int co = rem_p / p.splits;
    const int split = rem_p - co * p.splits;
    int unit = 2 * up + static_cast<int>(rank);
    const bool dummy = unit >= units;          // odd unit count: the last pair's peer only supplies its half of dY
    if (dummy) unit = 2 * up;
    const int a = unit / p.cin_pairs;
    const int cp = unit - a * p.cin_pairs;
    const int t_begin = split * p.steps_per_split;
    int t_end = t_begin + p.steps_per_split;
    if (t_end > p.num_steps) t_end = p.num_steps;
    const int n_iters = t_end - t_begin;     // host guarantees >= 1

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapX);
        tma_prefetch_desc(&mapDY);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < kStages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
            }
            mbar_init(&tfull[0], 1);
            fence_mbar_init();
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)),
                     "n"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();

    if (warp == 0) {
        // ------------------------------ TMA producer (both CTAs; bytes signalled on the leader's barrier) ----
        for (int it = 0; it < n_iters; ++it) {
            const int stage = it % kStages;
            const uint32_t phase = static_cast<uint32_t>((it / kStages) & 1);
            const int t = t_begin + it;
            const int n = W16 ? (t >> 2) : t;
            const int h0 = W16 ? (t & 3) * 4 : 0;
            mbar_wait(&empty[stage], phase ^ 1u, 61);
            if (elect_one_sync()) {
                uint8_t* sx = smem + stage * kStageBytes;
                if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * (2 * kXBytes + kDYSlab));
                const uint32_t bar = cluster_addr(&full[stage], 0);
                tma2_load_4d(cluster_addr(sx, rank), &mapX, bar, cp * 128, -1, h0 + a - 1, n);
                tma2_load_4d(cluster_addr(sx + kXSlab, rank), &mapX, bar, cp * 128 + 64, -1, h0 + a - 1, n);
                tma2_load_4d(cluster_addr(sx + 2 * kXSlab, rank), &mapDY, bar, co * 128 + static_cast<int>(rank) * 64, 0,
                             h0, n);
            }
            __syncwarp();
        }
    } else if (warp == 1 && rank == 0) {
        // ------------------------------ leader: MMA issuer for the pair (M = 256, N = 128) ------------------
        constexpr uint32_t idesc = make_idesc_bf16(256, 128, 1, 1);   // both operands MN-major
        constexpr uint32_t kSboA = W16 ? 1024u : 1280u;
        constexpr uint32_t kStepA = W16 ? (18u * 128u) >> 4 : (20u * 128u) >> 4;
        for (int it = 0; it < n_iters; ++it) {
            const int stage = it % kStages;
            const uint32_t phase = static_cast<uint32_t>((it / kStages) & 1);
            mbar_wait(&full[stage], phase, 63);
            tc_fence_after();
            if (elect_one_sync()) {
                const uint32_t sx = smem_u32(smem + stage * kStageBytes);
                const uint32_t sdy = sx + 2 * kXSlab;
                const uint64_t bd = make_smem_desc_sw128(sdy, kDYSlab, 1024);   // one 64-column slab per CTA
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    const uint64_t ad = make_smem_desc_sw128(sx + b * 128, kXSlab, kSboA);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        umma2_bf16(tmem_base + b * 128, ad + static_cast<uint64_t>(k * kStepA),
                                   bd + static_cast<uint64_t>(k * 128), idesc, (it | k) != 0 ? 1u : 0u);
                    }
                }
                umma2_commit_both(&empty[stage]);
            }
            __syncwarp();
        }
        if (elect_one_sync()) umma2_commit_both(&tfull[0]);
        __syncwarp();
    } else if (warp >= 2 && !dummy) {
        // ---------------------------------- epilogue (each CTA: its own 128 TMEM lanes = its unit's rows) ------
        const int q = warp & 3;
        const int row = q * 32 + lane;
        mbar_wait(&tfull[0], 0, 64);
        tc_fence_after();
        uint8_t* stg = smem + q * 4096;          // the pipeline's shared memory is idle by now
        const int piece = lane & 7;
#pragma unroll 1
        for (int b = 0; b < 3; ++b) {
            const long long mine = (static_cast<long long>(a * 3 + b) * p.cin + cp * 128 + row) * p.cout + co * 128;
            long long rowoff[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) rowoff[it] = __shfl_sync(0xffffffffu, mine, it * 4 + (lane >> 3));
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                float v[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(b * 128 + c * 32), v);
                __syncwarp();
                float4* rowp = reinterpret_cast<float4*>(stg + lane * 128);
                const int sw = lane & 7;
#pragma unroll
                for (int g = 0; g < 8; ++g) rowp[g ^ sw] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
                __syncwarp();
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = it * 4 + (lane >> 3);
                    const float4 val = *reinterpret_cast<const float4*>(stg + r * 128 + ((piece ^ (r & 7)) << 4));
                    float* dst = p.dw + rowoff[it] + c * 32 + piece * 4;
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(val.x), "f"(val.y),
                                 "f"(val.z), "f"(val.w)
                                 : "memory");
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();      // the peer's shared memory stays valid until the leader's MMAs are done with it
    tc_fence_after();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
}

}  // namespace

// p.splits / p.steps_per_split are per PAIR here; grid = 2 * ceil(3*cin_pairs / 2) * cout_slabs * splits
cudaError_t launch_wgrad_row2(const CUtensorMap& x, const CUtensorMap& dy, const WgradRowParams& p, int w16,
                              cudaStream_t stream) {
    static bool attr_done[2] = {false, false};
    auto k16 = wgrad_row2_kernel<true>;
    auto k8 = wgrad_row2_kernel<false>;
    if (!attr_done[w16 ? 1 : 0]) {
        cudaError_t e = w16 ? cudaFuncSetAttribute(k16, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes)
                            : cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[w16 ? 1 : 0] = true;
    }
    const int unit_pairs = (3 * p.cin_pairs + 1) / 2;
    const int grid = 2 * unit_pairs * p.cout_slabs * p.splits;
    if (w16)
        B2C_LAUNCH(k16, grid, 192, kSmemBytes, stream, x, dy, p);
    else
        B2C_LAUNCH(k8, grid, 192, kSmemBytes, stream, x, dy, p);
    return cudaGetLastError();
}

}  // namespace b2c
