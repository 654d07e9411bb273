// conv_wgrad_row.cu — weight gradient of 3x3 stride-1 convolutions, one kernel ROW of taps per CTA.
//
// Same GEMM as conv_wgrad.cu (reference site: autograd of nn.Conv2d, main.py:104; conv sites
// models/resnet.py:23-27), for 16x16 and 8x8 images with cin and cout multiples of 128. conv_wgrad.cu loads one
// activation box per (tap, 64-channel slab) and one dY tile per 128-row D block: 64 KB of operands per 8 MMAs, and ncu
// shows it bound by L2 latency x bytes in flight, not by the tensor pipe. Here a CTA owns the three taps of one kernel
// row `a` for a 128-channel pair of cin slabs and a 128-wide slice of cout:
//
//   k-step         64 pixels: 4 image rows x 16 (W = 16) or one whole 8x8 image (W = 8)
//   X operand      per cin slab ONE box with a one-pixel halo in W ({64 ch, W+2, rows, 1}, rows shifted by a-1);
//                  the taps b = 0..2 are MN-major descriptors starting b pixels into it (tcgen05 applies the
//                  128-byte swizzle to the absolute address: profiles/r1_umma_probe.log, P2)
//   dY operand     one {64 ch, W, rows, 1} box per 64-channel cout slab, shared by the three D blocks
//   accumulators   3 blocks (tap b) x 128 rows (cin pair) x 128 cout columns = 384 TMEM columns
//   per k-step     34 KB of operands feed 12 MMAs (was 64 KB per 8)
//   output         dW in [tap][cin][cout] order (B2cConvDesc.flags bit 5): coalesced red.global.add.v4.f32
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kXSlab = 10240;                 // one cin slab's box: <= 80 pixels x 128 B, rounded to 1024
constexpr int kDYSlab = 8192;                 // 64 pixels x 64 channels
constexpr int kStageBytes = 2 * kXSlab + 2 * kDYSlab;   // 36 KB
constexpr int kStages = 5;
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
constexpr int kTmemCols = 512;

// W16: true = 16x16 images (k-step = 4 rows x 16), false = 8x8 images (k-step = 8 rows x 8)
template <bool W16>
__global__ void __launch_bounds__(192, 1)
wgrad_row_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapDY,
                 const __grid_constant__ WgradRowParams p) {
    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* aux = smem + kStages * kStageBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty = full + kStages;
    uint64_t* tfull = empty + kStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull + 1);

    constexpr int kRows = W16 ? 4 : 8;        // image rows per k-step
    constexpr int kW = W16 ? 16 : 8;
    constexpr uint32_t kXBytes = static_cast<uint32_t>(kRows * (kW + 2) * 128);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // blockIdx = (unit, split); unit = (kernel row a, cin pair, cout slab)
    const int unit = blockIdx.x / p.splits;
    const int split = blockIdx.x - unit * p.splits;
    const int a = unit / (p.cin_pairs * p.cout_slabs);
    const int rem_u = unit - a * p.cin_pairs * p.cout_slabs;
    const int cp = rem_u / p.cout_slabs;
    const int co = rem_u - cp * p.cout_slabs;
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
        tmem_alloc<kTmemCols>(tmem_ptr);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        for (int it = 0; it < n_iters; ++it) {
            const int stage = it % kStages;
            const uint32_t phase = static_cast<uint32_t>((it / kStages) & 1);
            const int t = t_begin + it;
            const int n = W16 ? (t >> 2) : t;           // four k-steps per 16x16 image, one per 8x8 image
            const int h0 = W16 ? (t & 3) * 4 : 0;
            mbar_wait(&empty[stage], phase ^ 1u, 41);
            if (elect_one_sync()) {
                uint8_t* sx = smem + stage * kStageBytes;
                mbar_arrive_expect_tx(&full[stage], 2 * kXBytes + 2 * kDYSlab);
                tma_load_4d(sx, &mapX, &full[stage], cp * 128, -1, h0 + a - 1, n);
                tma_load_4d(sx + kXSlab, &mapX, &full[stage], cp * 128 + 64, -1, h0 + a - 1, n);
                tma_load_4d(sx + 2 * kXSlab, &mapDY, &full[stage], co * 128, 0, h0, n);
                tma_load_4d(sx + 2 * kXSlab + kDYSlab, &mapDY, &full[stage], co * 128 + 64, 0, h0, n);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer --------------------------------
        constexpr uint32_t idesc = make_idesc_bf16(128, 128, 1, 1);   // both operands MN-major
        // 8-pixel K groups: contiguous for W = 16 (1024 B apart), one haloed row (10 pixels) apart for W = 8;
        // a 16-pixel MMA step advances one haloed row of 18 pixels (W = 16) or two of 10 (W = 8)
        constexpr uint32_t kSboA = W16 ? 1024u : 1280u;
        constexpr uint32_t kStepA = W16 ? (18u * 128u) >> 4 : (20u * 128u) >> 4;
        for (int it = 0; it < n_iters; ++it) {
            const int stage = it % kStages;
            const uint32_t phase = static_cast<uint32_t>((it / kStages) & 1);
            mbar_wait(&full[stage], phase, 43);
            tc_fence_after();
            if (elect_one_sync()) {
                const uint32_t sx = smem_u32(smem + stage * kStageBytes);
                const uint32_t sdy = sx + 2 * kXSlab;
                const uint64_t bd = make_smem_desc_sw128(sdy, kDYSlab, 1024);
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    // D block b: rows 0..63 = cin slab 0, rows 64..127 = cin slab 1 (LBO = slab distance), both read
                    // b pixels into their haloed box
                    const uint64_t ad = make_smem_desc_sw128(sx + b * 128, kXSlab, kSboA);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        umma_bf16(tmem_base + b * 128, ad + static_cast<uint64_t>(k * kStepA),
                                  bd + static_cast<uint64_t>(k * 128), idesc, (it | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit(&empty[stage]);
            }
            __syncwarp();
        }
        if (elect_one_sync()) umma_commit(&tfull[0]);
        __syncwarp();
    } else {
        // ---------------------------------- epilogue --------------------------------------
        // TMEM lane = D row = cin channel (cp*128 + row); block b = tap a*3 + b; columns = cout. A row is contiguous in
        // the [tap][cin][cout] gradient: each 32x32 chunk goes through a 4 KB swizzled stage (the pipeline's shared
        // memory is idle by now) and leaves as red.global.add.v4.f32 covering 4 rows x 128 contiguous bytes.
        const int q = warp & 3;
        const int row = q * 32 + lane;
        mbar_wait(&tfull[0], 0, 44);
        tc_fence_after();
        uint8_t* stg = smem + q * 4096;
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
    tc_fence_after();
    if (warp == 1) tmem_dealloc<kTmemCols>(tmem_base);
}

}  // namespace

cudaError_t launch_wgrad_row(const CUtensorMap& x, const CUtensorMap& dy, const WgradRowParams& p, int w16,
                             cudaStream_t stream) {
    static bool attr_done[2] = {false, false};
    auto k16 = wgrad_row_kernel<true>;
    auto k8 = wgrad_row_kernel<false>;
    if (!attr_done[w16 ? 1 : 0]) {
        cudaError_t e = w16 ? cudaFuncSetAttribute(k16, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes)
                            : cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[w16 ? 1 : 0] = true;
    }
    const int grid = 3 * p.cin_pairs * p.cout_slabs * p.splits;
    if (w16)
        B2C_LAUNCH(k16, grid, 192, kSmemBytes, stream, x, dy, p);
    else
        B2C_LAUNCH(k8, grid, 192, kSmemBytes, stream, x, dy, p);
    return cudaGetLastError();
}

}  // namespace b2c
