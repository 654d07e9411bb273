// conv_wgrad.cu — convolution weight gradient on tcgen05 tensor cores.
//
// Replaces the cuDNN backward-filter kernel the reference reaches through autograd of nn.Conv2d
// (`loss.backward()` main.py:104; conv sites models/resnet.py:23-27,33-34,102-103).
//
//   GEMM view   dW[(tap, cin), cout] = sum_pixels X[pixel shifted by tap, cin] * dY[pixel, cout]
//   Both operands have the reduction (pixels) as the strided dimension and channels contiguous
//   (NHWC), so both are fed to tcgen05.mma as MN-major SWIZZLE_128B tiles straight from TMA:
//   a k-step is a box of 64 or 128 pixels x 64 channels (8 / 16 KB; 128 wherever the
//   image tiles that way and cout <= 128 — fewer, larger pipeline steps keep the single issuing warp ahead); D rows = two 64-channel (tap, cin-block)
//   slabs of X, D columns = a BN-wide slice of cout, taken from BN/64 slabs of dY.
//   The pixel reduction is split across CTAs (split-K); partial tiles are combined with fp32
//   `red.global.add` straight into the caller's zero-initialised gradient tensor, at the
//   strides the caller gives (OIHW for the reference's `.grad` layout).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 = epilogue.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

template <int BN, int KC>
struct WgCfg {
    static constexpr int kSlabBytes = KC * 128;  // KC pixels x 64 bf16 channels
    static constexpr int kNSlabs = BN / 64;
    static constexpr int kStageBytes = (2 + kNSlabs) * kSlabBytes;
    static constexpr int kStages = (208 * 1024) / kStageBytes < 8 ? (208 * 1024) / kStageBytes : 8;
    static constexpr int kTmemCols = BN;
    static constexpr int kBarBytes = 256;
    static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kBarBytes;
};

template <int BN, int KC>
__global__ void __launch_bounds__(192, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapDY,
             const __grid_constant__ WgradParams p) {
    using Cfg = WgCfg<BN, KC>;
    constexpr int kSlabBytes = Cfg::kSlabBytes;
    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* aux = smem + Cfg::kStages * Cfg::kStageBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty = full + Cfg::kStages;
    uint64_t* tfull = empty + Cfg::kStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // work decomposition
    int id = blockIdx.x;
    const int n_tile = id % p.num_n_tiles;
    id /= p.num_n_tiles;
    const int m_blk = id % p.num_m_blocks;
    const int split = id / p.num_m_blocks;
    const int k_begin = split * p.chunks_per_split;
    int k_end = k_begin + p.chunks_per_split;
    if (k_end > p.k_chunks_total) k_end = p.k_chunks_total;
    const int n_iters = k_end - k_begin;  // host guarantees >= 1

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapX);
        tma_prefetch_desc(&mapDY);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < Cfg::kStages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
            }
            mbar_init(&tfull[0], 1);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();      // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
    const int img = p.gH * p.gW;

    if (warp == 0) {
        {
            // ------------------------------ TMA producer (whole warp loops, one elected lane issues) ----
            int slab[2], tap[2], cb[2];
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                slab[b] = m_blk * 2 + b;
                if (slab[b] >= p.num_slabs) slab[b] = p.num_slabs - 1;  // duplicate; rows ignored later
                tap[b] = slab[b] / p.cin_blocks;
                cb[b] = slab[b] - tap[b] * p.cin_blocks;
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int kc = k_begin; kc < k_end; ++kc) {
                const int p0 = kc * KC;
                const int n0 = p0 / img;
                const int h0 = (p0 - n0 * img) / p.gW;
                mbar_wait(&empty[stage], phase ^ 1u, 11);
                if (elect_one_sync()) {
                    mbar_arrive_expect_tx(&full[stage], Cfg::kStageBytes);
                    uint8_t* sa = smem + stage * Cfg::kStageBytes;
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int cw = p.tap_dw[tap[b]];
                        if (p.x_split_w)     // stride 2: W axis split into (parity, w/2), see encode_nhwc_s2
                            tma_load_5d(sa + b * kSlabBytes, &mapX, &full[stage], cb[b] * 64, cw & 1, cw >> 1,
                                        h0 * p.in_stride + p.tap_dh[tap[b]], n0);
                        else
                            tma_load_4d(sa + b * kSlabBytes, &mapX, &full[stage], cb[b] * 64, cw,
                                        h0 * p.in_stride + p.tap_dh[tap[b]], n0);
                    }
#pragma unroll
                    for (int nb = 0; nb < Cfg::kNSlabs; ++nb) {
                        tma_load_4d(sa + (2 + nb) * kSlabBytes, &mapDY, &full[stage],
                                    (n_tile * Cfg::kNSlabs + nb) * 64, 0, h0, n0);
                    }
                }
                __syncwarp();
                if (++stage == Cfg::kStages) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
        }
    } else if (warp == 1) {
        {
            // ------------------------------ MMA issuer (whole warp loops, one elected lane issues) ------
            constexpr uint32_t idesc = make_idesc_bf16(128, BN, 1, 1);  // both operands MN-major
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < n_iters; ++it) {
                mbar_wait(&full[stage], phase, 13);
                tc_fence_after();
                if (elect_one_sync()) {
                    const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
                    // MN-major SW128: 64-element MN blocks kSlabBytes apart (LBO), 8-pixel K groups 1024 B apart (SBO)
                    const uint64_t ad = make_smem_desc_sw128(sa, kSlabBytes, 1024);
                    const uint64_t bd = make_smem_desc_sw128(sa + 2 * kSlabBytes, kSlabBytes, 1024);
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k) {
                        // advance 16 pixels = 16 rows x 128 B = 2048 B
                        umma_bf16(tmem_base, ad + 128u * k, bd + 128u * k, idesc, (it | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty[stage]);
                }
                __syncwarp();
                if (++stage == Cfg::kStages) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
            if (elect_one_sync()) umma_commit(&tfull[0]);
            __syncwarp();
        }
    } else {
        // ---------------------------------- epilogue --------------------------------------
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int b = row >> 6;
        const int slab = m_blk * 2 + b;
        const int tap = slab / p.cin_blocks;
        const int cin = (slab - tap * p.cin_blocks) * 64 + (row & 63);
        const bool valid = slab < p.num_slabs && cin < p.cin_valid;
        mbar_wait(&tfull[0], 0, 14);
        tc_fence_after();
        int gc = cin, gt = tap;
        if (p.kflat_cin > 0) {  // im2col'ed stem: unflatten k -> (tap, c)
            gt = cin / p.kflat_cin;
            gc = cin - gt * p.kflat_cin;
        }
        float* base = p.dw + static_cast<long long>(gc) * p.c_stride + static_cast<long long>(gt) * p.t_stride +
                      static_cast<long long>(n_tile) * BN * p.o_stride;
        if (p.o_stride == 1 && p.debug == 0 && !(p.cout_valid & 3)) {
            // dW laid out [tap][cin][cout]: a row of the accumulator tile is contiguous in memory. Each 32x32 fp32
            // chunk is transposed through a 4 KB per-warp stage (the pipeline's shared memory is idle by now) so that
            // one red.global.add.v4.f32 covers 4 rows x 128 contiguous bytes: 4 lines per instruction instead of one
            // scattered 4-byte atomic per lane and element.
            uint8_t* stg = smem + q * 4096;
            const long long mine = valid ? (static_cast<long long>(gc) * p.c_stride + static_cast<long long>(gt) * p.t_stride)
                                         : -1;
            long long rowoff[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) rowoff[it] = __shfl_sync(0xffffffffu, mine, it * 4 + (lane >> 3));
            const int piece = lane & 7;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                float v[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c * 32), v);
                __syncwarp();
                float4* rowp = reinterpret_cast<float4*>(stg + lane * 128);
                const int sw = lane & 7;
#pragma unroll
                for (int g = 0; g < 8; ++g) rowp[g ^ sw] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
                __syncwarp();
                const int col = n_tile * BN + c * 32 + piece * 4;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = it * 4 + (lane >> 3);
                    const float4 val = *reinterpret_cast<const float4*>(stg + r * 128 + ((piece ^ (r & 7)) << 4));
                    if (rowoff[it] >= 0 && col < p.cout_valid) {
                        float* dst = p.dw + rowoff[it] + col;
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(val.x),
                                     "f"(val.y), "f"(val.z), "f"(val.w)
                                     : "memory");
                    }
                }
            }
        } else {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            float v[32];
            tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c * 32), v);
            if (valid) {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    if (n_tile * BN + c * 32 + e < p.cout_valid) {
                        float* dst = base + static_cast<long long>(c * 32 + e) * p.o_stride;
                        if (p.debug == 0) atomicAdd(dst, v[e]);
                        else if (p.debug & 8) *dst = v[e];          // measurement only: plain store, wrong sums
                    }
                }
            }
        }
        }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <int BN, int KC>
cudaError_t launch_one(const CUtensorMap& x, const CUtensorMap& dy, const WgradParams& p, cudaStream_t stream) {
    using Cfg = WgCfg<BN, KC>;
    auto kern = wgrad_kernel<BN, KC>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int grid = p.num_n_tiles * p.num_m_blocks * p.splits;
    B2C_LAUNCH(kern, grid, 192, Cfg::kSmemBytes, stream, x, dy, p);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_wgrad(int bn, const CUtensorMap& x, const CUtensorMap& dy, const WgradParams& p,
                         cudaStream_t stream) {
    if (p.k_pixels == 128) {
        switch (bn) {
            case 64: return launch_one<64, 128>(x, dy, p, stream);
            case 128: return launch_one<128, 128>(x, dy, p, stream);
            default: return cudaErrorInvalidValue;
        }
    }
    switch (bn) {
        case 64: return launch_one<64, 64>(x, dy, p, stream);
        case 128: return launch_one<128, 64>(x, dy, p, stream);
        case 256: return launch_one<256, 64>(x, dy, p, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace b2c
