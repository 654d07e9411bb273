// conv_halo.cu — 3x3 stride-1 convolution forward / data-gradient with activation-halo reuse.
//
// Same GEMM view, operands and epilogue as conv_gemm.cu (reference sites: models/resnet.py:23-27 and the
// autograd dgrad of main.py:104), but built around the measured limiter of that kernel: bytes staged from
// L2 into shared memory per MMA (profiles/r1_ncu_tap_gemm_full.txt — 45 B/clk/SM, the L2->SM ceiling, with
// the tensor pipe 45 % active). Here the activation tile of a 64-channel chunk is loaded ONCE with its halo
// and the taps are issued as row-shifted shared-memory descriptors over it; tcgen05 applies the 128-byte
// swizzle to the absolute smem address, so a descriptor may start at any 128 B row and step between 8-row
// groups by any SBO (verified on hardware: profiles/r1_umma_probe.log, P1).
//
//   mode 0 (W >= 16): tile = 16 rows x 8 cols of one image; halo box {64 ch, 10, 18, 1}; tap (a, b) of the
//                     3x3 window starts at row a*10 + b of the tile; 8-row groups are 10 rows apart (SBO 1280).
//   mode 1 (W == 8) : tile = two whole 8x8 images; three boxes {64 ch, 10, 8, 2}, one per window row a,
//                     loaded at image row a-1; tap b starts at row b.
// Activation bytes staged per output tile and chunk: 23 KB (mode 0) or 60 KB (mode 1) instead of 144 KB.
// Weight tiles stream through their own ring; when all of them fit (cin = cout = 64: 72 KB) they are
// loaded once per CTA and stay resident.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue.
#include "b2c_kernels.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kAStageBytes = 23552;  // 180 rows x 128 B rounded up to 1024

template <int BN>
struct HaloCfg {
    static constexpr int kBBytes = BN * 128;
    static constexpr int kSA = BN == 64 ? 4 : 3;
    static constexpr int kSB = BN == 64 ? 9 : (BN == 128 ? 8 : 4);
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kBarBytes = 512;
    static constexpr int kStatBytes = 2 * 4 * BN * 4;
    static constexpr int kSmemBytes = 1024 + kSA * kAStageBytes + kSB * kBBytes + kBarBytes + kStatBytes;
};

template <int BN, bool kStats>
__global__ void __launch_bounds__(192, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const __grid_constant__ HaloParams p) {
    using Cfg = HaloCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + Cfg::kSA * kAStageBytes;
    uint8_t* aux = sB + Cfg::kSB * Cfg::kBBytes;
    uint64_t* full_a = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty_a = full_a + Cfg::kSA;
    uint64_t* full_b = empty_a + Cfg::kSA;
    uint64_t* empty_b = full_b + Cfg::kSB;
    uint64_t* tfull = empty_b + Cfg::kSB;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + 2);
    float* stat = reinterpret_cast<float*>(aux + Cfg::kBarBytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < Cfg::kSA; ++s) {
                mbar_init(&full_a[s], 1);
                mbar_init(&empty_a[s], 1);
            }
            for (int s = 0; s < Cfg::kSB; ++s) {
                mbar_init(&full_b[s], 1);
                mbar_init(&empty_b[s], 1);
            }
            mbar_init(&tfull[0], 1);
            mbar_init(&tfull[1], 1);
            mbar_init(&tempty[0], 4);
            mbar_init(&tempty[1], 4);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    const int total_tiles = p.num_m_tiles * p.num_n_tiles;
    const int loads = p.mode == 0 ? 1 : 3;          // A boxes per chunk
    const int taps_per_load = 9 / loads;
    const uint32_t a_bytes = p.mode == 0 ? 180u * 128u : 160u * 128u;
    const bool resident = p.b_resident != 0;

    if (warp == 0) {
        if (lane == 0) {
            // ------------------------------ TMA producer ------------------------------
            int sa = 0, sb = 0;
            uint32_t pa = 0, pb = 0;
            bool first = true;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int m_tile = tile / p.num_n_tiles;
                const int n_tile = tile - m_tile * p.num_n_tiles;
                int n0, h0, w0;
                if (p.mode == 0) {
                    const int per_img = p.tiles_h * p.tiles_w;
                    n0 = m_tile / per_img;
                    const int rem = m_tile - n0 * per_img;
                    const int th = rem / p.tiles_w;
                    h0 = th * 16;
                    w0 = (rem - th * p.tiles_w) * 8;
                } else {
                    n0 = m_tile * 2;
                    h0 = 0;
                    w0 = 0;
                }
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    for (int l = 0; l < loads; ++l) {
                        mbar_wait(&empty_a[sa], pa ^ 1u, 21);
                        mbar_arrive_expect_tx(&full_a[sa], a_bytes);
                        tma_load_4d(sA + sa * kAStageBytes, &mapA, &full_a[sa], kc * 64, w0 - 1,
                                    p.mode == 0 ? h0 - 1 : l - 1, n0);
                        if (++sa == Cfg::kSA) {
                            sa = 0;
                            pa ^= 1u;
                        }
                        if (!resident || first) {
                            for (int j = 0; j < taps_per_load; ++j) {
                                const int a = p.mode == 0 ? j / 3 : l;
                                const int b = p.mode == 0 ? j - a * 3 : j;
                                const int r = p.flip ? 2 - a : a;
                                const int s = p.flip ? 2 - b : b;
                                if (!resident) mbar_wait(&empty_b[sb], pb ^ 1u, 22);
                                mbar_arrive_expect_tx(&full_b[sb], Cfg::kBBytes);
                                tma_load_2d(sB + sb * Cfg::kBBytes, &mapB, &full_b[sb], kc * 64,
                                            (r * 3 + s) * p.b_rows_per_tap + n_tile * BN);
                                if (++sb == Cfg::kSB) {
                                    sb = 0;
                                    pb ^= 1u;
                                }
                            }
                        }
                    }
                }
                first = false;
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ------------------------------ MMA issuer --------------------------------
            constexpr uint32_t idesc = make_idesc_bf16(128, BN, 0, 0);
            int sa = 0, sb = 0;
            uint32_t pa = 0, pb = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            bool first = true;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                mbar_wait(&tempty[acc], acc_phase ^ 1u, 23);
                tc_fence_after();
                const uint32_t d = tmem_base + static_cast<uint32_t>(acc * BN);
                uint32_t accum = 0;
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    for (int l = 0; l < loads; ++l) {
                        mbar_wait(&full_a[sa], pa, 24);
                        tc_fence_after();
                        const uint32_t a_base = smem_u32(sA + sa * kAStageBytes);
                        for (int j = 0; j < taps_per_load; ++j) {
                            const int a = p.mode == 0 ? j / 3 : 0;
                            const int b = p.mode == 0 ? j - a * 3 : j;
                            if (!resident || first) {
                                mbar_wait(&full_b[sb], pb, 25);
                                tc_fence_after();
                            }
                            const uint64_t ad = make_smem_desc_sw128(a_base + (a * 10 + b) * 128, 16, 1280);
                            const uint64_t bd = make_smem_desc_sw128(smem_u32(sB + sb * Cfg::kBBytes), 16, 1024);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                umma_bf16(d, ad + 2u * k, bd + 2u * k, idesc, accum);
                                accum = 1;
                            }
                            if (!resident) umma_commit(&empty_b[sb]);
                            if (++sb == Cfg::kSB) {
                                sb = 0;
                                pb ^= 1u;
                            }
                        }
                        umma_commit(&empty_a[sa]);
                        if (++sa == Cfg::kSA) {
                            sa = 0;
                            pa ^= 1u;
                        }
                    }
                }
                if (resident) {  // the ring never advances past the resident slabs
                    sb = 0;
                    pb = 0;
                }
                umma_commit(&tfull[acc]);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1u;
                }
                first = false;
            }
        }
    } else {
        // ---------------------------------- epilogue --------------------------------------
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int g = row >> 3, j = row & 7;
        const int et = threadIdx.x - 64;
        if (kStats) {
            for (int c = lane; c < BN; c += 32) {
                stat[(0 * 4 + q) * BN + c] = 0.f;
                stat[(1 * 4 + q) * BN + c] = 0.f;
            }
        }
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int m_tile = tile / p.num_n_tiles;
            const int n_tile = tile - m_tile * p.num_n_tiles;
            int n, h, w;
            if (p.mode == 0) {
                const int per_img = p.tiles_h * p.tiles_w;
                n = m_tile / per_img;
                const int rem = m_tile - n * per_img;
                const int th = rem / p.tiles_w;
                h = th * 16 + g;
                w = (rem - th * p.tiles_w) * 8 + j;
            } else {
                n = m_tile * 2 + (g >> 3);
                h = g & 7;
                w = j;
            }
            const bool valid = n < p.N;
            const size_t off = (static_cast<size_t>(n * p.H + h) * p.W + w) * p.out_ld + static_cast<size_t>(n_tile) * BN;
            mbar_wait(&tfull[acc], acc_phase, 26);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                float v[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                  static_cast<uint32_t>(acc * BN + c * 32),
                              v);
                if (valid) {
                    if (p.add_src != nullptr) {
                        const uint4* src = reinterpret_cast<const uint4*>(p.add_src + off + c * 32);
#pragma unroll
                        for (int gg = 0; gg < 4; ++gg) {
                            const uint4 u = __ldg(src + gg);
                            v[gg * 8 + 0] += bf16lo(u.x);
                            v[gg * 8 + 1] += bf16hi(u.x);
                            v[gg * 8 + 2] += bf16lo(u.y);
                            v[gg * 8 + 3] += bf16hi(u.y);
                            v[gg * 8 + 4] += bf16lo(u.z);
                            v[gg * 8 + 5] += bf16hi(u.z);
                            v[gg * 8 + 6] += bf16lo(u.w);
                            v[gg * 8 + 7] += bf16hi(u.w);
                        }
                    }
                    uint4* dst = reinterpret_cast<uint4*>(p.out + off + c * 32);
#pragma unroll
                    for (int gg = 0; gg < 4; ++gg) {
                        uint4 u;
                        u.x = pack_bf16x2(v[gg * 8 + 0], v[gg * 8 + 1]);
                        u.y = pack_bf16x2(v[gg * 8 + 2], v[gg * 8 + 3]);
                        u.z = pack_bf16x2(v[gg * 8 + 4], v[gg * 8 + 5]);
                        u.w = pack_bf16x2(v[gg * 8 + 6], v[gg * 8 + 7]);
                        dst[gg] = u;
                    }
                }
                if (kStats) {
                    // rows of images past N were zero-filled by TMA, so they add nothing
                    float sq[32];
#pragma unroll
                    for (int e = 0; e < 32; ++e) sq[e] = v[e] * v[e];
                    const float s1 = warp_colsum32(v, lane);
                    const float s2 = warp_colsum32(sq, lane);
                    stat[(0 * 4 + q) * BN + c * 32 + lane] += s1;
                    stat[(1 * 4 + q) * BN + c * 32 + lane] += s2;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1u;
            }
        }
        if (kStats) {
            named_bar_sync(1, 128);
            const int n_tile = blockIdx.x % p.num_n_tiles;
            const int prow = blockIdx.x / p.num_n_tiles;
            for (int col = et; col < BN; col += 128) {
                const float s1 = stat[(0 * 4 + 0) * BN + col] + stat[(0 * 4 + 1) * BN + col] +
                                 stat[(0 * 4 + 2) * BN + col] + stat[(0 * 4 + 3) * BN + col];
                const float s2 = stat[(1 * 4 + 0) * BN + col] + stat[(1 * 4 + 1) * BN + col] +
                                 stat[(1 * 4 + 2) * BN + col] + stat[(1 * 4 + 3) * BN + col];
                const size_t o = static_cast<size_t>(prow) * p.cout + n_tile * BN + col;
                p.part_sum[o] = s1;
                p.part_sq[o] = s2;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <int BN, bool kStats>
cudaError_t launch_one(const CUtensorMap& a, const CUtensorMap& b, HaloParams p, int max_ctas, int* stat_rows,
                       cudaStream_t stream) {
    using Cfg = HaloCfg<BN>;
    auto kern = conv_halo_kernel<BN, kStats>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    p.b_resident = (9 * p.k_chunks <= Cfg::kSB && p.num_n_tiles == 1) ? 1 : 0;
    const int total = p.num_m_tiles * p.num_n_tiles;
    int grid = total < max_ctas ? total : max_ctas;
    grid -= grid % p.num_n_tiles;
    if (grid < p.num_n_tiles) grid = p.num_n_tiles;
    if (stat_rows != nullptr) *stat_rows = grid / p.num_n_tiles;
    kern<<<grid, 192, Cfg::kSmemBytes, stream>>>(a, b, p);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_conv_halo(int bn, const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int max_ctas,
                             int* stat_rows, cudaStream_t stream) {
    const bool stats = p.part_sum != nullptr;
    switch (bn) {
        case 64:
            return stats ? launch_one<64, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one<64, false>(a, b, p, max_ctas, stat_rows, stream);
        case 128:
            return stats ? launch_one<128, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one<128, false>(a, b, p, max_ctas, stat_rows, stream);
        case 256:
            return stats ? launch_one<256, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one<256, false>(a, b, p, max_ctas, stat_rows, stream);
        default:
            return cudaErrorInvalidValue;
    }
}

}  // namespace b2c
