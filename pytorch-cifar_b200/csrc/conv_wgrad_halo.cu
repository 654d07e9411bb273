// conv_wgrad_halo.cu — weight gradient of 3x3 stride-1 convolutions with activation-halo reuse.
//
// Same GEMM as conv_wgrad.cu (reference site: autograd of nn.Conv2d, main.py:104; conv sites
// models/resnet.py:23-27), for images of 16 or 32 pixels a side. conv_wgrad.cu loads the activation tile once per
// tap (ncu: 2.05 GB of L2->SM reads per launch on the 64->64 layer, profiles/r1b_ncu_conv_full.txt); here a
// k-step is a 16x8-pixel tile whose activation box is loaded ONCE with its halo ({64 ch, 10, 18, 1}) and all nine
// taps are MN-major descriptors shifted over it (tcgen05 applies the swizzle to the absolute smem address:
// profiles/r1_umma_probe.log, P2). Per k-step: 22.5 KB of X + 16 KB of dY feed 40 MMAs.
//
//   CTA work unit   one 64-channel slab of cin x one 64-channel slab of cout x a range of pixel tiles (split-K)
//   accumulators    5 blocks of 128 rows (two taps x 64 cin each; the tenth half-block is unused) x 64 cout columns
//                   = 320 TMEM columns
//   output          fp32 partial tiles to a caller-provided scratch buffer [unit][split][5][128][64]; a second,
//                   tiny kernel sums the splits into the OIHW gradient (no atomics, deterministic)
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kXBytes = 180 * 128;      // halo box
constexpr int kXStride = 23552;         // rounded up to 1024
constexpr int kDYBytes = 128 * 128;     // 16x8 pixels x 64 channels
constexpr int kStageBytes = kXStride + kDYBytes;
constexpr int kStages = 5;
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
constexpr int kTmemCols = 512;

__global__ void __launch_bounds__(192, 1)
wgrad_halo_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapDY,
                  const __grid_constant__ WgradHaloParams p) {
    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* aux = smem + kStages * kStageBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty = full + kStages;
    uint64_t* tfull = empty + kStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // work decomposition: blockIdx = (unit, split); unit = (cin slab, cout slab)
    const int unit = blockIdx.x / p.splits;
    const int split = blockIdx.x - unit * p.splits;
    const int ci = unit / p.cout_slabs;
    const int co = unit - ci * p.cout_slabs;
    const int t_begin = split * p.tiles_per_split;
    int t_end = t_begin + p.tiles_per_split;
    if (t_end > p.num_tiles) t_end = p.num_tiles;
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
    pdl_wait();      // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
    const int per_img = p.tiles_h * p.tiles_w;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        for (int it = 0; it < n_iters; ++it) {
            const int stage = it % kStages;
            const uint32_t phase = static_cast<uint32_t>((it / kStages) & 1);
            const int t = t_begin + it;
            const int n = t / per_img;
            const int rem = t - n * per_img;
            const int th = rem / p.tiles_w;
            const int h0 = th * 16, w0 = (rem - th * p.tiles_w) * 8;
            mbar_wait(&empty[stage], phase ^ 1u, 31);
            if (elect_one_sync()) {
                uint8_t* sx = smem + stage * kStageBytes;
                mbar_arrive_expect_tx(&full[stage], kXBytes + kDYBytes);
                tma_load_4d(sx, &mapX, &full[stage], ci * 64, w0 - 1, h0 - 1, n);
                tma_load_4d(sx + kXStride, &mapDY, &full[stage], co * 64, w0, h0, n);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer --------------------------------
        constexpr uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);   // both operands MN-major
        for (int it = 0; it < n_iters; ++it) {
            const int stage = it % kStages;
            const uint32_t phase = static_cast<uint32_t>((it / kStages) & 1);
            mbar_wait(&full[stage], phase, 33);
            tc_fence_after();
            if (elect_one_sync()) {
                const uint32_t sx = smem_u32(smem + stage * kStageBytes);
                const uint32_t sdy = sx + kXStride;
#pragma unroll
                for (int mb = 0; mb < 5; ++mb) {
                    // rows 0..63 = tap 2*mb, rows 64..127 = tap 2*mb+1 (the last block repeats tap 8; ignored later)
                    const int t0 = 2 * mb, t1 = mb == 4 ? 8 : 2 * mb + 1;
                    const int off0 = (t0 / 3) * 10 + (t0 % 3), off1 = (t1 / 3) * 10 + (t1 % 3);
                    // MN-major A: 64-element M slabs (off1-off0) rows apart (LBO), 8-pixel K groups one halo row
                    // (10 pixels = 1280 B) apart (SBO); a k-step of 16 pixels = two tile rows = 2560 B
                    const uint64_t ad = make_smem_desc_sw128(sx + off0 * 128, (off1 - off0) * 128, 1280);
                    const uint64_t bd = make_smem_desc_sw128(sdy, 0, 1024);
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        umma_bf16(tmem_base + mb * 64, ad + static_cast<uint64_t>(k * 160), bd + static_cast<uint64_t>(k * 128),
                                  idesc, (it | k) != 0 ? 1u : 0u);
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
        const int q = warp & 3;
        const int row = q * 32 + lane;
        mbar_wait(&tfull[0], 0, 34);
        tc_fence_after();
        float* dst = p.partial + (static_cast<size_t>(blockIdx.x) * 5 * 128 + row) * 64;
#pragma unroll 1
        for (int mb = 0; mb < 5; ++mb) {
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                float v[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(mb * 64 + c * 32), v);
                float4* o = reinterpret_cast<float4*>(dst + static_cast<size_t>(mb) * 128 * 64 + c * 32);
#pragma unroll
                for (int g = 0; g < 8; ++g) o[g] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc<kTmemCols>(tmem_base);
}

// dw[cout, cin, tap] += sum over splits of the partial tiles. Block = 64 consecutive (tap, cin, cout) elements
// (cout fastest = the contiguous axis of the partial tiles) x 4 split lanes; every lane sums its quarter of the
// splits with eight loads in flight, the four partial sums meet in shared memory in a fixed order (deterministic).
__global__ void __launch_bounds__(256)
wgrad_halo_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int cin, int cout, int cout_slabs,
                         int splits, int tco) {
    pdl_trigger_tiny();
    pdl_wait();
    __shared__ float red[4][64];
    const int x = threadIdx.x & 63, y = threadIdx.x >> 6;
    const long long total = 9LL * cin * cout;
    const size_t split_stride = static_cast<size_t>(5) * 128 * 64;
    for (long long base = blockIdx.x * 64LL; base < total; base += static_cast<long long>(gridDim.x) * 64) {
        const long long i = base + x;
        float s = 0.f;
        int o = 0, c = 0, t = 0;
        if (i < total) {
            o = static_cast<int>(i % cout);
            c = static_cast<int>((i / cout) % cin);
            t = static_cast<int>(i / (static_cast<long long>(cout) * cin));
            const int unit = (c >> 6) * cout_slabs + (o >> 6);
            const int mb = t >> 1, row = (t & 1) * 64 + (c & 63), col = o & 63;
            const float* src = partial + ((static_cast<size_t>(unit) * splits * 5 + mb) * 128 + row) * 64 + col;
            int k = y;
            float a[8];
            for (; k + 28 < splits; k += 32) {
#pragma unroll
                for (int u = 0; u < 8; ++u) a[u] = src[static_cast<size_t>(k + 4 * u) * split_stride];
#pragma unroll
                for (int u = 0; u < 8; ++u) s += a[u];
            }
            for (; k < splits; k += 4) s += src[static_cast<size_t>(k) * split_stride];
        }
        red[y][x] = s;
        __syncthreads();
        if (y == 0 && i < total) {
            // i = (t * cin + c) * cout + o is itself the [tap][cin][cout] index
            const size_t di = tco ? static_cast<size_t>(i) : (static_cast<size_t>(o) * cin + c) * 9 + t;
            dw[di] += (red[0][x] + red[1][x]) + (red[2][x] + red[3][x]);
        }
        __syncthreads();
    }
}

}  // namespace

size_t wgrad_halo_scratch_floats(int cin, int cout, int splits) {
    return static_cast<size_t>(cin / 64) * (cout / 64) * splits * 5 * 128 * 64;
}

cudaError_t launch_wgrad_halo(const CUtensorMap& x, const CUtensorMap& dy, const WgradHaloParams& p, float* dw,
                              cudaStream_t stream) {
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int grid = p.cin_slabs * p.cout_slabs * p.splits;
    B2C_LAUNCH(wgrad_halo_kernel, grid, 192, kSmemBytes, stream, x, dy, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const long long total = 9LL * p.cin_slabs * 64 * p.cout_slabs * 64;
    int rgrid = static_cast<int>((total + 63) / 64);
    if (rgrid > 148 * 8) rgrid = 148 * 8;
    B2C_LAUNCH(wgrad_halo_reduce_kernel, rgrid, 256, 0, stream, p.partial, dw, p.cin_slabs * 64, p.cout_slabs * 64, p.cout_slabs,
                                                        p.splits, p.tco);
    return cudaGetLastError();
}

}  // namespace b2c
