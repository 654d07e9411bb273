// umma2_probe.cu — standalone hardware probe (no torch) for the next tensor-core step: tcgen05.mma.cta_group::2.
// A measurement tool, not part of the product path.
//
//   Q1  Numerics / addressing: a CTA pair (cluster of 2) computes D[256 x N] = A[256 x 64] * B[N x 64]^T with ONE
//       tcgen05.mma.cta_group::2 stream issued by the leader CTA. Each CTA holds its 128 rows of A and its N/2 rows
//       of B at the same shared-memory offsets (K-major, SWIZZLE_128B, filled by hand), TMEM allocated with
//       cta_group::2, completion multicast to both CTAs' mbarriers. Each CTA reads back its 128 D rows.
//   Q2  Throughput: cycles per M=256 MMA (K=16) for N = 64 / 128 / 256 issued back to back, against the single-CTA
//       M=128 figures of profiles/r1_umma_probe.log (cycles ~ 64 + N/2).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../pytorch-cifar_b200/csrc/b2c_ptx.cuh"

using namespace b2c;

#define CK(x)                                                                              \
    do {                                                                                   \
        cudaError_t e_ = (x);                                                              \
        if (e_ != cudaSuccess) {                                                           \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                       \
        }                                                                                  \
    } while (0)

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
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

// K-major SWIZZLE_128B tile [rows][64 bf16]: element (r, k) at r*128 + ((k/8) ^ (r%8))*16 + (k%8)*2
__device__ __forceinline__ void fill_sw128(uint8_t* tile, const __nv_bfloat16* src, int rows) {
    for (int i = threadIdx.x; i < rows * 64; i += blockDim.x) {
        const int r = i >> 6, k = i & 63;
        *reinterpret_cast<__nv_bfloat16*>(tile + r * 128 + (((k >> 3) ^ (r & 7)) << 4) + (k & 7) * 2) = src[i];
    }
}

struct Probe2Args {
    int N;          // total N of the pair (each CTA holds N/2 rows of B)
    int reps;       // > 0: timing mode, that many back-to-back K=64 passes
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
probe2_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B, float* __restrict__ D,
              long long* __restrict__ cycles, Probe2Args a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;                      // 128 rows x 128 B
    uint8_t* sB = smem + 16384;              // N/2 rows x 128 B (<= 16 KB)
    uint64_t* done = reinterpret_cast<uint64_t*>(smem + 32768);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(done + 1);
    const uint32_t rank = cluster_rank();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    fill_sw128(sA, A + static_cast<size_t>(rank) * 128 * 64, 128);
    fill_sw128(sB, B + static_cast<size_t>(rank) * (a.N / 2) * 64, a.N / 2);
    if (threadIdx.x == 0) {
        mbar_init(done, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)),
                     "n"(256)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> tensor core reads
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // both CTAs' operands, barriers and TMEM are ready
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (rank == 0 && warp == 0) {
        if (elect_one_sync()) {
            const uint32_t idesc = make_idesc_bf16(256, a.N, 0, 0);
            const uint64_t ad = make_smem_desc_sw128(smem_u32(sA), 16, 1024);
            const uint64_t bd = make_smem_desc_sw128(smem_u32(sB), 16, 1024);
            const int passes = a.reps > 0 ? a.reps : 1;
            const long long t0 = clock64();
            for (int p = 0; p < passes; ++p) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma2_bf16(tmem_base, ad + 2u * k, bd + 2u * k, idesc, (p | k) != 0 ? 1u : 0u);
            }
            umma2_commit_both(done);
            mbar_wait(done, 0, 91);
            if (cycles != nullptr) cycles[0] = clock64() - t0;
        }
        __syncwarp();
    }
    mbar_wait(done, 0, 90);
    tc_fence_after();
    // read back this CTA's 128 rows (TMEM lane = row), N columns
    float* out = D + (static_cast<size_t>(rank) * 128 + warp * 32 + lane) * a.N;
    for (int c = 0; c < a.N / 32; ++c) {
        float v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + static_cast<uint32_t>(c * 32), v);
#pragma unroll
        for (int e = 0; e < 32; ++e) out[c * 32 + e] = v[e];
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(256) : "memory");
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
    CK(cudaSetDevice(0));
    const int smem = 1024 + 32768 + 64;
    CK(cudaFuncSetAttribute(probe2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    for (int N : {64, 128, 256}) {
        std::vector<__nv_bfloat16> hA(256 * 64), hB(N * 64);
        std::vector<float> fA(256 * 64), fB(N * 64);
        unsigned s = 12345u + N;
        auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 9) & 0xffff) / 65536.0f - 0.5f; };
        for (size_t i = 0; i < hA.size(); ++i) { fA[i] = bf(rnd()); hA[i] = __float2bfloat16(fA[i]); }
        for (size_t i = 0; i < hB.size(); ++i) { fB[i] = bf(rnd()); hB[i] = __float2bfloat16(fB[i]); }
        __nv_bfloat16 *dA, *dB;
        float* dD;
        long long* dC;
        CK(cudaMalloc(&dA, hA.size() * 2));
        CK(cudaMalloc(&dB, hB.size() * 2));
        CK(cudaMalloc(&dD, 256 * N * 4));
        CK(cudaMalloc(&dC, 8));
        CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0xff, 256 * N * 4));
        Probe2Args a{N, 0};
        probe2_kernel<<<2, 128, smem>>>(dA, dB, dD, dC, a);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        std::vector<float> hD(256 * N);
        CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
        double maxerr = 0, maxref = 0;
        for (int m = 0; m < 256; ++m)
            for (int n = 0; n < N; ++n) {
                double r = 0;
                for (int k = 0; k < 64; ++k) r += (double)fA[m * 64 + k] * fB[n * 64 + k];
                maxerr = fmax(maxerr, fabs(r - hD[m * N + n]));
                maxref = fmax(maxref, fabs(r));
            }
        printf("Q1 cta_group::2 M=256 N=%d K=64: max|err| = %.3e (max|ref| = %.3f)\n", N, maxerr, maxref);
        // timing: 2000 passes of 4 MMAs
        Probe2Args t{N, 2000};
        probe2_kernel<<<2, 128, smem>>>(dA, dB, dD, dC, t);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        long long cyc = 0;
        CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
        const double per = (double)cyc / (2000.0 * 4.0);
        printf("Q2 cta_group::2 M=256 N=%d: %.1f cycles per K=16 MMA -> %.0f flop/clk per SM (single-CTA M=128: ~%d cycles, %.0f)\n",
               N, per, 256.0 * N * 16 * 2 / per / 2, 64 + N / 2, 128.0 * N * 16 * 2 / (64 + N / 2));
        cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dC);
    }
    return 0;
}
