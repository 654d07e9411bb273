// umma_probe.cu — standalone hardware probe (no torch) for questions the conv kernels' next
// optimisation step depends on. It is a measurement tool, not part of the product path.
//
//   P1  K-major SWIZZLE_128B operand whose descriptor start address is shifted by whole 128 B rows
//       (not a multiple of 8 rows) and/or whose 8-row groups are SBO != 1024 B apart: does tcgen05.mma
//       read rows (shift + (r/8)*SBO/128 + r%8) with the swizzle phase of the ABSOLUTE smem address?
//       (=> a conv halo tile can be loaded once and reused by all 9 taps)
//   P2  same question for MN-major operands shifted along K (wgrad halo reuse).
//   P3  TMA tiled 4-D load with elementStrides = 2 and negative start coordinates.
//
// Prints one line per configuration: max |error| against the host model of the hypothesis.
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

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_enc;

static CUtensorMap map2d(void* base, int rows, int cols, int brows) {
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t str[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)brows};
    cuuint32_t es[2] = {1, 1};
    CUresult r = g_enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode2d failed %d\n", (int)r); exit(2); }
    return m;
}

// mode 0: K-major A (rows x 64 k) shifted by rows;  B K-major [64 n][64 k]
// mode 1: MN-major A ([k rows][128 m] as two 64-wide slabs), shifted along k; B MN-major [k rows][64 n]
struct ProbeArgs {
    int mode, shift_rows, sbo, use_base_offset, a_rows;
};

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapA2,
             const __grid_constant__ CUtensorMap mapB, float* out, ProbeArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                // up to 256 rows x 128 B = 32 KB (mode 1: two slabs of a_rows x 128 B)
    uint8_t* sB = smem + 65536;        // 64 rows x 128 B (mode 1: a_rows? no: 64 k rows x 64 n)
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 65536 + 16384);
    uint64_t* bar2 = bar + 1;
    uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(bar2, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        __syncwarp();
        tmem_alloc<64>(tptr);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tptr;
    if (threadIdx.x == 0) {
        if (a.mode == 0) {
            mbar_arrive_expect_tx(bar, a.a_rows * 128 + 64 * 128);
            tma_load_2d(sA, &mapA, bar, 0, 0);
            tma_load_2d(sB, &mapB, bar, 0, 0);
        } else {
            mbar_arrive_expect_tx(bar, 2 * a.a_rows * 128 + 64 * 128);
            tma_load_2d(sA, &mapA, bar, 0, 0);                      // m 0..63
            tma_load_2d(sA + a.a_rows * 128, &mapA2, bar, 64, 0);   // m 64..127
            tma_load_2d(sB, &mapB, bar, 0, 0);
        }
        mbar_wait(bar, 0, 100);
        tc_fence_after();
        if (a.mode == 0) {
            const uint32_t start = smem_u32(sA) + a.shift_rows * 128;
            const uint32_t bo = a.use_base_offset ? ((start >> 7) & 7u) : 0u;
            const uint64_t ad = make_smem_desc_sw128(start, 16, a.sbo, bo);
            const uint64_t bd = make_smem_desc_sw128(smem_u32(sB), 16, 1024);
            const uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
            for (int k = 0; k < 4; ++k) umma_bf16(tmem, ad + 2u * k, bd + 2u * k, idesc, k != 0);
        } else {
            const uint32_t start = smem_u32(sA) + a.shift_rows * 128;
            const uint32_t bo = a.use_base_offset ? ((start >> 7) & 7u) : 0u;
            const uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);
            for (int k = 0; k < 4; ++k) {
                const uint64_t ad = make_smem_desc_sw128(start + k * 2048, a.a_rows * 128, a.sbo, bo);
                const uint64_t bd = make_smem_desc_sw128(smem_u32(sB) + k * 2048, 8192, 1024);
                umma_bf16(tmem, ad, bd, idesc, k != 0);
            }
        }
        umma_commit(bar2);
    }
    __syncwarp();
    mbar_wait(bar2, 0, 101);
    tc_fence_after();
    for (int c = 0; c < 2; ++c) {
        float v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, v);
        for (int e = 0; e < 32; ++e) out[(warp * 32 + lane) * 64 + c * 32 + e] = v[e];
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) tmem_dealloc<64>(tmem);
}

// P3: dump a 4-D TMA box (128 pixels x 64 ch, SWIZZLE_128B) to global memory, un-swizzled.
__global__ void tma_dump_kernel(const __grid_constant__ CUtensorMap map, __nv_bfloat16* out, int c1, int c2, int c3) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384);
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar, 16384);
        tma_load_4d(smem, &map, bar, 0, c1, c2, c3);
    }
    mbar_wait(bar, 0, 102);
    const int row = threadIdx.x;  // 128 threads
    for (int ch = 0; ch < 8; ++ch) {
        const int phys = ch ^ (row & 7);
        const uint4 v = *reinterpret_cast<const uint4*>(smem + row * 128 + phys * 16);
        *reinterpret_cast<uint4*>(out + row * 64 + ch * 8) = v;
    }
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
    CK(cudaSetDevice(0));
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    g_enc = (EncodeTiledFn)p;

    const int AR = 256;
    std::vector<float> hA(AR * 128), hB(64 * 64);
    std::vector<__nv_bfloat16> bA(AR * 128), bB(64 * 64);
    srand(1);
    for (size_t i = 0; i < hA.size(); ++i) { hA[i] = bf((rand() % 2001 - 1000) / 1000.f); bA[i] = __float2bfloat16(hA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { hB[i] = bf((rand() % 2001 - 1000) / 1000.f); bB[i] = __float2bfloat16(hB[i]); }
    __nv_bfloat16 *dA, *dB;
    float* dD;
    CK(cudaMalloc(&dA, bA.size() * 2));
    CK(cudaMalloc(&dB, bB.size() * 2));
    CK(cudaMalloc(&dD, 128 * 64 * 4));
    CK(cudaMemcpy(dA, bA.data(), bA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, bB.data(), bB.size() * 2, cudaMemcpyHostToDevice));
    const int smem_bytes = 1024 + 65536 + 16384 + 64;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    std::vector<float> hD(128 * 64);

    // ---- P1: K-major. A viewed as [256 rows][64 k] (first 64 columns of the 128-wide host array: use ld 128)
    {
        // device matrix has 128 columns; K-major tile uses columns 0..63
        CUtensorMap mA = map2d(dA, AR, 128, 256);
        CUtensorMap mB = map2d(dB, 64, 64, 64);
        const int cfgs[][3] = {{0, 1024, 0}, {1, 1024, 0}, {1, 1024, 1}, {3, 1024, 0}, {3, 1024, 1}, {8, 1024, 0},
                               {34, 1024, 0}, {34, 1024, 1}, {0, 1280, 0}, {0, 1280, 1}, {1, 1280, 0}, {1, 1280, 1},
                               {11, 1280, 0}, {11, 1280, 1}, {0, 2048, 0}, {5, 2048, 0}, {5, 2048, 1}};
        for (auto& c : cfgs) {
            ProbeArgs a{0, c[0], c[1], c[2], AR};
            CK(cudaMemset(dD, 0, 128 * 64 * 4));
            probe_kernel<<<1, 128, smem_bytes>>>(mA, mA, mB, dD, a);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("P1 shift=%d sbo=%d bo=%d : CUDA error %s\n", c[0], c[1], c[2], cudaGetErrorString(e)); return 3; }
            CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
            double maxerr = 0;
            bool inrange = true;
            for (int r = 0; r < 128; ++r) {
                const int row = c[0] + (r / 8) * (c[1] / 128) + r % 8;
                if (row >= AR) { inrange = false; continue; }
                for (int n = 0; n < 64; ++n) {
                    double s = 0;
                    for (int k = 0; k < 64; ++k) s += (double)hA[row * 128 + k] * hB[n * 64 + k];
                    maxerr = fmax(maxerr, fabs(s - hD[r * 64 + n]));
                }
            }
            printf("P1 kmajor shift_rows=%d sbo=%d base_offset=%d : max_abs_err=%.5f %s%s\n", c[0], c[1], c[2], maxerr,
                   maxerr < 1e-2 ? "MATCH" : "MISMATCH", inrange ? "" : " (partial)");
        }
    }
    // ---- P2: MN-major. A = [k rows][128 m] (two 64-col slabs), shifted along k rows; B = [64 k rows][64 n]
    {
        const int a_rows = 128;  // k rows loaded per slab
        CUtensorMap mA = map2d(dA, AR, 128, a_rows);
        CUtensorMap mB = map2d(dB, 64, 64, 64);
        const int cfgs[][3] = {{0, 1024, 0}, {1, 1024, 0}, {1, 1024, 1}, {5, 1024, 0}, {5, 1024, 1}, {8, 1024, 0}, {34, 1024, 0}, {34, 1024, 1}};
        for (auto& c : cfgs) {
            ProbeArgs a{1, c[0], c[1], c[2], a_rows};
            CK(cudaMemset(dD, 0, 128 * 64 * 4));
            probe_kernel<<<1, 128, smem_bytes>>>(mA, mA, mB, dD, a);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("P2 shift=%d : CUDA error %s\n", c[0], cudaGetErrorString(e)); return 3; }
            CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
            double maxerr = 0;
            for (int m = 0; m < 128; ++m)
                for (int n = 0; n < 64; ++n) {
                    double s = 0;
                    for (int k = 0; k < 64; ++k) s += (double)hA[(k + c[0]) * 128 + m] * hB[k * 64 + n];
                    maxerr = fmax(maxerr, fabs(s - hD[m * 64 + n]));
                }
            printf("P2 mnmajor shift_k=%d sbo=%d base_offset=%d : max_abs_err=%.5f %s\n", c[0], c[1], c[2], maxerr,
                   maxerr < 1e-2 ? "MATCH" : "MISMATCH");
        }
    }
    // ---- P3: 4-D TMA, element stride 2, negative coordinates. Tensor [N=2, H=16, W=16, C=64].
    {
        const int N = 2, H = 16, W = 16, C = 64;
        std::vector<__nv_bfloat16> hx(N * H * W * C);
        for (int i = 0; i < N * H * W * C; ++i) hx[i] = __float2bfloat16((float)((i / C) % 251) + (i % C) * 0.0f + 1.f);
        __nv_bfloat16 *dx, *dout;
        CK(cudaMalloc(&dx, hx.size() * 2));
        CK(cudaMalloc(&dout, 128 * 64 * 2));
        CK(cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice));
        // output grid 8x8 per image, box = 8 x 8 x 2 images = 128 pixels; stride 2
        CUtensorMap m;
        cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
        cuuint64_t str[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
        cuuint32_t box[4] = {64, 16, 16, 2};
        cuuint32_t es[4] = {1, 2, 2, 1};
        CUresult r = g_enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, dx, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("P3 encode failed: %d\n", (int)r); return 0; }
        CK(cudaFuncSetAttribute(tma_dump_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + 16384 + 64));
        std::vector<__nv_bfloat16> ho(128 * 64);
        for (int off = -1; off <= 1; ++off) {
            CK(cudaMemset(dout, 0xff, 128 * 64 * 2));
            tma_dump_kernel<<<1, 128, 1024 + 16384 + 64>>>(m, dout, off, off, 0);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("P3 off=%d : CUDA error %s\n", off, cudaGetErrorString(e)); return 3; }
            CK(cudaMemcpy(ho.data(), dout, ho.size() * 2, cudaMemcpyDeviceToHost));
            int bad = 0;
            for (int row = 0; row < 128; ++row) {
                const int n = row / 64, i = (row / 8) % 8, j = row % 8;
                const int y = 2 * i + off, x = 2 * j + off;
                float exp = 0.f;
                if (y >= 0 && y < H && x >= 0 && x < W) exp = __bfloat162float(hx[((n * H + y) * W + x) * C]);
                const float got = __bfloat162float(ho[row * 64]);
                if (got != exp) { if (bad < 4) printf("   row %d (n%d i%d j%d): got %.1f exp %.1f\n", row, n, i, j, got, exp); ++bad; }
            }
            printf("P3 tma stride2 start=%d : %s (%d bad rows)\n", off, bad ? "MISMATCH" : "MATCH", bad);
        }
    }
    printf("probe done\n");
    return 0;
}
