// Experiment: does a K-major SWIZZLE_128B UMMA operand descriptor work when its start address is offset by an
// arbitrary number of ROWS (not a multiple of the 8-row swizzle atom) inside a taller swizzled buffer?
// D[128x128] = A_view(o)[128x64] . B[128x64]^T, A_view(o) = rows o .. o+127 of a 256-row buffer.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I audio_style_transfer_b200/csrc \
//        scripts/experiments/umma_row_offset.cu -o gpurun_out/umma_row_offset && gpurun_out/umma_row_offset
#include "ptx.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>

__global__ void __launch_bounds__(128) k(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int o, int use_base_offset) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    uint8_t* gen = raw + (base - smem_u32(raw));
    uint8_t* sa = gen;                 // 256 rows x 128 B
    uint8_t* sb = gen + 32768;         // 128 rows x 128 B
    const uint32_t bar = base + 49152;
    for (int i = threadIdx.x; i < 256 * 8; i += 128) {
        const int r = i >> 3, c = i & 7;
        *reinterpret_cast<uint4*>(sa + sw128_off(r, c)) = *reinterpret_cast<const uint4*>(A + r * 64 + c * 8);
    }
    for (int i = threadIdx.x; i < 128 * 8; i += 128) {
        const int r = i >> 3, c = i & 7;
        *reinterpret_cast<uint4*>(sb + sw128_off(r, c)) = *reinterpret_cast<const uint4*>(B + r * 64 + c * 8);
    }
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init_fence(); }
    if (threadIdx.x < 32) tmem_alloc<128>(bar + 8);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + 49152 + 8);
    if (threadIdx.x == 0) {
        const uint32_t a0 = base + o * 128;
        for (int kk = 0; kk < 4; ++kk) {
            uint64_t ad = umma_desc_kmajor_sw128(a0 + kk * 32);
            if (use_base_offset) ad |= (uint64_t)((a0 >> 7) & 7) << 49;
            const uint64_t bd = umma_desc_kmajor_sw128(base + 32768 + kk * 32);
            umma_f16(tmem, ad, bd, umma_idesc_bf16(128, 128), kk > 0);
        }
        umma_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int cg = 0; cg < 4; ++cg) {
        float v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + cg * 32, v);
        for (int e = 0; e < 32; ++e) D[(warp * 32 + lane) * 128 + cg * 32 + e] = v[e];
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<128>(tmem);
}

int main() {
    std::vector<__nv_bfloat16> hA(256 * 64), hB(128 * 64);
    std::vector<float> fA(256 * 64), fB(128 * 64);
    srand(1);
    for (size_t i = 0; i < hA.size(); ++i) { float x = (rand() % 17 - 8) / 8.0f; hA[i] = __float2bfloat16(x); fA[i] = __bfloat162float(hA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { float x = (rand() % 13 - 6) / 4.0f; hB[i] = __float2bfloat16(x); fB[i] = __bfloat162float(hB[i]); }
    __nv_bfloat16 *dA, *dB; float* dD;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * 128 * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
    const int offs[] = {0, 8, 64, 1, 3, 4, 13, 100, 127};
    for (int ubo = 0; ubo < 2; ++ubo)
        for (int o : offs) {
            cudaMemset(dD, 0, 128 * 128 * 4);
            k<<<1, 128, 60000>>>(dA, dB, dD, o, ubo);
            cudaError_t e = cudaDeviceSynchronize();
            std::vector<float> hD(128 * 128);
            cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
            double maxerr = 0;
            for (int m = 0; m < 128; ++m)
                for (int n = 0; n < 128; ++n) {
                    double ref = 0;
                    for (int q = 0; q < 64; ++q) ref += (double)fA[(m + o) * 64 + q] * fB[n * 64 + q];
                    maxerr = fmax(maxerr, fabs(ref - hD[m * 128 + n]));
                }
            printf("base_offset_field=%d row_offset=%3d  max|err|=%g  %s\n", ubo, o, maxerr, cudaGetErrorString(e));
        }
    return 0;
}
