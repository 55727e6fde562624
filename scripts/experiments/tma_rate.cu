// Experiment: per-SM TMA load throughput for one [128 rows x 64 bf16] operand tile (16 KB), L2-resident source.
//   A: tensor map over [rows][128] bf16 (row pitch 256 B), box {64,128}: every other 128-byte line  (what the layer
//      kernels do today for activations AND weights)
//   B: tensor map over [rows][64] bf16 (row pitch 128 B), box {64,128}: one contiguous 16 KB block
//   C: cp.async.bulk (1-D), 16 KB contiguous
// DEPTH loads in flight per CTA, 148 CTAs (1 per SM), each walks its own region of an 8 MB buffer.
#include "ptx.cuh"
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int MODE, int DEPTH>
__global__ void __launch_bounds__(128) k(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                        const uint8_t* buf, int iters, int rows_total, unsigned long long* out) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t bar = base + DEPTH * 16384;
    if (threadIdx.x == 0) {
        for (int i = 0; i < DEPTH; ++i) mbar_init(bar + 8 * i, 1);
        mbar_init_fence();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const int row_base = (blockIdx.x * 416) % (rows_total - 128 * 8);   // each CTA its own region, unaligned to tiles
        unsigned long long t0 = clock64();
        for (int it = 0; it < iters + DEPTH; ++it) {
            const int s = it % DEPTH;
            if (it >= DEPTH) mbar_wait(bar + 8 * s, ((it / DEPTH) - 1) & 1);
            if (it < iters) {
                mbar_expect_tx(bar + 8 * s, 16384);
                const int row = row_base + (it % 7) * 128 + (it % 5);       // shifted rows like the dilated taps
                if (MODE == 0) tma_load_2d(base + s * 16384, &tmA, bar + 8 * s, (it & 1) * 64, row);
                if (MODE == 1) tma_load_2d(base + s * 16384, &tmB, bar + 8 * s, 0, row + (it & 1) * (rows_total / 2));
                if (MODE == 2) bulk_load_1d(base + s * 16384, buf + (size_t)(row + (it & 1) * (rows_total / 2)) * 128, 16384, bar + 8 * s);
            }
        }
        unsigned long long t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    }
}

int main() {
    void* fn = nullptr; cudaDriverEntryPointQueryResult qres;
    cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &fn, 12000, cudaEnableDefault, &qres);
    auto enc = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    const int rows = 32768;                       // [32768][128] bf16 = 8 MB
    uint8_t* buf; cudaMalloc(&buf, (size_t)rows * 256); cudaMemset(buf, 1, (size_t)rows * 256);
    unsigned long long* out; cudaMalloc(&out, 148 * 8);
    CUtensorMap tmA, tmB;
    { const uint64_t dims[2] = {128, (uint64_t)rows}; const uint64_t str[1] = {256}; const uint32_t box[2] = {64, 128}; const uint32_t es[2] = {1, 1};
      enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
    { const uint64_t dims[2] = {64, (uint64_t)rows * 2}; const uint64_t str[1] = {128}; const uint32_t box[2] = {64, 128}; const uint32_t es[2] = {1, 1};
      enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
    const int iters = 2000;
    auto run = [&](auto kern, const char* name, int depth, int grid) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, depth * 16384 + 2048);
        for (int rep = 0; rep < 2; ++rep) kern<<<grid, 128, depth * 16384 + 2048>>>(tmA, tmB, buf, iters, rows, out);
        cudaError_t e = cudaDeviceSynchronize();
        unsigned long long h[148]; cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid;
        printf("%-34s depth %d grid %3d: %7.0f cycles per 16 KB tile = %6.1f B/clk/SM   (%s)\n", name, depth, grid, avg / iters, 16384.0 * iters / avg, cudaGetErrorString(e));
    };
    for (int grid : {148, 1}) {
        run(k<0, 1>, "A: pitch 256 B (strided lines)", 1, grid); run(k<0, 2>, "A: pitch 256 B (strided lines)", 2, grid);
        run(k<0, 4>, "A: pitch 256 B (strided lines)", 4, grid); run(k<0, 8>, "A: pitch 256 B (strided lines)", 8, grid);
        run(k<1, 1>, "B: pitch 128 B (contiguous)", 1, grid); run(k<1, 2>, "B: pitch 128 B (contiguous)", 2, grid);
        run(k<1, 4>, "B: pitch 128 B (contiguous)", 4, grid); run(k<1, 8>, "B: pitch 128 B (contiguous)", 8, grid);
        run(k<2, 1>, "C: cp.async.bulk 1-D 16 KB", 1, grid); run(k<2, 2>, "C: cp.async.bulk 1-D 16 KB", 2, grid);
        run(k<2, 4>, "C: cp.async.bulk 1-D 16 KB", 4, grid); run(k<2, 8>, "C: cp.async.bulk 1-D 16 KB", 8, grid);
    }
    return 0;
}
