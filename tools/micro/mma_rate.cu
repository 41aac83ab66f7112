// Microbenchmark: cycles per tcgen05.mma (SS, bf16, M=128, K=16) for several N, issued back to
// back from one elected lane on fixed shared-memory operands. Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I celebrity-image-denoiser_b200/csrc -o /tmp/mma_rate tools/micro/mma_rate.cu
#include <cstdio>
#include "ptx.cuh"
using namespace cid;

template <int N, int DISTINCT>
__global__ void __launch_bounds__(128, 1) rate_kernel(long long* out, int iters) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(&slot, 512);
    fence_proxy_async();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tmem = slot;
    if (warp == 0) {
        const uint32_t a_lo = umma_desc_lo(smem_u32(smem));
        const uint32_t b_lo = umma_desc_lo(smem_u32(smem + 32768));
        constexpr uint32_t idesc = umma_idesc_bf16(128, N);
        long long t0 = 0, t1 = 0;
        for (int rep = 0; rep < 2; ++rep) {
            t0 = clock64();
            if (elect_one()) {
                for (int i = 0; i < iters; ++i) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        // DISTINCT: walk A through different 1 KB row groups like the conv taps do
                        const uint32_t ao = DISTINCT ? ((i & 7) * 64 + 2 * k) : 2 * k;
                        umma_bf16_ss(tmem + (i & 1) * N, umma_desc_join<1024>(a_lo + ao), umma_desc_join<1024>(b_lo + 2 * k), idesc, 1u);
                    }
                }
                umma_commit(&bar);
            }
            __syncwarp();
            mbar_wait(&bar, rep & 1);
            t1 = clock64();
        }
        if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

template <int N, int D>
void run(int blocks, int iters) {
    long long* d; cudaMalloc(&d, blocks * 8);
    auto k = rate_kernel<N, D>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    k<<<blocks, 128, 100 * 1024>>>(d, iters);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, blocks * 8, cudaMemcpyDeviceToHost);
    double cyc = double(h[0]) / (iters * 4);
    printf("N=%3d distinctA=%d blocks=%3d: %.1f cycles/MMA (ideal %d) -> %.0f%% of peak  [%s]\n", N, D, blocks, cyc, N / 2,
           100.0 * (N / 2) / cyc, cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    for (int blocks : {1, 148}) {
        run<16, 0>(blocks, 2000); run<64, 0>(blocks, 2000); run<128, 0>(blocks, 2000); run<256, 0>(blocks, 2000);
        run<64, 1>(blocks, 2000); run<256, 1>(blocks, 2000);
    }
    return 0;
}
