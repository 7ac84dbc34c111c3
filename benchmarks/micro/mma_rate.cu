// Issue rate of tcgen05.mma (kind::f16, bf16 operands resident in shared memory, 128-byte swizzle) for the tile shapes the
// conv kernels use: one CTA (M = 128) and a CTA pair (cta_group::2, M = 256), N in {64, 128, 256}.  One CTA / pair per
// SM, every SM busy; operands are never reloaded, so what is measured is the tensor pipe + its shared-memory reads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I mlmicroservicetemplate_b200/csrc -o gpurun_out/mma_rate benchmarks/micro/mma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include "ptx.cuh"
using namespace b2p;

template <int N, bool PAIR>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, int a_stride_tiles, int a_shift_rows, long long* out) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    if (threadIdx.x < 32) {
        if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
        __syncwarp();
        if (PAIR) { tmem_alloc_pair(&slot, 256); tmem_relinquish_pair(); } else { tmem_alloc(&slot, 256); tmem_relinquish(); }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tm = slot;
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0 && rank == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(PAIR ? 256 : 128, N);
        const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 96 * 1024);
        // walk 4 A tiles (16 KB each = one 64-wide k-block of 128 rows) and 2 B tiles like a k loop does
        uint64_t ad[4], bd[2];
        for (int i = 0; i < 4; ++i) ad[i] = umma_desc_kmajor(a0 + static_cast<uint32_t>((i % a_stride_tiles) * 16384 + a_shift_rows * 128), 128);
        for (int i = 0; i < 2; ++i) bd[i] = umma_desc_kmajor(b0 + static_cast<uint32_t>(i * 32768), 128);
        t0 = clock64();
        for (int it = 0; it < iters; it += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (PAIR) umma_bf16_pair(tm, ad[u] + 2 * k, bd[u & 1] + 2 * k, idesc, 1u);
                    else umma_bf16(tm, ad[u] + 2 * k, bd[u & 1] + 2 * k, idesc, 1u);
                }
            }
        }
        if (PAIR) umma_commit_pair(&bar); else umma_commit(&bar);
        mbar_wait(&bar, 0);
        t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    } else if (PAIR && threadIdx.x == 0) {
        mbar_wait(&bar, 0);
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();
    if (threadIdx.x < 32) {
        tc_fence_after();
        if (PAIR) tmem_dealloc_pair(tm, 256); else tmem_dealloc(tm, 256);
    }
}

template <int N, bool PAIR>
void run(int sms, int iters, int shift = 0) {
    auto k = rate_kernel<N, PAIR>;
    const int smem = 161 * 1024 + 1024;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    long long* d;
    cudaMalloc(&d, sizeof(long long) * sms);
    cudaMemset(d, 0, sizeof(long long) * sms);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(PAIR ? (sms & ~1) : sms);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PAIR ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    for (int rep = 0; rep < 2; ++rep) {
        cudaError_t e = cudaLaunchKernelEx(&cfg, k, iters, 4, shift, d);
        if (e != cudaSuccess) { printf("launch: %s\n", cudaGetErrorString(e)); return; }
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("sync: %s\n", cudaGetErrorString(e)); return; }
    }
    long long h[256] = {0};
    cudaMemcpy(h, d, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
    const double per = static_cast<double>(mx) / (4.0 * iters);
    const double macs_per_sm = 128.0 * N * 16;      // per CTA and MMA (a pair's MMA does twice that over two SMs)
    printf("%s M=%d N=%3d A start +%2d rows: %7.1f cycles per K=16 MMA  -> %6.0f MAC/clk/SM  (smem operand bytes per SM and MMA: %d)\n",
           PAIR ? "pair  " : "single", PAIR ? 256 : 128, N, shift, per, macs_per_sm / per, 128 * 32 + (PAIR ? N / 2 : N) * 32);
    cudaFree(d);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 4000;
    run<32, false>(sms, iters);
    run<64, false>(sms, iters);
    run<96, false>(sms, iters);
    run<128, false>(sms, iters);
    run<160, false>(sms, iters);
    run<192, false>(sms, iters);
    run<256, false>(sms, iters);
    run<64, true>(sms, iters);
    run<128, true>(sms, iters);
    run<192, true>(sms, iters);
    // the patch-resident 3x3 kernels start the A descriptor s rows (s * 128 bytes) into a swizzled tile
    for (int shift : {1, 2, 4, 8, 33, 66}) run<128, true>(sms, iters, shift);
    for (int shift : {1, 2, 8}) run<64, false>(sms, iters, shift);
    for (int shift : {1, 2}) run<128, false>(sms, iters, shift);
    run<256, true>(sms, iters);
    return 0;
}
