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

// The same MMA stream (single CTA, N = 128 or pair N = 256 style operand footprints) while eight other warps of the CTA
// move `side_bytes_per_mma` bytes per MMA slot through (1) st.shared + ld.shared, (2) st.global (128 B per thread,
// one line each, like a row-per-thread epilogue), (3) ld.global.nc of lines that hit the L2, (4) TMA bulk loads
// global -> shared.  Shows which epilogue / refill paths steal the tensor pipe's shared-memory reads.
template <int N>
__global__ void __launch_bounds__(384, 1) contend_kernel(int iters, int mode, int side_iters, uint4* gbuf, long long* out) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar, tbar;
    __shared__ uint32_t slot;
    for (int i = threadIdx.x; i < 192 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x < 32) {
        if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&tbar, 1); fence_mbar_init(); }
        __syncwarp();
        tmem_alloc(&slot, 256);
        tmem_relinquish();
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(128, N);
        const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 64 * 1024);
        uint64_t ad[4], bd[2];
        for (int i = 0; i < 4; ++i) ad[i] = umma_desc_kmajor(a0 + static_cast<uint32_t>(i * 16384), 128);
        for (int i = 0; i < 2; ++i) bd[i] = umma_desc_kmajor(b0 + static_cast<uint32_t>(i * 32768), 128);
        const long long t0 = clock64();
        for (int it = 0; it < iters; it += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16(tm, ad[u] + 2 * k, bd[u & 1] + 2 * k, idesc, 1u);
        }
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        out[blockIdx.x] = clock64() - t0;
    } else if (warp >= 4 && mode > 0) {
        // side traffic: thread-private 128-byte rows
        const int tid = threadIdx.x - 128;                     // 0..255
        uint8_t* sside = smem + 128 * 1024;                    // 32 KB scratch for modes 1 / 4
        uint4* grow = gbuf + (static_cast<size_t>(blockIdx.x) * 256 + tid) * 8;      // 128 B per thread per sweep
        const size_t sweep = static_cast<size_t>(gridDim.x) * 256 * 8;               // uint4 per sweep of the grid
        uint4 acc4 = make_uint4(tid, 1, 2, 3);
        const long long s0 = clock64();
        for (int it = 0; it < side_iters; ++it) {
            if (mode == 1) {
                const uint32_t base = smem_u32(sside) + static_cast<uint32_t>(tid * 128);
#pragma unroll
                for (int j = 0; j < 8; ++j) st_shared_v4(base + ((static_cast<uint32_t>(j) ^ (tid & 7)) << 4), acc4);
#pragma unroll
                for (int j = 0; j < 8; ++j) { const uint4 r = ld_shared_v4(base + ((static_cast<uint32_t>(j) ^ (tid & 7)) << 4)); acc4.x ^= r.y; }
            } else if (mode == 2) {
                uint4* dstp = grow + (it & 7) * sweep;
#pragma unroll
                for (int j = 0; j < 8; ++j) st_v4(dstp + j, acc4);
            } else if (mode == 3) {
                const uint4* srcp = grow + (it & 7) * sweep;
#pragma unroll
                for (int j = 0; j < 8; ++j) { const uint4 r = ld_nc_v4(srcp + j); acc4.x ^= r.y; }
            } else if (mode == 4) {
                if (tid == 0) {                                 // one 32 KB bulk copy global -> shared per iteration
                    mbar_arrive_expect_tx(&tbar, 32768);
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                     smem_u32(sside)), "l"(gbuf + (static_cast<size_t>(blockIdx.x) * 8 + (it & 7)) * 2048), "r"(32768u),
                                 "r"(smem_u32(&tbar)) : "memory");
                    mbar_wait(&tbar, static_cast<uint32_t>(it & 1));
                }
            }
        }
        if (tid == 0) out[gridDim.x + blockIdx.x] = clock64() - s0;
        if (acc4.x == 0x12345678u) out[2 * gridDim.x + threadIdx.x] = acc4.x;   // keep the loads alive
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) {
        tc_fence_after();
        tmem_dealloc(tm, 256);
    }
}

template <int N>
void run_contend(int sms, int iters, int mode, int side_iters, const char* what) {
    auto k = contend_kernel<N>;
    const int smem = 193 * 1024 + 1024;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    long long* d;
    cudaMalloc(&d, sizeof(long long) * (2 * sms + 512));
    uint4* g;
    const size_t gbytes = static_cast<size_t>(sms) * 256 * 128 * 8;   // 8 sweeps of 128 B per side thread: 38 MB, L2 resident
    cudaMalloc(&g, gbytes);
    cudaMemset(g, 1, gbytes);
    for (int rep = 0; rep < 2; ++rep) {
        k<<<sms, 384, smem>>>(iters, mode, side_iters, g, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("contend: %s\n", cudaGetErrorString(e)); return; }
    }
    long long h[512] = {0};
    cudaMemcpy(h, d, sizeof(long long) * 2 * sms, cudaMemcpyDeviceToHost);
    long long mx = 0, smx = 1;
    for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
    for (int i = 0; i < sms; ++i) smx = h[sms + i] > smx ? h[sms + i] : smx;
    const double total = static_cast<double>(mx);
    const double side_bytes = mode == 4 ? 32768.0 * side_iters : (mode == 1 ? 2.0 : 1.0) * 256 * 128.0 * side_iters;
    printf("single M=128 N=%3d + %-34s: %6.1f cycles per MMA over %7.0f cycles; side traffic ran %7.0f cycles at %5.1f B/clk/SM\n", N, what,
           total / (4.0 * iters), total, mode ? static_cast<double>(smx) : 0.0, mode ? side_bytes / static_cast<double>(smx) : 0.0);
    cudaFree(d);
    cudaFree(g);
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
    // side traffic sized to roughly 32-64 B/clk if it ran unhindered for the MMA stream's duration
    for (int n : {128, 256}) {
        const int side = n == 128 ? 1500 : 3000;
        auto go = [&](int mode, int si, const char* w) { if (n == 128) run_contend<128>(sms, iters, mode, si, w); else run_contend<256>(sms, iters, mode, si, w); };
        go(0, 0, "nothing");
        go(1, side, "st.shared + ld.shared rows");
        go(2, side * 2, "st.global rows (128 B / thread)");
        go(3, side * 2, "ld.global.nc rows (L2 hits)");
        go(4, side * 2, "bulk copy global->shared 32 KB");
    }
    return 0;
}
