// Where the 28x28x128 3x3 kernel's time goes: the kernel compiled with clock stamps (C3B_TRACE), launched alone on
// random data, per-role waiting time printed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DC3B_TRACE -I include -I mlmicroservicetemplate_b200/csrc \
//        -o benchmarks/micro/_build/c3b_trace benchmarks/micro/c3b_trace.cu -lcuda
#include "../../mlmicroservicetemplate_b200/csrc/conv3x3_c128.cu"
#include <vector>
namespace b2p {
cudaError_t launch_dummy() { return cudaSuccess; }
}
int main(int argc, char** argv) {
    using namespace b2p;
    const int n = argc > 1 ? atoi(argv[1]) : 256;
    size_t act = static_cast<size_t>(n) * 28 * 28 * 128 * 2;
    void *x, *y, *w;
    float* bias;
    cudaMalloc(&x, act); cudaMalloc(&y, act); cudaMalloc(&w, 128 * 9 * 128 * 2); cudaMalloc(&bias, 512);
    cudaMemset(x, 0x3c, act); cudaMemset(w, 0x3c, 128 * 9 * 128 * 2); cudaMemset(bias, 0, 512);
    Conv3x3bPlan p;
    p.weight = w; p.bias = bias;
    char err[256] = "";
    if (!conv3x3_c128_plan_set(p, x, y, n, err, sizeof err)) { printf("plan: %s\n", err); return 1; }
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        cudaError_t e = conv3x3_c128_launch(p, n, 1, 0, sms, 0);
        cudaEventRecord(e1);
        if (e != cudaSuccess || (e = cudaDeviceSynchronize()) != cudaSuccess) { printf("launch: %s\n", cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("launch %d: %.1f us\n", rep, ms * 1000);
    }
    static long long h[256][8];
    cudaMemcpyFromSymbol(h, g_c3b_trace, sizeof h);
    printf("CTA  tiles | MMA thread: total  wait tmem_empty  wait planes  wait weights | epilogue warp: total  wait tmem_full\n");
    for (int b = 0; b < 2 * (sms / 2); b += (b < 8 ? 1 : 24))
        printf("%3d  %5lld | %17lld %16lld %12lld %13lld | %20lld %15lld\n", b, h[b][6], h[b][3], h[b][0], h[b][1], h[b][2], h[b][5], h[b][4]);
    return 0;
}
