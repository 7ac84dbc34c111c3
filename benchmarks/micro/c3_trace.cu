// Where the 56x56x64 3x3 kernel's time goes: the kernel compiled with clock stamps (C3_TRACE), launched alone.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DC3_TRACE -I include -I mlmicroservicetemplate_b200/csrc \
//        -o benchmarks/micro/_build/c3_trace benchmarks/micro/c3_trace.cu -lcuda
#include "../../mlmicroservicetemplate_b200/csrc/conv3x3_c64.cu"
int main(int argc, char** argv) {
    using namespace b2p;
    const int n = argc > 1 ? atoi(argv[1]) : 256;
    size_t act = static_cast<size_t>(n) * 56 * 56 * 64 * 2;
    void *x, *y, *w;
    float* bias;
    cudaMalloc(&x, act); cudaMalloc(&y, act); cudaMalloc(&w, 64 * 9 * 64 * 2); cudaMalloc(&bias, 256);
    cudaMemset(x, 0x3c, act); cudaMemset(w, 0x3c, 64 * 9 * 64 * 2); cudaMemset(bias, 0, 256);
    Conv3x3Plan p;
    p.weight = w; p.bias = bias;
    char err[256] = "";
    if (!conv3x3_c64_plan_set(p, x, y, nullptr, n, err, sizeof err)) { printf("plan: %s\n", err); return 1; }
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    static long long h[256][8];
    for (int rep = 0; rep < 3; ++rep) {
        memset(h, 0, sizeof h);
        cudaMemcpyToSymbol(g_c3_trace, h, sizeof h);
        cudaEventRecord(e0);
        cudaError_t e = conv3x3_c64_launch(p, n, 1, 0, sms, 0);
        cudaEventRecord(e1);
        if (e != cudaSuccess || (e = cudaDeviceSynchronize()) != cudaSuccess) { printf("launch: %s\n", cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("launch %d: %.1f us\n", rep, ms * 1000);
    }
    cudaMemcpyFromSymbol(h, g_c3_trace, sizeof h);
    printf("CTA tiles | MMA thread: total  wait tmem_empty  wait patch | epilogue groups: wait tmem_full g0, g1\n");
    for (int b = 0; b < sms; b += (b < 4 ? 1 : 36))
        printf("%3d %5lld | %17lld %16lld %11lld | %20lld %10lld\n", b, h[b][2], h[b][3], h[b][0], h[b][1], h[b][4], h[b][5]);
    return 0;
}
