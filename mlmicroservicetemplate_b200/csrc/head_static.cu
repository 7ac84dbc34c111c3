// b2p_head_static: the classifier head (global average pool -> linear) for callers that are not
// allowed to allocate device memory.
//
// The regular path runs these two launches out of a layer program: b2p_create() cudaMalloc's a
// workspace slot for the pooled features that sit between them.  That is a problem for a host
// process that has handed all of the GPU's memory to its own pool (a raw cudaMalloc from inside a
// library then fails even though the pool has room), for one that must not touch the allocator on
// its request path because cudaMalloc/cudaFree serialise against everything running on the device,
// and for one that works under a fixed memory budget agreed at start-up.  For those callers this
// entry point keeps the pooled features in a fixed array that is part of the library image itself
// and performs no allocation of any kind: no cudaMalloc/cudaFree, no handle, no torch.
//
// Limits: n <= B2P_HEAD_STATIC_MAX_IMAGES (256) images and c <= B2P_HEAD_STATIC_MAX_CHANNELS (2048)
// channels per call, i.e. 1 MiB of bf16 scratch.  Every GPU that loads the library carries that
// 1 MiB until the process exits, whether or not this function is ever called.
//
// Sharing: there is ONE array per device and every call on that device uses it.  It is never
// cleared and keeps whatever the previous call left; that is harmless here because a call
// overwrites rows [0, n) before reading them and never reads beyond row n (the GEMM's operand
// descriptor is encoded for exactly n rows, so the tail of the last 128-row tile is zero-filled by
// the TMA unit, not read from the array).  What is NOT safe is overlap: two calls in flight on
// the same device at the same time (different streams, or different host threads) race on the
// array.  Order them on one stream or with events.
//
// First call on a device: resolves the device address of the array (cudaGetSymbolAddress, with
// that device current), looks up the driver's tensor-map encoder and raises the kernel's dynamic
// shared-memory limit.  Make that first call outside of any stream capture.
#include <mutex>
#include "kernels.h"

// Pooled features, bf16 [n][c].  In host code this name is only a stand-in for the symbol — it is
// NOT a device address; the address each device gave the array is looked up once below.
static __device__ __align__(256) uint16_t g_head_scratch[B2P_HEAD_STATIC_MAX_IMAGES * B2P_HEAD_STATIC_MAX_CHANNELS];

namespace {

constexpr int kMaxImages = B2P_HEAD_STATIC_MAX_IMAGES;
constexpr int kMaxChannels = B2P_HEAD_STATIC_MAX_CHANNELS;
constexpr int kMaxDevices = 64;

struct DeviceSetup {
    bool ready = false;
    void* scratch = nullptr;    // device address of g_head_scratch on this device
    int num_sms = 0;
};
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_mutex;

int32_t setup_current_device(int dev, DeviceSetup** out) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    DeviceSetup& s = g_setup[dev];
    if (!s.ready) {
        int major = 0;
        cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (e != cudaSuccess || major != 10) {
            b2p::set_global_error("b2p_head_static: device %d is not sm_100 (no fallback exists)", dev);
            return B2P_E_DEVICE;
        }
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_head_scratch);      // device `dev` is current here
        if (e != cudaSuccess || ptr == nullptr) {
            b2p::set_global_error("b2p_head_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                                  cudaGetErrorString(e));
            return B2P_E_CUDA;
        }
        e = cudaDeviceGetAttribute(&s.num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) {
            b2p::set_global_error("b2p_head_static: cannot query the SM count: %s", cudaGetErrorString(e));
            return B2P_E_CUDA;
        }
        char eb[256] = "";
        if (!b2p::tma_init(eb, sizeof eb)) {
            b2p::set_global_error("b2p_head_static: %s", eb);
            return B2P_E_CUDA;
        }
        s.scratch = ptr;
        s.ready = true;
    }
    *out = &s;
    return B2P_OK;
}

}  // namespace

extern "C" int32_t b2p_head_static(const void* feat, int32_t n, int32_t hw, int32_t c, const void* weight,
                                   int32_t weight_rows, const float* bias, int32_t classes, float* logits,
                                   void* stream) {
    if (!feat || !weight || !bias || !logits || n <= 0 || hw <= 0 || c <= 0 || classes <= 0) {
        b2p::set_global_error("b2p_head_static: null pointer or non-positive size");
        return B2P_E_ARG;
    }
    if (n > kMaxImages || c > kMaxChannels) {
        b2p::set_global_error(
            "b2p_head_static: n=%d, c=%d exceed the fixed scratch (at most %d images x %d channels per call); "
            "split the batch or use b2p_forward, which allocates its workspace in b2p_create",
            n, c, kMaxImages, kMaxChannels);
        return B2P_E_LIMIT;
    }
    if (c % 64 != 0 || weight_rows % 128 != 0 || weight_rows < classes) {
        b2p::set_global_error("b2p_head_static: need c %% 64 == 0 and weight_rows a multiple of 128 >= classes");
        return B2P_E_ARG;
    }
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        b2p::set_global_error("b2p_head_static: no current CUDA device");
        return B2P_E_DEVICE;
    }
    DeviceSetup* s = nullptr;
    int32_t rc = setup_current_device(dev, &s);
    if (rc != B2P_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // 1) global average pool: feat [n, hw, c] bf16 -> scratch [n, c] bf16
    cudaError_t e = b2p::launch_avgpool(feat, s->scratch, n, hw, c, st);
    if (e != cudaSuccess) {
        b2p::set_global_error("b2p_head_static: pooling launch failed: %s", cudaGetErrorString(e));
        return B2P_E_CUDA;
    }
    // 2) linear: scratch [n, c] x weight[weight_rows, c]^T + bias -> logits [n, classes] fp32
    b2p_op op = {};
    op.kind = B2P_OP_LINEAR;
    op.in_c = c;
    op.out_c = classes;
    op.src2 = B2P_SLOT_NONE;
    b2p::GemmPlan plan;
    char eb[256] = "";
    if (!b2p::gemm_plan_init(plan, op, weight, static_cast<int64_t>(weight_rows) * c * 2, bias, n, s->num_sms, eb,
                             sizeof eb) ||
        !b2p::gemm_plan_set_src(plan, s->scratch, n, eb, sizeof eb)) {
        b2p::set_global_error("b2p_head_static: %s", eb);
        return B2P_E_CUDA;
    }
    e = b2p::gemm_plan_launch(plan, n, false, logits, s->num_sms, st);
    if (e != cudaSuccess) {
        b2p::set_global_error("b2p_head_static: GEMM launch failed: %s", cudaGetErrorString(e));
        return B2P_E_CUDA;
    }
    return B2P_OK;
}
