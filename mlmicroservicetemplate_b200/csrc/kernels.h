// Internal C++ interface between the C-ABI layer (api.cu) and the kernel files.
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../include/b200predict.h"

namespace b2p {

// Launch with programmatic stream serialization: the kernel may become resident while its
// predecessor in the stream is still draining.  Every kernel launched through this helper MUST
// execute griddepcontrol.wait (ptx.cuh: pdl_wait) before it touches global memory.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ------------------------------------------------------------------ conv / GEMM (tcgen05)
struct GemmPlan {
    CUtensorMap tmap_a;   // activations: 2-D tiled [rows, K] or 4-D im2col over NHWC
    CUtensorMap tmap_b;   // weights [out_c_pad, K] bf16, K-major
    CUtensorMap tmap_out; // bf16 output [rows, out_c], 128 x 64 store boxes (unused by the fp32 epilogue)
    CUtensorMap tmap_res; // bf16 residual [rows, out_c], same boxes (only when the op has a residual)
    CUtensorMap tmap_a2;  // optional second A source (projection shortcut input), im2col 1x1 / src2_stride
    CUtensorMap tmap_b64; // weights again with 64-row boxes: the 64-column pieces of a split last wave
    int tail_split = 0;   // allow cutting the tiles of a nearly empty last wave into 64-column pieces
    int src2_c = 0, src2_stride = 1;
    int block_n = 0;      // 64 / 128 / 256
    int force_block_n = 0;  // set before gemm_plan_init to pin the tile width (autotuner)
    int cta2 = 0;         // run as CTA pairs: 256-row tiles, each CTA loads half of the weight tile (cta_group::2)
    int block_k = 64;     // 64 (128-byte swizzle) or 16 (32-byte swizzle, the stem)
    int im2col = 0;       // A operand through TMA im2col mode
    int out_f32 = 0;      // fp32 [M, ldc] output (LINEAR) instead of bf16 NHWC
    int taps_h = 1, taps_w = 1, k_chunks = 1;
    int in_h = 0, in_w = 0, in_c = 0;
    int out_h = 0, out_w = 0;
    int stride = 1, pad_lo = 0, pad_hi = 0;
    int n_real = 0;       // real output channels (columns >= n_real are not stored)
    int n_pad = 0;        // weight rows (multiple of block_n)
    int ldc = 0;          // output row pitch in elements
    int relu = 0;
    int reverse = 0;      // process tiles last-to-first: consecutive layers alternate direction so that a layer
                          // starts on the rows its predecessor wrote last (still resident in the 126 MB L2)
    const float* bias = nullptr;
    const void* weight = nullptr;
    // L2 cache policies of the activation / residual loads and the output store (ptx.cuh L2_EVICT_*; the default is
    // the ordinary policy).  api.cu marks operands at their LAST use as evict-first.
    uint64_t pol_a = 0x1000000000000000ull, pol_res = 0x1000000000000000ull, pol_out = 0x1000000000000000ull;
};

// Resolves the driver's tensor-map encoders once per process. Returns false if unavailable.
bool tma_init(char* err, size_t err_len);

// Fills everything except tmap_a (which depends on the source pointer).
bool gemm_plan_init(GemmPlan& p, const b2p_op& op, const void* weight_dev, int64_t weight_bytes,
                    const float* bias_dev, int rows_at_capacity, int num_sms, char* err,
                    size_t err_len);
// (Re)encodes the A-operand tensor map for a source buffer holding `batch_cap` images.
bool gemm_plan_set_src(GemmPlan& p, const void* src, int batch_cap, char* err, size_t err_len);
// (Re)encode the output / residual tensor maps for buffers with `rows` rows of out_c bf16 each.
// TMA clips stores beyond `rows`, so `rows` must not exceed what the buffer really holds.
bool gemm_plan_set_src2(GemmPlan& p, const void* src2, int batch_cap, char* err, size_t err_len);
bool gemm_plan_set_dst(GemmPlan& p, const void* dst, int rows, char* err, size_t err_len);
bool gemm_plan_set_res(GemmPlan& p, const void* res, int rows, char* err, size_t err_len);
// Launches the persistent kernel for `m_rows` output rows.  `out_f32` is only read by LINEAR plans.
cudaError_t gemm_plan_launch(const GemmPlan& p, int m_rows, bool has_res, void* out_f32, int num_sms,
                             cudaStream_t stream);

// ------------------------------------------------------------------ dedicated stem (+ max pool) kernel
struct StemPlan {
    CUtensorMap tmap_in;    // s2d input [n,112,112,16] bf16, 4-D tiled, 8-channel boxes
    CUtensorMap tmap_out;   // [rows, 64] bf16: one conv row (112) or one pooled row (56) per store
    const uint8_t* w_image = nullptr;   // 32 KB shared-memory image of the weights (device)
    const float* bias = nullptr;
    int pool = 0;
};
// Re-lays the packed stem weight [64][4][4][16] into the [tap][k half][out ch][8] shared-memory image.
void stem_pack_weight_image(const uint16_t* w_packed, uint16_t* image);
bool stem_plan_init(StemPlan& p, const void* src_s2d, void* dst, int batch_cap, int pool, char* err, size_t err_len);
cudaError_t stem_plan_launch(const StemPlan& p, int n, int relu, int num_sms, cudaStream_t stream);

// ------------------------------------------------------------------ patch-resident 3x3 conv, 56x56x64 -> 64
struct Conv3x3Plan {
    CUtensorMap tmap_in, tmap_out, tmap_res, tmap_w;
    const void* weight = nullptr;        // packed [64][3][3][64] bf16 (device); nine 64x64 tiles stay in smem
    const float* bias = nullptr;
    int has_res = 0;
};
bool conv3x3_c64_supported(const b2p_op& o);
bool conv3x3_c64_plan_set(Conv3x3Plan& p, const void* src, void* dst, const void* res, int batch_cap, char* err,
                          size_t err_len);
cudaError_t conv3x3_c64_launch(const Conv3x3Plan& p, int n, int relu, int reverse, int num_sms, cudaStream_t s);

// ------------------------------------------------------------------ patch-resident 3x3 conv, 28x28x128 -> 128
struct Conv3x3bPlan {
    CUtensorMap tmap_in, tmap_w;
    const void* weight = nullptr;        // packed [128][3][3][128] bf16 (device); each CTA of a pair keeps 64 rows resident
    void* out = nullptr;                 // NHWC bf16, written with plain stores
    const float* bias = nullptr;
};
bool conv3x3_c128_supported(const b2p_op& o);
bool conv3x3_c128_plan_set(Conv3x3bPlan& p, const void* src, void* dst, int batch_cap, char* err, size_t err_len);
cudaError_t conv3x3_c128_launch(const Conv3x3bPlan& p, int n, int relu, int reverse, int num_sms, cudaStream_t s);

// ------------------------------------------------------------------ fused conv3(+res+ReLU) -> next conv1(+ReLU)
struct PairPlan {
    CUtensorMap tmap_a, tmap_w3, tmap_res, tmap_out1, tmap_w1, tmap_out2, tmap_a2;
    int k1 = 0, k2 = 0, n1 = 0, n2 = 0;
    int stride2 = 1, out_h = 1, out_w = 1, has_res = 1;
    int sub = 1;            // 2: out1 keeps only even rows / columns (compact tensor at out1)
    void* out1 = nullptr;
    const float* bias1 = nullptr;
    const float* bias2 = nullptr;
};
bool pair_plan_supported(int k1, int n1, int n2);
bool pair_plan_init(PairPlan& p, int k1, int k2, int n1, int n2, const void* w3, const float* b3, const void* w1,
                    const float* b1, char* err, size_t err_len);
// Second GEMM1 source (folded projection shortcut): NHWC [batch_cap, out_h*stride2, out_w*stride2, k2] bf16.
bool pair_plan_set_src2(PairPlan& p, const void* src2, int batch_cap, int out_h, int out_w, int stride2, char* err,
                        size_t err_len);
// sub = 2: out1 is a compact [n, out_h/2, out_w/2, n1] tensor holding the even rows / columns only.
bool pair_plan_set_buffers(PairPlan& p, const void* a, const void* res, void* out1, void* out2, int rows, int sub, char* err,
                           size_t err_len);
cudaError_t pair_plan_launch(const PairPlan& p, int m_rows, int reverse, int num_sms, cudaStream_t s);

// ------------------------------------------------------------------ whole bottleneck block, 56x56 stage (bottleneck56.cu)
struct BneckPlan {
    CUtensorMap tmap_x;      // block input [n,56,56,cin] bf16, 4-D tiled, [64 ch x 64 px x 2 rows] boxes
    CUtensorMap tmap_w1;     // [64][cin], 64 x 64 boxes (one K slice travels with each input box)
    CUtensorMap tmap_w2;     // [64][9*64], one 64 x 64 box per tap
    CUtensorMap tmap_w3;     // [256][64], 128-row boxes
    CUtensorMap tmap_out;    // [n*56*56, 256] block output
    const void* x = nullptr; // the block input again (residual, read by the output epilogue)
    const float* b1 = nullptr;
    const float* b2 = nullptr;
    const float* b3 = nullptr;
    int cin = 0, proj = 0;
};
bool bneck56_supported(int in_h, int in_w, int cin, int mid, int cout, bool proj);
bool bneck56_plan_init(BneckPlan& p, int cin, bool proj, const void* w1, const float* b1, const void* w2, const float* b2,
                       const void* w3, const float* b3, char* err, size_t err_len);
bool bneck56_plan_set(BneckPlan& p, const void* src, void* dst, int batch_cap, char* err, size_t err_len);
cudaError_t bneck56_launch(const BneckPlan& p, int n, int num_sms, cudaStream_t s);

// ------------------------------------------------------------------ ViT pieces (vit.cu)
struct AttentionPlan {
    CUtensorMap tmap_q;     // packed QKV [n*seq, 3*dim] bf16, 64 x 128 boxes (query tiles)
    CUtensorMap tmap_kv;    // same tensor, 64 x 208 boxes (all keys / values of one head)
    int seq = 0, dim = 0, heads = 0;
};
cudaError_t launch_patchify(const void* x_nchw, void* out, int n, cudaStream_t s);
cudaError_t launch_tokens(const void* proj, const float* cls, const float* pos, void* out, int n, int seq, int dim,
                          cudaStream_t s);
cudaError_t launch_layernorm(const void* x, const float* gamma, const float* beta, void* out, long rows, int dim,
                             long in_row_stride_elems, float eps, cudaStream_t s);
bool attention_plan_init(AttentionPlan& p, const void* qkv, int batch_cap, int seq, int dim, int heads, char* err,
                         size_t err_len);
cudaError_t attention_plan_launch(const AttentionPlan& p, void* out, int n, cudaStream_t s);

// Records an error that is not tied to a handle; b2p_last_error(NULL) returns it.
void set_global_error(const char* fmt, ...);

// ------------------------------------------------------------------ small memory-bound kernels
cudaError_t launch_stem_pack(const void* x_nchw, void* out_nhwc16, int n, int h, int w,
                             cudaStream_t s);
cudaError_t launch_nchw_to_nhwc(const void* x, void* out, int n, int c, int h, int w,
                                cudaStream_t s);
cudaError_t launch_maxpool(const void* x, void* out, int n, int in_h, int in_w, int c, int out_h,
                           int out_w, int k, int stride, int pad, cudaStream_t s);
cudaError_t launch_avgpool(const void* x, void* out, int n, int hw, int c, cudaStream_t s);
cudaError_t launch_softmax_topk(const float* logits, int n, int classes, int k, float* probs,
                                int32_t* idx, cudaStream_t s);

// ------------------------------------------------------------------ K1 preprocess
struct PreprocessLaunch {
    int max_taps_h = 0;   // largest horizontal tap count over the batch
    int max_row_bytes = 0;
    int all_identity = 0;
};
// `imgs_dev` is the device copy of the descriptors; `imgs_host` is used to size the launch.
// `norm_lut` = device table [3][256] of (v/255 - mean)/std built by launch_norm_lut (may be null: the
// crop-only fast path is then not used).
cudaError_t launch_norm_lut(float* lut, cudaStream_t s);
cudaError_t launch_preprocess(const uint8_t* slab, const b2p_image* imgs_dev,
                              const b2p_image* imgs_host, int n, void* out, int out_f32,
                              const float* norm_lut, cudaStream_t s, int* launches);

// Crop-only batches (every image already has its resized size) written directly as the stem's 2x2 space-to-depth
// input [n,112,112,16] bf16; bit-identical to launch_preprocess (bf16 NCHW) followed by launch_stem_pack.
bool preprocess_all_identity(const b2p_image* imgs_host, int n);
cudaError_t launch_preprocess_s2d(const uint8_t* slab, const b2p_image* imgs_dev, int n, void* out_s2d,
                                  const float* norm_lut, cudaStream_t s);

}  // namespace b2p
