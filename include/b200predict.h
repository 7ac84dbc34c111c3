/*
 * b200predict.h — C ABI of libb200predict.so, the B200 (sm_100a) implementation of the
 * batched image `predict` path of CodyRichter/MLMicroserviceTemplate.
 *
 * Every entry point is `extern "C"`, takes plain pointers / sizes / a cudaStream_t passed as
 * `void*`, returns 0 on success or a negative B2P_E_* code, never throws and never calls back
 * into the host language.  A handle is bound to ONE device and is NOT re-entrant: exactly one
 * host thread drives a given handle; different handles (different GPUs) are independent.
 * All compute entry points are stream-ordered and asynchronous; the caller owns every buffer it
 * passes and keeps it alive until the stream work has completed.
 *
 * There is no CPU fallback anywhere behind this header: b2p_create() fails with
 * B2P_E_DEVICE on anything that is not compute capability 10.x.
 *
 * Reference citations are file:line in CodyRichter/MLMicroserviceTemplate.
 */
#ifndef B200PREDICT_H_
#define B200PREDICT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2P_ABI_VERSION 3

/* error codes */
#define B2P_OK 0
#define B2P_E_ARG -1      /* bad argument (null pointer, size out of range, malformed program) */
#define B2P_E_DEVICE -2   /* not an sm_100 device / device mismatch */
#define B2P_E_CUDA -3     /* a CUDA runtime/driver call failed; see b2p_last_error() */
#define B2P_E_NOMEM -4    /* device or pinned allocation failed */
#define B2P_E_LIMIT -5    /* a size exceeds a documented fixed maximum */
#define B2P_E_UNSUPPORTED -6 /* well-formed input outside this path's scope (e.g. a progressive JPEG): the
                                caller keeps it on the host decoder; nothing was launched              */

typedef struct b2p_handle b2p_handle;

/* ---- layer program -----------------------------------------------------------------------
 * The classifier forward is handed over as a flat program of ops over numbered activation
 * slots.  Slot B2P_SLOT_INPUT is the caller's input tensor, B2P_SLOT_OUTPUT the caller's output;
 * every other slot is handle-owned workspace sized `slot_elems[i] * max_batch` bf16 elements.
 * Activations are NHWC bf16.  The host mirror builds the program by walking the torchvision
 * module tree (ResNet: SP/torchvision/models/resnet.py:266-282).                              */
#define B2P_SLOT_INPUT 0
#define B2P_SLOT_OUTPUT -2
#define B2P_SLOT_NONE -1

enum b2p_op_kind {
    B2P_OP_CONV = 1,       /* implicit-GEMM conv (+bias, +residual, +ReLU), bf16 NHWC -> bf16 NHWC  */
    B2P_OP_LINEAR = 2,     /* [n, in_c] bf16 x [out_c, in_c]^T + bias -> fp32 [n, out_c]            */
    B2P_OP_MAXPOOL = 3,    /* kh x kw / stride / pad_lo max pool, NHWC bf16                          */
    B2P_OP_AVGPOOL = 4,    /* global average pool NHWC bf16 -> [n, c] bf16                           */
    B2P_OP_STEM_PACK = 5,  /* NCHW bf16 [n,3,H,W] -> 2x2 space-to-depth NHWC [n,H/2,W/2,16] bf16    */
    B2P_OP_NCHW_TO_NHWC = 6,/* NCHW bf16 [n,c,H,W] -> NHWC bf16 (test helper, c % 8 == 0)            */
    B2P_OP_STEM_CONV = 7,  /* ResNet stem on the s2d image [n,112,112,16]: 4x4 conv (pad 2/1) + bias +
                              ReLU -> [n,112,112,64]; with reserved[0] = 1 the 3x3/s2/p1 max pool is
                              fused and the output is [n,56,56,64]. Weight blob as for OP_CONV.       */
    /* ViT (SP/torchvision/models/vision_transformer.py); a token matrix is an "image" of in_h = tokens,
       in_w = 1, so linear layers are OP_CONV 1x1 (relu = 2 selects exact-erf GELU).                   */
    B2P_OP_PATCHIFY = 8,   /* NCHW bf16 [n,3,224,224] -> 16x16 patches [n*196, 768] bf16 (c,ky,kx order)*/
    B2P_OP_TOKENS = 9,     /* [cls | src[n,in_h-1,c]] + pos -> [n,in_h,c]; weight blob = pos f32 [in_h*c],
                              bias blob = class token f32 [c]                                            */
    B2P_OP_LAYERNORM = 10, /* per-token LayerNorm over c, eps = 1e-6; weight/bias blobs = gamma/beta f32;
                              reserved[0] = 1: only token 0 of each image -> [n, c]                      */
    B2P_OP_ATTENTION = 11, /* softmax(Q K^T / sqrt(64)) V per head on packed QKV [n,in_h,3c] -> [n,in_h,c];
                              reserved[0] = number of heads (head dim 64)                                */
    B2P_OP_CONV_PAIR = 12, /* two chained 1x1/s1 convs in one kernel (a bottleneck's conv3 and the next block's
                              conv1): dst = relu(conv(src) + bias + res) as OP_CONV, and additionally
                              slot reserved[0] = relu(conv(dst) + bias2) with weight blob reserved[1] and bias blob
                              reserved[2], reserved[3] = its output channels (64, 128 or 256).  reserved2 = 2: dst keeps only
                              the pixels with even row and column, as a compact [n, out_h/2, out_w/2, out_c] tensor - for a
                              dst whose only other reader is a 1x1/stride-2 conv (ResNet's projection shortcut), which then
                              reads it with stride 1                                                                    */
    B2P_OP_BOTTLENECK = 13 /* a whole identity bottleneck block of the 56x56 stage in one kernel
                              (SP/torchvision/models/resnet.py:143-163): t1 = relu(conv1x1(src)), t2 = relu(conv3x3(t1)),
                              dst = relu(conv1x1(t2) + src).  in_c = 256, reserved2 = bottleneck width (64), out_c = 256,
                              res == src.  Blobs: weight / bias = conv1 [64][256], reserved[0] / [1] = conv2 [64][3][3][64]
                              and its bias, reserved[2] / [3] = conv3 [256][64] and its bias.  src and dst are
                              workspace slots.  (res == B2P_SLOT_NONE would describe a projection block with the
                              shortcut folded into conv3; b2p_create refuses it: no kernel covers it.)              */
};

typedef struct b2p_op {
    int32_t kind;
    int32_t src, dst, res;          /* slot ids; res = residual added before ReLU, or B2P_SLOT_NONE */
    int32_t in_h, in_w, in_c;
    int32_t out_h, out_w, out_c;
    int32_t kh, kw, stride;
    int32_t pad_lo, pad_hi;         /* zero padding before / after, same on both axes              */
    int32_t relu;                   /* 0 none, 1 ReLU, 2 GELU (exact erf)                          */
    int32_t weight, bias;           /* blob indices: weight bf16 [out_c_pad][kh][kw][in_c], bias f32 */
    int32_t reserved[4];
    /* Optional second GEMM source for OP_CONV / OP_CONV_PAIR (src2 = B2P_SLOT_NONE when unused): the op also
       adds a 1x1 / stride src2_stride conv of slot src2 (src2_c channels, out_h*stride x out_w*stride pixels)
       into the same accumulator - a bottleneck's projection shortcut folded into its conv3, K = in_c + src2_c.
       The weight blob then holds [out_c_pad][kh*kw*in_c + src2_c] and the bias blob the sum of both biases.  */
    int32_t src2, src2_c, src2_stride;
    int32_t reserved2;
} b2p_op;

typedef struct b2p_model_desc {
    int32_t abi_version;            /* B2P_ABI_VERSION */
    int32_t max_batch;              /* largest n any call will pass                                */
    int32_t sub_batch;              /* >0: run the program over chunks of this many images so that
                                       inter-layer activations stay L2-resident; 0 = whole batch   */
    int32_t n_ops;
    const b2p_op* ops;
    int32_t n_slots;                /* slots 0..n_slots-1; slot 0 is the external input            */
    const int64_t* slot_elems;      /* per-image element count of each slot (bf16 elements)        */
    int32_t n_blobs;
    const void* const* blob_ptrs;   /* HOST pointers; copied to the device by b2p_create            */
    const int64_t* blob_bytes;
    int64_t in_elems_per_image;     /* elements per image of the external input  (bf16)            */
    int64_t out_elems_per_image;    /* elements per image of the external output (fp32 or bf16)    */
} b2p_model_desc;

/* One decoded image inside a batch slab (uint8, HWC, 3 channels, rows densely packed).
 * The resize geometry follows SP/torchvision/transforms/_presets.py:58-65: the image is resized
 * (antialiased bilinear) to res_h x res_w, then the 224x224 window at (top, left) is kept.      */
typedef struct b2p_image {
    int64_t offset;                 /* byte offset of pixel (0,0,0) from the slab base             */
    int32_t h, w;                   /* source size                                                 */
    int32_t res_h, res_w;           /* size after resize                                           */
    int32_t top, left;              /* crop origin in the resized image                            */
} b2p_image;

#define B2P_OUT_HW 224              /* model input is 224 x 224                                     */
#define B2P_MAX_TOPK 8

typedef struct b2p_stats {
    float ms_h2d, ms_preprocess, ms_forward, ms_epilogue, ms_d2h;
                                    /* CUDA-event times.  ms_preprocess / ms_forward / ms_epilogue (K1 / K2 / K3): the
                                       MEAN over the SAMPLED b2p_predict and b2p_predict_host calls issued since the
                                       previous b2p_stats_get whose work has completed (at most the 64 most recent).
                                       Every 4th call is sampled, the first one included (environment B2P_STATS_EVERY=n
                                       at create; 1 = every call): the events that delimit the stages sit between
                                       kernels and cost those kernels their launch overlap;
                                       ms_h2d / ms_d2h: the most recent b2p_predict_host call                        */
    int64_t calls;
    int64_t kernel_launches;        /* kernels launched by this handle since create                */
} b2p_stats;

/* Library-wide: ABI version and a static build string ("sm_100a ..."). */
int32_t b2p_abi_version(void);
const char* b2p_build_info(void);

/* Replaces init() (src/model/model.py:6-13): builds one model replica on `device`: copies the
 * packed weights, allocates workspace, encodes the TMA descriptors.  Synchronous.               */
int32_t b2p_create(int32_t device, const b2p_model_desc* desc, b2p_handle** out);
int32_t b2p_destroy(b2p_handle* h);
const char* b2p_last_error(const b2p_handle* h);   /* h may be NULL: last create() failure      */
int32_t b2p_stats_get(b2p_handle* h, b2p_stats* out);

/* Stage P (K1), the arithmetic predict() would do after Image.open (src/model/model.py:23):
 * antialiased resize + crop + /255 + (x-mean)/std + HWC->NCHW.  `slab` is a DEVICE pointer to the
 * uint8 images, `imgs` a HOST array of n descriptors (copied before return).  `out` is a device
 * tensor [n,3,224,224], bf16 when out_f32 == 0, fp32 otherwise (the fp32 form exists for the
 * 1e-6 parity gate).  Precondition (not checkable here, the slab size is unknown): every
 * [offset, offset + h*w*3) lies inside the caller's slab.  b2p_predict_host and b2p_jpeg_decode,
 * which are told the slab size, check it and fail with B2P_E_ARG.                              */
int32_t b2p_preprocess(b2p_handle* h, const uint8_t* slab, const b2p_image* imgs, int32_t n,
                       void* out, int32_t out_f32, void* stream);

/* Stage F (K2): runs the layer program on n images. `x` = external input, `out` = external
 * output (fp32 logits [n, classes] for a classifier program).                                  */
int32_t b2p_forward(b2p_handle* h, const void* x, int32_t n, void* out, void* stream);

/* Stage E (K3): fp32 softmax over `classes` + top-k (k <= B2P_MAX_TOPK), ties -> lowest index.
 * logits [n, classes] fp32 device -> probs [n, k] fp32, idx [n, k] int32 (device).             */
int32_t b2p_softmax_topk(b2p_handle* h, const float* logits, int32_t n, int32_t classes,
                         int32_t k, float* probs, int32_t* idx, void* stream);

/* The three stages chained on one stream, device buffers in, device buffers out.  This is what
 * predict() (src/model/model.py:16-31) computes for a batch of n already-decoded images.       */
int32_t b2p_predict(b2p_handle* h, const uint8_t* slab, const b2p_image* imgs, int32_t n,
                    int32_t k, float* probs, int32_t* idx, void* stream);

/* Same, end to end from HOST memory: copies `slab_bytes` of (pinned) host slab to the device,
 * runs the chain, copies probs/idx back to (pinned) host memory.  Asynchronous: synchronise the
 * stream before reading probs_host / idx_host.  Records per-stage CUDA-event times (b2p_stats). */
int32_t b2p_predict_host(b2p_handle* h, const uint8_t* slab_host, int64_t slab_bytes,
                         const b2p_image* imgs, int32_t n, int32_t k, float* probs_host,
                         int32_t* idx_host, void* stream);

/* Device logits of the most recent b2p_predict / b2p_predict_host call ([n, classes] fp32,
 * handle-owned, valid until the next call).  Used by the parity tests.                         */
const float* b2p_last_logits(b2p_handle* h);

/* No counterpart in the reference (it has no device code at all).  The classifier head — global
 * average pool of feat [n, hw, c] bf16 (NHWC) followed by logits = pooled x weight^T + bias — for
 * callers that must not allocate device memory: a process whose own pool already holds all of the
 * GPU's memory (a library-side cudaMalloc would fail), or one that keeps the allocator off its
 * request path because cudaMalloc/cudaFree stall the device.  It takes no handle and calls no
 * cudaMalloc/cudaFree or any allocating routine; the pooled features live in a fixed 1 MiB array
 * inside the library image, which every GPU that loads the library holds until exit.
 *   weight : device, bf16 [weight_rows][c], weight_rows a multiple of 128 and >= classes (zero rows
 *            pad the tail), c a multiple of 64;  bias : device, fp32 [weight_rows];
 *   logits : device, fp32 [n][classes].
 * Fails with B2P_E_LIMIT when n > B2P_HEAD_STATIC_MAX_IMAGES or c > B2P_HEAD_STATIC_MAX_CHANNELS.
 * All calls on one device share the array and it keeps the previous call's contents (nothing in it
 * needs to start at zero: rows [0, n) are rewritten before they are read and no other row is read);
 * calls that overlap on the GPU race on it.  The first call on each device does a one-time set-up
 * (symbol address, driver entry points, kernel attributes) and should come before any stream
 * capture.  Errors are reported through b2p_last_error(NULL).                                    */
#define B2P_HEAD_STATIC_MAX_IMAGES 256
#define B2P_HEAD_STATIC_MAX_CHANNELS 2048
int32_t b2p_head_static(const void* feat, int32_t n, int32_t hw, int32_t c, const void* weight,
                        int32_t weight_rows, const float* bias, int32_t classes, float* logits,
                        void* stream);

/* ---- JPEG decode on the GPU (the step right before the path) -------------
 * For the files it covers this produces the pixels `Image.open(image_file.file)` + `.convert('RGB')`
 * (src/model/model.py:23; Pillow / libjpeg-turbo defaults: accurate integer IDCT, fancy chroma upsampling,
 * fixed-point YCbCr->RGB) would: uint8 RGB, HWC, rows densely packed — bit for bit.  Scope: baseline /
 * extended-sequential Huffman JPEG (SOF0, SOF1), 8-bit, ONE interleaved scan, greyscale or YCbCr with luma
 * sampling 1x1 (4:4:4), 2x1 (4:2:2) or 2x2 (4:2:0), restart intervals allowed, no Adobe marker, at most 64 Mpixel.  Everything
 * else reports B2P_E_UNSUPPORTED and stays on the caller's host decoder.  The decoder object is independent
 * of b2p_handle; like a handle it is bound to one device and driven by one host thread, and consecutive
 * calls must be issued on the same stream (its workspace is reused in stream order).                     */
typedef struct b2p_jpeg b2p_jpeg;

typedef struct b2p_jpeg_info {
    int32_t width, height;
    int32_t components;             /* 1 = greyscale, 3 = YCbCr                                     */
    int32_t h_samp, v_samp;         /* luma sampling factors: 1x1, 2x1 or 2x2                      */
    int32_t restart_interval;       /* MCUs per restart interval, 0 = none                         */
    int32_t segments;               /* independently decodable entropy-coded segments              */
    int32_t reserved;
    int64_t scan_bytes;             /* entropy-coded bytes                                         */
} b2p_jpeg_info;

/* Host only, no device needed: walks the markers of one file.  B2P_OK (info filled), B2P_E_UNSUPPORTED
 * (fine for the host decoder, not for this one) or B2P_E_ARG (not a decodable JPEG); the reason is in
 * b2p_jpeg_last_error(NULL).                                                                            */
int32_t b2p_jpeg_inspect(const uint8_t* file, int64_t len, b2p_jpeg_info* info);

typedef struct b2p_jpeg_stats {
    float ms_copy, ms_huffman, ms_idct, ms_colour;   /* CUDA-event times of the most recent b2p_jpeg_decode  */
    int64_t calls;
} b2p_jpeg_stats;

int32_t b2p_jpeg_create(int32_t device, b2p_jpeg** out);
int32_t b2p_jpeg_destroy(b2p_jpeg* d);
const char* b2p_jpeg_last_error(const b2p_jpeg* d);   /* d may be NULL: the calling THREAD's last inspect / create failure */
int32_t b2p_jpeg_stats_get(b2p_jpeg* d, b2p_jpeg_stats* out);   /* waits for the most recent call       */

/* Decodes n files (HOST pointers, any memory) into the DEVICE slab: image i becomes uint8 [h][w][3] at
 * slab + offsets[i], ready to be described by a b2p_image and handed to b2p_preprocess / b2p_predict on
 * the same stream.  `status` (device, n x int32, may be NULL) receives per image 0 or a bit mask:
 * 1 = the entropy-coded data ended early, 2 = an invalid Huffman code was met (pixels are then undefined,
 * as with any corrupt file).  Headers are parsed on the calling thread; the entropy-coded bytes and the
 * tables travel in one host->device copy; three kernels follow.  Asynchronous on `stream`.  The workspace
 * grows (with a stream synchronisation) the first time a larger batch is seen.  Returns B2P_E_UNSUPPORTED /
 * B2P_E_ARG without launching anything if any file is outside the scope — inspect first.                */
int32_t b2p_jpeg_decode(b2p_jpeg* d, const uint8_t* const* files, const int64_t* lens, int32_t n,
                        uint8_t* slab, int64_t slab_bytes, const int64_t* offsets, int32_t* status,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200PREDICT_H_ */
