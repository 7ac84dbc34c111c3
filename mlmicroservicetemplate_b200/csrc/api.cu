// C-ABI layer of libb200predict.so: handle lifetime, the layer-program executor and the chained
// predict entry points declared in include/b200predict.h.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include "kernels.h"

using namespace b2p;

namespace {

// Error text of calls that have no handle (b2p_create, b2p_head_static): one buffer per calling THREAD, because
// init() builds the replicas of several GPUs from several threads at once.
thread_local char g_create_error[512] = "";

struct OpPlan {
    b2p_op op;
    GemmPlan gemm;        // valid for CONV / LINEAR
    StemPlan stem;        // valid for STEM_CONV
    AttentionPlan att;    // valid for ATTENTION
    PairPlan pair;        // valid for CONV_PAIR
    Conv3x3Plan c3;       // valid for CONV ops taken by the patch-resident 3x3 kernel
    BneckPlan bneck;      // valid for BOTTLENECK
    bool use_c3 = false;
    Conv3x3bPlan c3b;     // same idea for 28x28x128 -> 128
    bool use_c3b = false;
    void* stem_image = nullptr;
    bool is_gemm = false;
    bool src_external = false;
    bool dst_external = false;
    bool res_external = false;
};

}  // namespace

struct b2p_handle {
    int device = -1;
    int num_sms = 0;
    int max_batch = 0;
    int chunk = 0;                       // images per pass through the program
    bool alternate = true;               // alternate the tile direction layer by layer (B2P_NO_ALTERNATE, read at create)
    int fused_input_slot = -1;           // >= 0: op 0 is the stem's space-to-depth pack of the external input and nothing
                                         // else reads that input, so the predict chain may let K1 write this slot directly
    int64_t in_elems = 0, out_elems = 0;
    int classes = 0;
    std::vector<OpPlan> ops;
    std::vector<void*> slots;            // device workspace per slot (slot 0 unused)
    std::vector<int64_t> slot_elems;
    std::vector<void*> blobs;            // device copies of the weight blobs
    std::vector<int64_t> blob_bytes;
    // predict-chain staging (handle-owned)
    // predict_host input staging: two device slabs filled on a private copy stream, so the H2D copy
    // of call i+1 overlaps the kernels of call i (which run on the caller's stream)
    uint8_t* slab_dev[2] = {nullptr, nullptr};
    int64_t slab_cap[2] = {0, 0};
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {};      // H2D into slab i finished (copy stream)
    cudaEvent_t ev_consumed[2] = {};    // K1 finished reading slab i (caller's stream)
    bool slab_used[2] = {false, false};
    int slab_idx = 0;
    // descriptor staging: a ring of pinned + device slots so that back-to-back asynchronous calls
    // never overwrite a descriptor block whose host->device copy is still queued
    static constexpr int RING = 8;
    b2p_image* imgs_dev[RING] = {};
    b2p_image* imgs_pinned[RING] = {};
    cudaEvent_t imgs_done[RING] = {};
    bool imgs_used[RING] = {};
    int ring_pos = 0;
    b2p_image* cur_imgs_dev = nullptr;
    void* x_dev = nullptr;               // [max_batch,3,224,224] bf16
    float* logits_dev = nullptr;         // [max_batch, out_elems]
    float* norm_lut = nullptr;           // [3][256] normalisation table for K1's crop-only path
    float* probs_dev = nullptr;
    int32_t* idx_dev = nullptr;
    cudaEvent_t ev[7] = {};   // [0],[1] copy begin/end (copy stream), [5] D2H begin, [6] D2H end (caller's stream); [2..4] unused
    bool ev_valid = false;
    // per-stage events of the predict chains: K1 begin, K1 / K2 / K3 end, one set per call in a ring, so that
    // b2p_stats_get can average the stages over every call since the previous read
    static constexpr int SRING = 64;
    cudaEvent_t sev[SRING][4] = {};
    int sev_pos = 0;          // next set to record into
    int sev_unread = 0;       // sets recorded since the last b2p_stats_get (saturates at SRING)
    int sev_last = -1;
    // Stage events sit BETWEEN kernels of the chain and keep the next kernel's programmatic launch from overlapping
    // its predecessor's tail, so only every stage_every-th call records them (B2P_STATS_EVERY, default 4; 1 = every
    // call); b2p_stats_get averages over the sampled calls.
    int stage_every = 4;
    long predict_calls = 0;
    b2p_stats stats = {};
    char err[512] = "";
};

namespace {

void set_error(b2p_handle* h, const char* fmt, ...) {
    char* dst = h ? h->err : g_create_error;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
}

#define B2P_CUDA(h, call)                                                                  \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess) {                                                          \
            set_error(h, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                      __LINE__);                                                           \
            return B2P_E_CUDA;                                                             \
        }                                                                                  \
    } while (0)

void* slot_ptr(b2p_handle* h, int slot, const void* ext_in, void* ext_out, int64_t chunk_off_imgs) {
    if (slot == B2P_SLOT_INPUT)
        return const_cast<uint8_t*>(static_cast<const uint8_t*>(ext_in)) + chunk_off_imgs * h->in_elems * 2;
    if (slot == B2P_SLOT_OUTPUT) return ext_out;   // offset applied by the caller (element size varies)
    if (slot < 0 || slot >= static_cast<int>(h->slots.size())) return nullptr;
    return h->slots[slot];
}

int validate_desc(const b2p_model_desc* d) {
    if (!d || d->abi_version != B2P_ABI_VERSION) return B2P_E_ARG;
    if (d->max_batch <= 0 || d->max_batch > 4096) return B2P_E_LIMIT;
    if (d->n_ops <= 0 || !d->ops || d->n_slots <= 0 || !d->slot_elems) return B2P_E_ARG;
    if (d->n_blobs < 0 || (d->n_blobs > 0 && (!d->blob_ptrs || !d->blob_bytes))) return B2P_E_ARG;
    if (d->sub_batch < 0) return B2P_E_ARG;
    for (int i = 0; i < d->n_ops; ++i) {
        const b2p_op& o = d->ops[i];
        if (o.kind < B2P_OP_CONV || o.kind > B2P_OP_BOTTLENECK) return B2P_E_ARG;
        auto ok_slot = [&](int s, bool allow_none) {
            if (s == B2P_SLOT_NONE) return allow_none;
            if (s == B2P_SLOT_OUTPUT) return true;
            return s >= 0 && s < d->n_slots;
        };
        if (!ok_slot(o.src, false) || o.src == B2P_SLOT_OUTPUT) return B2P_E_ARG;
        if (!ok_slot(o.dst, false) || o.dst == B2P_SLOT_INPUT) return B2P_E_ARG;
        if (!ok_slot(o.res, true) || o.res == B2P_SLOT_OUTPUT) return B2P_E_ARG;
        if (o.src2_c > 0 && (o.src2 <= 0 || o.src2 >= d->n_slots || (o.kind != B2P_OP_CONV && o.kind != B2P_OP_CONV_PAIR)))
            return B2P_E_ARG;
        if (o.kind == B2P_OP_CONV_PAIR) {
            if (o.reserved[0] <= 0 || o.reserved[0] >= d->n_slots || o.reserved[1] < 0 || o.reserved[1] >= d->n_blobs ||
                o.reserved[2] < 0 || o.reserved[2] >= d->n_blobs || o.src == B2P_SLOT_INPUT ||
                o.dst == B2P_SLOT_OUTPUT || o.res == B2P_SLOT_INPUT || (o.res == B2P_SLOT_NONE && o.src2_c <= 0))
                return B2P_E_ARG;
        }
        if (o.kind == B2P_OP_BOTTLENECK) {
            for (int r = 0; r < 4; ++r)
                if (o.reserved[r] < 0 || o.reserved[r] >= d->n_blobs) return B2P_E_ARG;
            if (o.src <= 0 || o.dst <= 0 || o.src == o.dst || (o.res != B2P_SLOT_NONE && o.res != o.src)) return B2P_E_ARG;
        }
        if (o.kind == B2P_OP_CONV || o.kind == B2P_OP_LINEAR || o.kind == B2P_OP_STEM_CONV ||
            o.kind == B2P_OP_TOKENS || o.kind == B2P_OP_LAYERNORM || o.kind == B2P_OP_CONV_PAIR ||
            o.kind == B2P_OP_BOTTLENECK) {
            if (o.weight < 0 || o.weight >= d->n_blobs || o.bias < 0 || o.bias >= d->n_blobs) return B2P_E_ARG;
        }
        // every op: the slots it touches must hold the tensors its shape fields describe (a malformed program
        // would otherwise read or write out of bounds on the device)
        if (o.in_h <= 0 || o.in_w <= 0 || o.in_c <= 0 || o.out_h <= 0 || o.out_w <= 0 || o.out_c <= 0) return B2P_E_ARG;
        auto holds = [&](int s, int64_t need) {
            if (s == B2P_SLOT_INPUT) return need <= d->in_elems_per_image;
            if (s == B2P_SLOT_OUTPUT) return need <= d->out_elems_per_image;
            return need <= d->slot_elems[s];
        };
        int64_t in_need = static_cast<int64_t>(o.in_h) * o.in_w * o.in_c;
        int64_t out_need = static_cast<int64_t>(o.out_h) * o.out_w * o.out_c;
        if (o.kind == B2P_OP_TOKENS) in_need = static_cast<int64_t>(o.in_h - 1) * o.in_w * o.in_c;   // the class token is added
        if (o.kind == B2P_OP_ATTENTION) in_need *= 3;                                               // packed Q | K | V
        int64_t dst_need = out_need;
        if (o.kind == B2P_OP_CONV_PAIR && o.reserved2 == 2) {      // dst keeps the even rows / columns only
            if ((o.out_h | o.out_w) & 1) return B2P_E_ARG;
            dst_need = static_cast<int64_t>(o.out_h / 2) * (o.out_w / 2) * o.out_c;
        }
        if (!holds(o.src, in_need) || !holds(o.dst, dst_need)) return B2P_E_ARG;
        if (o.res != B2P_SLOT_NONE && !holds(o.res, out_need)) return B2P_E_ARG;
        if (o.src2_c > 0) {
            const int64_t st = o.src2_stride > 0 ? o.src2_stride : 1;
            if (!holds(o.src2, static_cast<int64_t>(o.out_h) * st * o.out_w * st * o.src2_c)) return B2P_E_ARG;
        }
    }
    return B2P_OK;
}

int run_program(b2p_handle* h, const void* x, int n, void* out, cudaStream_t st, size_t first_op = 0) {
    // Tile direction: a GEMM layer walks its row tiles in the direction OPPOSITE to the last op that
    // touched its input, so it starts on the rows that are still resident in L2 (126 MB) instead of on
    // the ones that were evicted first.  Slot -> direction of the last toucher (0 = ascending).
    const bool alternate = h->alternate;
    std::vector<int> touch(h->slots.size() + 1, 0);
    auto touch_of = [&](int slot) -> int& { return touch[slot >= 0 && slot < (int)h->slots.size() ? slot : h->slots.size()]; };
    for (int off = 0; off < n; off += h->chunk) {
        const int nb = (n - off) < h->chunk ? (n - off) : h->chunk;
        for (size_t op_i = first_op; op_i < h->ops.size(); ++op_i) {
            OpPlan& p = h->ops[op_i];
            int op_dir = 0;
            {
                const bool directed = p.is_gemm || p.op.kind == B2P_OP_CONV_PAIR;
                const int dir = (directed && alternate) ? !touch_of(p.op.src) : 0;
                if (p.is_gemm) p.gemm.reverse = dir;
                op_dir = dir;
                touch_of(p.op.src) = dir;
                if (p.op.res != B2P_SLOT_NONE) touch_of(p.op.res) = dir;
                if (p.op.src2_c > 0) touch_of(p.op.src2) = dir;
                touch_of(p.op.dst) = dir;
                if (p.op.kind == B2P_OP_CONV_PAIR) touch_of(p.op.reserved[0]) = dir;
            }
            const b2p_op& o = p.op;
            const void* src = slot_ptr(h, o.src, x, out, off);
            void* dst = slot_ptr(h, o.dst, x, out, off);
            if (dst && o.dst == B2P_SLOT_OUTPUT)    // chunk `off` of the caller's output (fp32 logits for LINEAR, else bf16)
                dst = static_cast<uint8_t*>(dst) + static_cast<size_t>(off) * h->out_elems * (o.kind == B2P_OP_LINEAR ? 4 : 2);
            const void* res = o.res == B2P_SLOT_NONE ? nullptr : slot_ptr(h, o.res, x, out, off);
            if (!src || !dst) { set_error(h, "program references an unallocated slot"); return B2P_E_ARG; }
            cudaError_t e = cudaSuccess;
            switch (o.kind) {
                case B2P_OP_CONV:
                case B2P_OP_LINEAR: {
                    if (p.use_c3) {
                        if (p.src_external || p.dst_external || p.res_external) {
                            char eb[256] = "";
                            if (!conv3x3_c64_plan_set(p.c3, src, dst, res, nb, eb, sizeof eb)) {
                                set_error(h, "%s", eb);
                                return B2P_E_CUDA;
                            }
                        }
                        e = conv3x3_c64_launch(p.c3, nb, o.relu, op_dir, h->num_sms, st);
                        break;
                    }
                    if (p.use_c3b) {
                        if (p.src_external || p.dst_external) {
                            char eb[256] = "";
                            if (!conv3x3_c128_plan_set(p.c3b, src, dst, nb, eb, sizeof eb)) {
                                set_error(h, "%s", eb);
                                return B2P_E_CUDA;
                            }
                        }
                        e = conv3x3_c128_launch(p.c3b, nb, o.relu, op_dir, h->num_sms, st);
                        break;
                    }
                    if (p.src_external) {
                        // A-operand map points into the caller's buffer: encode it for this chunk
                        char eb[256] = "";
                        if (!gemm_plan_set_src(p.gemm, src, nb, eb, sizeof eb)) {
                            set_error(h, "%s", eb);
                            return B2P_E_CUDA;
                        }
                    }
                    const int rows = o.kind == B2P_OP_LINEAR ? nb : nb * o.out_h * o.out_w;
                    char eb2[256] = "";
                    if ((p.dst_external && !gemm_plan_set_dst(p.gemm, dst, rows, eb2, sizeof eb2)) ||
                        (res && p.res_external && !gemm_plan_set_res(p.gemm, res, rows, eb2, sizeof eb2))) {
                        set_error(h, "%s", eb2);
                        return B2P_E_CUDA;
                    }
                    e = gemm_plan_launch(p.gemm, rows, res != nullptr, dst, h->num_sms, st);
                    break;
                }
                case B2P_OP_STEM_CONV: {
                    if (o.src == B2P_SLOT_INPUT || o.dst == B2P_SLOT_OUTPUT) {
                        char eb[256] = "";
                        if (!stem_plan_init(p.stem, src, dst, nb, o.reserved[0], eb, sizeof eb)) {
                            set_error(h, "%s", eb);
                            return B2P_E_CUDA;
                        }
                    }
                    e = stem_plan_launch(p.stem, nb, o.relu, h->num_sms, st);
                    break;
                }
                case B2P_OP_BOTTLENECK:
                    e = bneck56_launch(p.bneck, nb, h->num_sms, st);
                    break;
                case B2P_OP_CONV_PAIR:
                    e = pair_plan_launch(p.pair, nb * o.out_h * o.out_w, op_dir, h->num_sms, st);
                    break;
                case B2P_OP_PATCHIFY:
                    e = launch_patchify(src, dst, nb, st);
                    break;
                case B2P_OP_TOKENS:
                    e = launch_tokens(src, static_cast<const float*>(h->blobs[o.bias]),
                                      static_cast<const float*>(h->blobs[o.weight]), dst, nb, o.in_h, o.in_c, st);
                    break;
                case B2P_OP_LAYERNORM: {
                    const bool cls_only = o.reserved[0] != 0;
                    e = launch_layernorm(src, static_cast<const float*>(h->blobs[o.weight]),
                                         static_cast<const float*>(h->blobs[o.bias]), dst,
                                         cls_only ? nb : static_cast<long>(nb) * o.in_h, o.in_c,
                                         cls_only ? static_cast<long>(o.in_h) * o.in_c : o.in_c, 1e-6f, st);
                    break;
                }
                case B2P_OP_ATTENTION: {
                    if (o.src == B2P_SLOT_INPUT) {
                        char eb[256] = "";
                        if (!attention_plan_init(p.att, src, nb, o.in_h, o.in_c, o.reserved[0], eb, sizeof eb)) {
                            set_error(h, "%s", eb);
                            return B2P_E_CUDA;
                        }
                    }
                    e = attention_plan_launch(p.att, dst, nb, st);
                    break;
                }
                case B2P_OP_MAXPOOL:
                    e = launch_maxpool(src, dst, nb, o.in_h, o.in_w, o.in_c, o.out_h, o.out_w, o.kh, o.stride,
                                       o.pad_lo, st);
                    break;
                case B2P_OP_AVGPOOL:
                    e = launch_avgpool(src, dst, nb, o.in_h * o.in_w, o.in_c, st);
                    break;
                case B2P_OP_STEM_PACK:
                    e = launch_stem_pack(src, dst, nb, o.in_h, o.in_w, st);
                    break;
                case B2P_OP_NCHW_TO_NHWC:
                    e = launch_nchw_to_nhwc(src, dst, nb, o.in_c, o.in_h, o.in_w, st);
                    break;
                default:
                    set_error(h, "unknown op kind %d", o.kind);
                    return B2P_E_ARG;
            }
            if (e != cudaSuccess) {
                set_error(h, "op kind %d launch failed: %s", o.kind, cudaGetErrorString(e));
                return B2P_E_CUDA;
            }
            h->stats.kernel_launches += 1;
        }
    }
    return B2P_OK;
}

int stage_images(b2p_handle* h, const b2p_image* imgs, int n, cudaStream_t st) {
    const int r = h->ring_pos;
    h->ring_pos = (r + 1) % b2p_handle::RING;
    if (h->imgs_used[r]) B2P_CUDA(h, cudaEventSynchronize(h->imgs_done[r]));
    memcpy(h->imgs_pinned[r], imgs, sizeof(b2p_image) * static_cast<size_t>(n));
    B2P_CUDA(h, cudaMemcpyAsync(h->imgs_dev[r], h->imgs_pinned[r], sizeof(b2p_image) * static_cast<size_t>(n),
                                cudaMemcpyHostToDevice, st));
    B2P_CUDA(h, cudaEventRecord(h->imgs_done[r], st));
    h->imgs_used[r] = true;
    h->cur_imgs_dev = h->imgs_dev[r];
    return B2P_OK;
}

int check_images(b2p_handle* h, const b2p_image* imgs, int n) {
    for (int i = 0; i < n; ++i) {
        const b2p_image& im = imgs[i];
        if (im.h <= 0 || im.w <= 0 || im.res_h < B2P_OUT_HW || im.res_w < B2P_OUT_HW || im.top < 0 ||
            im.left < 0 || im.top + B2P_OUT_HW > im.res_h || im.left + B2P_OUT_HW > im.res_w || im.offset < 0) {
            set_error(h, "image %d: bad geometry h=%d w=%d res=%dx%d crop=(%d,%d)", i, im.h, im.w, im.res_h,
                      im.res_w, im.top, im.left);
            return B2P_E_ARG;
        }
    }
    return B2P_OK;
}

}  // namespace

namespace b2p {
void set_global_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_create_error, sizeof g_create_error, fmt, ap);
    va_end(ap);
}
}  // namespace b2p

extern "C" {

int32_t b2p_abi_version(void) { return B2P_ABI_VERSION; }

const char* b2p_build_info(void) {
    return "libb200predict sm_100a (tcgen05/TMA/TMEM), built " __DATE__ " " __TIME__;
}

const char* b2p_last_error(const b2p_handle* h) { return h ? h->err : g_create_error; }

int32_t b2p_destroy(b2p_handle* h) {
    if (!h) return B2P_OK;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (OpPlan& op : h->ops) {
        if (op.stem_image) cudaFree(op.stem_image);
    }
    for (void* p : h->slots) if (p) cudaFree(p);
    for (void* p : h->blobs) if (p) cudaFree(p);
    for (int i = 0; i < 2; ++i) {
        if (h->slab_dev[i]) cudaFree(h->slab_dev[i]);
        if (h->ev_copied[i]) cudaEventDestroy(h->ev_copied[i]);
        if (h->ev_consumed[i]) cudaEventDestroy(h->ev_consumed[i]);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (int r = 0; r < b2p_handle::RING; ++r) {
        if (h->imgs_dev[r]) cudaFree(h->imgs_dev[r]);
        if (h->imgs_pinned[r]) cudaFreeHost(h->imgs_pinned[r]);
        if (h->imgs_done[r]) cudaEventDestroy(h->imgs_done[r]);
    }
    if (h->x_dev) cudaFree(h->x_dev);
    if (h->norm_lut) cudaFree(h->norm_lut);
    if (h->logits_dev) cudaFree(h->logits_dev);
    if (h->probs_dev) cudaFree(h->probs_dev);
    if (h->idx_dev) cudaFree(h->idx_dev);
    if (h->ev_valid) for (auto& e : h->ev) cudaEventDestroy(e);
    for (auto& set : h->sev) for (auto& e : set) if (e) cudaEventDestroy(e);
    delete h;
    return B2P_OK;
}

int32_t b2p_create(int32_t device, const b2p_model_desc* desc, b2p_handle** out) {
    g_create_error[0] = 0;
    if (!out) return B2P_E_ARG;
    *out = nullptr;
    int rc = validate_desc(desc);
    if (rc != B2P_OK) { set_error(nullptr, "invalid model description (code %d)", rc); return rc; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        set_error(nullptr, "no such CUDA device %d (%d visible)", device, ndev);
        return B2P_E_DEVICE;
    }
    cudaDeviceProp prop;
    B2P_CUDA(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error(nullptr, "device %d is sm_%d%d; this library only runs on sm_100 (B200) and has no fallback",
                  device, prop.major, prop.minor);
        return B2P_E_DEVICE;
    }
    B2P_CUDA(nullptr, cudaSetDevice(device));
    char eb[256] = "";
    if (!tma_init(eb, sizeof eb)) { set_error(nullptr, "%s", eb); return B2P_E_CUDA; }

    b2p_handle* h = new (std::nothrow) b2p_handle();
    if (!h) return B2P_E_NOMEM;
    h->device = device;
    h->num_sms = prop.multiProcessorCount;
    h->max_batch = desc->max_batch;
    h->chunk = (desc->sub_batch > 0 && desc->sub_batch < desc->max_batch) ? desc->sub_batch : desc->max_batch;
    h->in_elems = desc->in_elems_per_image;
    h->out_elems = desc->out_elems_per_image;
    h->alternate = getenv("B2P_NO_ALTERNATE") == nullptr;
    if (const char* e = getenv("B2P_STATS_EVERY")) h->stage_every = atoi(e) > 0 ? atoi(e) : 1;

#define B2P_FAIL(code, ...)                         \
    do {                                            \
        set_error(nullptr, __VA_ARGS__);            \
        b2p_destroy(h);                             \
        return code;                                \
    } while (0)

    // weights
    h->blobs.assign(desc->n_blobs, nullptr);
    h->blob_bytes.assign(desc->blob_bytes, desc->blob_bytes + desc->n_blobs);
    for (int i = 0; i < desc->n_blobs; ++i) {
        if (desc->blob_bytes[i] <= 0 || !desc->blob_ptrs[i]) B2P_FAIL(B2P_E_ARG, "blob %d is empty", i);
        if (cudaMalloc(&h->blobs[i], static_cast<size_t>(desc->blob_bytes[i])) != cudaSuccess)
            B2P_FAIL(B2P_E_NOMEM, "cudaMalloc(%lld) for blob %d failed", (long long)desc->blob_bytes[i], i);
        if (cudaMemcpy(h->blobs[i], desc->blob_ptrs[i], static_cast<size_t>(desc->blob_bytes[i]),
                       cudaMemcpyHostToDevice) != cudaSuccess)
            B2P_FAIL(B2P_E_CUDA, "weight upload of blob %d failed", i);
    }
    // workspace slots (sized for one chunk)
    h->slots.assign(desc->n_slots, nullptr);
    h->slot_elems.assign(desc->slot_elems, desc->slot_elems + desc->n_slots);
    for (int i = 1; i < desc->n_slots; ++i) {
        const size_t bytes = static_cast<size_t>(desc->slot_elems[i]) * 2 * h->chunk;
        if (bytes == 0) continue;
        if (cudaMalloc(&h->slots[i], bytes) != cudaSuccess)
            B2P_FAIL(B2P_E_NOMEM, "cudaMalloc(%zu) for slot %d failed", bytes, i);
        cudaMemset(h->slots[i], 0, bytes);
    }
    // op plans
    const bool autotune = getenv("B2P_AUTOTUNE") != nullptr && h->chunk >= 16;   // opt-in: measured gain is within noise
    h->ops.resize(desc->n_ops);
    for (int i = 0; i < desc->n_ops; ++i) {
        OpPlan& p = h->ops[i];
        p.op = desc->ops[i];
        const b2p_op& o = p.op;
        if (o.kind == B2P_OP_CONV || o.kind == B2P_OP_LINEAR) {
            p.is_gemm = true;
            p.src_external = (o.src == B2P_SLOT_INPUT);
            const int rows_cap = o.kind == B2P_OP_LINEAR ? h->chunk : h->chunk * o.out_h * o.out_w;
            if (!gemm_plan_init(p.gemm, o, h->blobs[o.weight], h->blob_bytes[o.weight],
                                static_cast<const float*>(h->blobs[o.bias]), rows_cap, h->num_sms, eb, sizeof eb))
                B2P_FAIL(B2P_E_ARG, "op %d: %s", i, eb);
            if (h->blob_bytes[o.bias] < static_cast<int64_t>(p.gemm.n_pad) * 4)
                B2P_FAIL(B2P_E_ARG, "op %d: bias blob shorter than padded out_c", i);
            if (!p.src_external) {
                if (!h->slots[o.src]) B2P_FAIL(B2P_E_ARG, "op %d reads an empty slot", i);
                if (!gemm_plan_set_src(p.gemm, h->slots[o.src], h->chunk, eb, sizeof eb))
                    B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
            }
            if (conv3x3_c64_supported(o) && h->blob_bytes[o.weight] == 64 * 9 * 64 * 2 && getenv("B2P_NO_C3") == nullptr) {
                // 56x56x64 3x3: patch-resident kernel instead of the nine-fold im2col re-read
                p.c3.weight = h->blobs[o.weight];
                p.c3.bias = static_cast<const float*>(h->blobs[o.bias]);
                p.use_c3 = true;
                const bool ext = (o.src == B2P_SLOT_INPUT) || (o.dst == B2P_SLOT_OUTPUT) || (o.res == B2P_SLOT_INPUT);
                if (!ext) {
                    const void* r_ = o.res == B2P_SLOT_NONE ? nullptr : h->slots[o.res];
                    if (!h->slots[o.src] || !h->slots[o.dst] || (o.res != B2P_SLOT_NONE && !r_))
                        B2P_FAIL(B2P_E_ARG, "op %d uses an empty slot", i);
                    if (!conv3x3_c64_plan_set(p.c3, h->slots[o.src], h->slots[o.dst], r_, h->chunk, eb, sizeof eb))
                        B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
                }
            }
            if (conv3x3_c128_supported(o) && h->blob_bytes[o.weight] == 128 * 9 * 128 * 2 && getenv("B2P_NO_C3B") == nullptr) {
                // 28x28x128 3x3: the patch is fetched once as tiled boxes instead of nine im2col gathers
                p.c3b.weight = h->blobs[o.weight];
                p.c3b.bias = static_cast<const float*>(h->blobs[o.bias]);
                p.use_c3b = true;
                if (o.src != B2P_SLOT_INPUT && o.dst != B2P_SLOT_OUTPUT) {
                    if (!h->slots[o.src] || !h->slots[o.dst]) B2P_FAIL(B2P_E_ARG, "op %d uses an empty slot", i);
                    if (!conv3x3_c128_plan_set(p.c3b, h->slots[o.src], h->slots[o.dst], h->chunk, eb, sizeof eb))
                        B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
                }
            }
            if (p.gemm.src2_c > 0) {
                if (!h->slots[o.src2]) B2P_FAIL(B2P_E_ARG, "op %d reads an empty second source", i);
                if (!gemm_plan_set_src2(p.gemm, h->slots[o.src2], h->chunk, eb, sizeof eb))
                    B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
            }
            p.dst_external = (o.dst == B2P_SLOT_OUTPUT);
            p.res_external = (o.res == B2P_SLOT_INPUT);
            if (o.kind == B2P_OP_LINEAR && o.res != B2P_SLOT_NONE) B2P_FAIL(B2P_E_ARG, "op %d: LINEAR has no residual", i);
            if (!p.dst_external) {
                if (!h->slots[o.dst]) B2P_FAIL(B2P_E_ARG, "op %d writes an empty slot", i);
                if (h->slot_elems[o.dst] < static_cast<int64_t>(o.out_h) * o.out_w * o.out_c && o.kind == B2P_OP_CONV)
                    B2P_FAIL(B2P_E_ARG, "op %d: destination slot too small", i);
                if (!gemm_plan_set_dst(p.gemm, h->slots[o.dst], rows_cap, eb, sizeof eb))
                    B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
            }
            if (o.res != B2P_SLOT_NONE && !p.res_external) {
                if (!h->slots[o.res]) B2P_FAIL(B2P_E_ARG, "op %d adds an empty slot", i);
                if (!gemm_plan_set_res(p.gemm, h->slots[o.res], rows_cap, eb, sizeof eb))
                    B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
            }
            // Tile-width autotune: when every operand lives in handle-owned slots, time each legal tile
            // width on the real shapes (contents are irrelevant) and keep the fastest.
            if (autotune && o.kind == B2P_OP_CONV && !p.src_external && !p.dst_external && !p.res_external &&
                p.gemm.block_k == 64) {
                const bool has_res = o.res != B2P_SLOT_NONE;
                float best_ms = 0.f;
                int best_bn = 0;
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                for (int bn : {256, 128, 64}) {
                    GemmPlan cand;
                    cand.force_block_n = bn;
                    char e2[256] = "";
                    if (!gemm_plan_init(cand, o, h->blobs[o.weight], h->blob_bytes[o.weight],
                                        static_cast<const float*>(h->blobs[o.bias]), rows_cap, h->num_sms, e2, sizeof e2) ||
                        !gemm_plan_set_src(cand, h->slots[o.src], h->chunk, e2, sizeof e2) ||
                        (cand.src2_c > 0 && !gemm_plan_set_src2(cand, h->slots[o.src2], h->chunk, e2, sizeof e2)) ||
                        !gemm_plan_set_dst(cand, h->slots[o.dst], rows_cap, e2, sizeof e2) ||
                        (has_res && !gemm_plan_set_res(cand, h->slots[o.res], rows_cap, e2, sizeof e2)))
                        continue;
                    float ms_best_of = 1e30f;
                    bool ok = true;
                    for (int rep = 0; rep < 4 && ok; ++rep) {
                        cudaEventRecord(e0, 0);
                        ok = gemm_plan_launch(cand, rows_cap, has_res, nullptr, h->num_sms, 0) == cudaSuccess;
                        cudaEventRecord(e1, 0);
                        ok = ok && cudaEventSynchronize(e1) == cudaSuccess;
                        float ms = 0.f;
                        cudaEventElapsedTime(&ms, e0, e1);
                        if (rep > 0 && ms < ms_best_of) ms_best_of = ms;
                    }
                    if (ok && (best_bn == 0 || ms_best_of < best_ms)) { best_bn = bn; best_ms = ms_best_of; }
                }
                cudaEventDestroy(e0);
                cudaEventDestroy(e1);
                if (best_bn != 0 && best_bn != p.gemm.block_n) {
                    GemmPlan tuned;
                    tuned.force_block_n = best_bn;
                    if (gemm_plan_init(tuned, o, h->blobs[o.weight], h->blob_bytes[o.weight],
                                       static_cast<const float*>(h->blobs[o.bias]), rows_cap, h->num_sms, eb, sizeof eb) &&
                        gemm_plan_set_src(tuned, h->slots[o.src], h->chunk, eb, sizeof eb) &&
                        (tuned.src2_c == 0 || gemm_plan_set_src2(tuned, h->slots[o.src2], h->chunk, eb, sizeof eb)) &&
                        gemm_plan_set_dst(tuned, h->slots[o.dst], rows_cap, eb, sizeof eb) &&
                        (!has_res || gemm_plan_set_res(tuned, h->slots[o.res], rows_cap, eb, sizeof eb)))
                        p.gemm = tuned;
                }
                if (getenv("B2P_VERBOSE"))
                    fprintf(stderr, "[b2p] op %d %dx%dx%d->%d k%d s%d res%d: tile width %d (%.1f us)\n", i, o.in_h, o.in_w,
                            o.in_c, o.out_c, o.kh, o.stride, (int)has_res, p.gemm.block_n, best_ms * 1e3f);
            }
        } else if (o.kind == B2P_OP_STEM_CONV) {
            if (o.in_h != 112 || o.in_w != 112 || o.in_c != 16 || o.out_c != 64 || o.kh != 4 || o.kw != 4 ||
                h->blob_bytes[o.weight] != 64 * 16 * 16 * 2 || h->blob_bytes[o.bias] < 64 * 4)
                B2P_FAIL(B2P_E_ARG, "op %d: STEM_CONV expects the 112x112x16 s2d stem with a [64][4][4][16] weight", i);
            std::vector<uint16_t> image(64 * 16 * 16);
            stem_pack_weight_image(static_cast<const uint16_t*>(desc->blob_ptrs[o.weight]), image.data());
            if (cudaMalloc(&p.stem_image, image.size() * 2) != cudaSuccess ||
                cudaMemcpy(p.stem_image, image.data(), image.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess)
                B2P_FAIL(B2P_E_NOMEM, "op %d: stem weight image upload failed", i);
            p.stem.w_image = static_cast<const uint8_t*>(p.stem_image);
            p.stem.bias = static_cast<const float*>(h->blobs[o.bias]);
            p.stem.pool = o.reserved[0];
            if (o.src != B2P_SLOT_INPUT && o.dst != B2P_SLOT_OUTPUT) {
                if (!h->slots[o.src] || !h->slots[o.dst]) B2P_FAIL(B2P_E_ARG, "op %d uses an empty slot", i);
                if (!stem_plan_init(p.stem, h->slots[o.src], h->slots[o.dst], h->chunk, o.reserved[0], eb, sizeof eb))
                    B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
            }
        } else if (o.kind == B2P_OP_CONV_PAIR) {
            const int n2 = o.reserved[3];
            const int64_t rows_cap = static_cast<int64_t>(h->chunk) * o.out_h * o.out_w;
            const int k2 = o.src2_c > 0 ? o.src2_c : 0;
            if (o.kh != 1 || o.kw != 1 || o.stride != 1 || o.pad_lo != 0 || !pair_plan_supported(o.in_c, o.out_c, n2) ||
                h->blob_bytes[o.weight] != static_cast<int64_t>(o.out_c) * (o.in_c + k2) * 2 ||
                h->blob_bytes[o.reserved[1]] != static_cast<int64_t>(n2) * o.out_c * 2 ||
                h->blob_bytes[o.bias] < o.out_c * 4 || h->blob_bytes[o.reserved[2]] < n2 * 4)
                B2P_FAIL(B2P_E_ARG, "op %d: malformed CONV_PAIR", i);
            void* s_src = h->slots[o.src]; void* s_dst = h->slots[o.dst];
            void* s_res = o.res == B2P_SLOT_NONE ? nullptr : h->slots[o.res];
            void* s_dst2 = h->slots[o.reserved[0]];
            void* s_src2 = k2 > 0 ? h->slots[o.src2] : nullptr;
            if (!s_src || (o.res != B2P_SLOT_NONE && !s_res) || !s_dst || !s_dst2 || (k2 > 0 && !s_src2))
                B2P_FAIL(B2P_E_ARG, "op %d uses an empty slot", i);
            const int sub = o.reserved2 == 2 ? 2 : 1;     // 2: dst keeps the even rows / columns only
            if (o.reserved2 != 0 && o.reserved2 != 1 && o.reserved2 != 2) B2P_FAIL(B2P_E_ARG, "op %d: CONV_PAIR reserved2 must be 0, 1 or 2", i);
            if (sub == 2 && ((o.out_h | o.out_w) & 1)) B2P_FAIL(B2P_E_ARG, "op %d: a subsampled CONV_PAIR output needs even out_h / out_w", i);
            if (h->slot_elems[o.dst] < static_cast<int64_t>(o.out_h / sub) * (o.out_w / sub) * o.out_c ||
                h->slot_elems[o.reserved[0]] < static_cast<int64_t>(o.out_h) * o.out_w * n2)
                B2P_FAIL(B2P_E_ARG, "op %d: destination slot too small", i);
            if (!pair_plan_init(p.pair, o.in_c, k2, o.out_c, n2, h->blobs[o.weight], static_cast<const float*>(h->blobs[o.bias]),
                                h->blobs[o.reserved[1]], static_cast<const float*>(h->blobs[o.reserved[2]]), eb, sizeof eb) ||
                !pair_plan_set_src2(p.pair, s_src2, h->chunk, o.out_h, o.out_w, o.src2_stride > 0 ? o.src2_stride : 1, eb, sizeof eb) ||
                !pair_plan_set_buffers(p.pair, s_src, s_res, s_dst, s_dst2, static_cast<int>(rows_cap), sub, eb, sizeof eb))
                B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
        } else if (o.kind == B2P_OP_BOTTLENECK) {
            const bool proj = o.res == B2P_SLOT_NONE;
            const int mid = o.reserved2;
            const int64_t k3 = proj ? mid + o.in_c : mid;
            if (!bneck56_supported(o.in_h, o.in_w, o.in_c, mid, o.out_c, proj) || o.out_h != o.in_h || o.out_w != o.in_w ||
                h->blob_bytes[o.weight] != static_cast<int64_t>(mid) * o.in_c * 2 ||
                h->blob_bytes[o.reserved[0]] != static_cast<int64_t>(mid) * 9 * mid * 2 ||
                h->blob_bytes[o.reserved[2]] != static_cast<int64_t>(o.out_c) * k3 * 2 ||
                h->blob_bytes[o.bias] < mid * 4 || h->blob_bytes[o.reserved[1]] < mid * 4 ||
                h->blob_bytes[o.reserved[3]] < o.out_c * 4)
                B2P_FAIL(B2P_E_ARG, "op %d: malformed BOTTLENECK (56x56, width 64, 256 outputs; identity needs in_c 256, projection in_c 64)", i);
            if (!h->slots[o.src] || !h->slots[o.dst]) B2P_FAIL(B2P_E_ARG, "op %d uses an empty slot", i);
            if (!bneck56_plan_init(p.bneck, o.in_c, proj, h->blobs[o.weight], static_cast<const float*>(h->blobs[o.bias]),
                                   h->blobs[o.reserved[0]], static_cast<const float*>(h->blobs[o.reserved[1]]),
                                   h->blobs[o.reserved[2]], static_cast<const float*>(h->blobs[o.reserved[3]]), eb, sizeof eb) ||
                !bneck56_plan_set(p.bneck, h->slots[o.src], h->slots[o.dst], h->chunk, eb, sizeof eb))
                B2P_FAIL(B2P_E_CUDA, "op %d: %s", i, eb);
        } else if (o.kind == B2P_OP_ATTENTION) {
            if (o.src != B2P_SLOT_INPUT) {
                if (!h->slots[o.src]) B2P_FAIL(B2P_E_ARG, "op %d reads an empty slot", i);
                if (!attention_plan_init(p.att, h->slots[o.src], h->chunk, o.in_h, o.in_c, o.reserved[0], eb, sizeof eb))
                    B2P_FAIL(B2P_E_ARG, "op %d: %s", i, eb);
            }
        } else if (o.kind == B2P_OP_LAYERNORM || o.kind == B2P_OP_TOKENS) {
            if (o.in_c != 768) B2P_FAIL(B2P_E_ARG, "op %d: token width must be 768", i);
            const int64_t need_w = o.kind == B2P_OP_TOKENS ? static_cast<int64_t>(o.in_h) * o.in_c * 4 : o.in_c * 4;
            if (h->blob_bytes[o.weight] < need_w || h->blob_bytes[o.bias] < o.in_c * 4)
                B2P_FAIL(B2P_E_ARG, "op %d: parameter blob too small", i);
        } else if ((o.kind == B2P_OP_MAXPOOL || o.kind == B2P_OP_AVGPOOL) && o.in_c % 8 != 0) {
            B2P_FAIL(B2P_E_ARG, "op %d: pooling needs channels %% 8 == 0", i);
        }
        if (o.kind == B2P_OP_LINEAR && o.dst == B2P_SLOT_OUTPUT) h->classes = o.out_c;
    }
    // L2 cache policies (B2P_L2_HINTS: 0 = none, 1 = default, 2 = also stream very large outputs).  An operand at
    // its LAST use (no later op reads the slot before it is overwritten) is loaded evict-first: it is dead afterwards,
    // and without the hint it would push the tensor being WRITTEN - which the next layer reads - out of the L2.
    // Applied to residuals and to the inputs of 1x1 layers (read once); 3x3 inputs are re-read per tap and stay normal.
    {
        const char* e = getenv("B2P_L2_HINTS");
        const int level = e ? atoi(e) : 1;
        const uint64_t EVICT_FIRST = 0x12F0000000000000ull;
        auto last_use = [&](int i, int slot) {
            if (slot == B2P_SLOT_NONE || slot == B2P_SLOT_OUTPUT) return false;
            for (int j = i + 1; j < desc->n_ops; ++j) {
                const b2p_op& o = desc->ops[j];
                if (o.src == slot || o.res == slot || (o.src2_c > 0 && o.src2 == slot)) return false;
                if (o.dst == slot || (o.kind == B2P_OP_CONV_PAIR && o.reserved[0] == slot)) return true;
            }
            return true;
        };
        for (int i = 0; i < desc->n_ops && level > 0; ++i) {
            OpPlan& p = h->ops[i];
            const b2p_op& o = p.op;
            if (!p.is_gemm || p.use_c3 || p.use_c3b) continue;
            if (o.res != B2P_SLOT_NONE && last_use(i, o.res)) p.gemm.pol_res = EVICT_FIRST;
            if (!p.gemm.im2col && o.src != o.res && last_use(i, o.src)) p.gemm.pol_a = EVICT_FIRST;
            if (level >= 2 && o.kind == B2P_OP_CONV &&
                static_cast<int64_t>(h->chunk) * o.out_h * o.out_w * o.out_c * 2 > (110ll << 20))
                p.gemm.pol_out = EVICT_FIRST;
        }
    }
    // K1 -> stem fusion for the predict chain: possible when the program opens with the space-to-depth pack of the
    // 224x224 input into a workspace slot, no later op reads the external input, and the batch is not chunked
    if (desc->n_ops > 1 && getenv("B2P_NO_FUSED_INPUT") == nullptr) {
        const b2p_op& o0 = desc->ops[0];
        bool ok = o0.kind == B2P_OP_STEM_PACK && o0.src == B2P_SLOT_INPUT && o0.dst > 0 && o0.in_h == B2P_OUT_HW &&
                  o0.in_w == B2P_OUT_HW && h->chunk == h->max_batch;
        for (int i = 1; i < desc->n_ops && ok; ++i) {
            const b2p_op& o = desc->ops[i];
            ok = o.src != B2P_SLOT_INPUT && o.res != B2P_SLOT_INPUT && !(o.src2_c > 0 && o.src2 == B2P_SLOT_INPUT);
        }
        if (ok) h->fused_input_slot = o0.dst;
    }
    // predict-chain staging
    const size_t nb = static_cast<size_t>(h->max_batch);
    for (int r = 0; r < b2p_handle::RING; ++r) {
        if (cudaMalloc(&h->imgs_dev[r], sizeof(b2p_image) * nb) != cudaSuccess ||
            cudaMallocHost(&h->imgs_pinned[r], sizeof(b2p_image) * nb) != cudaSuccess ||
            cudaEventCreateWithFlags(&h->imgs_done[r], cudaEventDisableTiming) != cudaSuccess)
            B2P_FAIL(B2P_E_NOMEM, "descriptor staging allocation failed");
    }
    if (cudaMalloc(&h->x_dev, nb * 3 * B2P_OUT_HW * B2P_OUT_HW * 2) != cudaSuccess ||
        cudaMalloc(&h->logits_dev, nb * static_cast<size_t>(h->out_elems > 0 ? h->out_elems : 1) * 4) != cudaSuccess ||
        cudaMalloc(&h->probs_dev, nb * B2P_MAX_TOPK * 4) != cudaSuccess ||
        cudaMalloc(&h->idx_dev, nb * B2P_MAX_TOPK * 4) != cudaSuccess)
        B2P_FAIL(B2P_E_NOMEM, "staging allocation failed");
    if (cudaMalloc(&h->norm_lut, 3 * 256 * sizeof(float)) != cudaSuccess || launch_norm_lut(h->norm_lut, 0) != cudaSuccess)
        B2P_FAIL(B2P_E_NOMEM, "normalisation table set-up failed");
    for (auto& e : h->ev)
        if (cudaEventCreate(&e) != cudaSuccess) B2P_FAIL(B2P_E_CUDA, "cudaEventCreate failed");
    for (auto& set : h->sev)
        for (auto& e : set)
            if (cudaEventCreate(&e) != cudaSuccess) B2P_FAIL(B2P_E_CUDA, "cudaEventCreate failed");
    if (cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) != cudaSuccess)
        B2P_FAIL(B2P_E_CUDA, "copy stream creation failed");
    for (int i = 0; i < 2; ++i)
        if (cudaEventCreateWithFlags(&h->ev_copied[i], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&h->ev_consumed[i], cudaEventDisableTiming) != cudaSuccess)
            B2P_FAIL(B2P_E_CUDA, "cudaEventCreate failed");
    h->ev_valid = true;
    if (cudaDeviceSynchronize() != cudaSuccess) B2P_FAIL(B2P_E_CUDA, "device sync after create failed");
#undef B2P_FAIL
    *out = h;
    return B2P_OK;
}

// Per-stage times of the last b2p_predict_host call are resolved here, after the caller has
// synchronised the stream; events that have not completed yet leave the previous values in place.
int32_t b2p_stats_get(b2p_handle* h, b2p_stats* out) {
    if (!h || !out) return B2P_E_ARG;
    if (h->ev_valid && h->sev_unread > 0) {
        // stage times: mean over the calls since the previous read whose work has completed
        double sum[3] = {0, 0, 0};
        int got = 0;
        for (int j = 0; j < h->sev_unread; ++j) {
            const int r = ((h->sev_pos - 1 - j) % b2p_handle::SRING + b2p_handle::SRING) % b2p_handle::SRING;
            float ms[3];
            bool ok = true;
            for (int i = 0; i < 3 && ok; ++i) ok = cudaEventElapsedTime(&ms[i], h->sev[r][i], h->sev[r][i + 1]) == cudaSuccess;
            if (!ok) { cudaGetLastError(); continue; }
            for (int i = 0; i < 3; ++i) sum[i] += ms[i];
            ++got;
        }
        if (got > 0) {
            h->stats.ms_preprocess = static_cast<float>(sum[0] / got);
            h->stats.ms_forward = static_cast<float>(sum[1] / got);
            h->stats.ms_epilogue = static_cast<float>(sum[2] / got);
            h->sev_unread = 0;
        }
    }
    if (h->ev_valid && h->stats.calls > 0 && h->sev_last >= 0) {
        float a = 0.f, b = 0.f;
        if (cudaEventElapsedTime(&a, h->ev[0], h->ev[1]) == cudaSuccess &&
            cudaEventElapsedTime(&b, h->ev[5], h->ev[6]) == cudaSuccess) {
            h->stats.ms_h2d = a;
            h->stats.ms_d2h = b;
        } else {
            cudaGetLastError();
        }
    }
    *out = h->stats;
    return B2P_OK;
}

int32_t b2p_preprocess(b2p_handle* h, const uint8_t* slab, const b2p_image* imgs, int32_t n, void* out,
                       int32_t out_f32, void* stream) {
    if (!h || !slab || !imgs || !out || n <= 0) return B2P_E_ARG;
    if (n > h->max_batch) { set_error(h, "n=%d exceeds max_batch=%d", n, h->max_batch); return B2P_E_LIMIT; }
    B2P_CUDA(h, cudaSetDevice(h->device));
    int rc = check_images(h, imgs, n);
    if (rc != B2P_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    rc = stage_images(h, imgs, n, st);
    if (rc != B2P_OK) return rc;
    int launches = 0;
    B2P_CUDA(h, launch_preprocess(slab, h->cur_imgs_dev, imgs, n, out, out_f32, h->norm_lut, st, &launches));
    h->stats.kernel_launches += launches;
    return B2P_OK;
}

int32_t b2p_forward(b2p_handle* h, const void* x, int32_t n, void* out, void* stream) {
    if (!h || !x || !out || n <= 0) return B2P_E_ARG;
    if (n > h->max_batch) { set_error(h, "n=%d exceeds max_batch=%d", n, h->max_batch); return B2P_E_LIMIT; }
    B2P_CUDA(h, cudaSetDevice(h->device));
    return run_program(h, x, n, out, static_cast<cudaStream_t>(stream));
}

int32_t b2p_softmax_topk(b2p_handle* h, const float* logits, int32_t n, int32_t classes, int32_t k,
                         float* probs, int32_t* idx, void* stream) {
    if (!h || !logits || !probs || !idx || n <= 0 || classes <= 0) return B2P_E_ARG;
    if (k <= 0 || k > B2P_MAX_TOPK || k > classes) { set_error(h, "k=%d out of range", k); return B2P_E_LIMIT; }
    B2P_CUDA(h, cudaSetDevice(h->device));
    B2P_CUDA(h, launch_softmax_topk(logits, n, classes, k, probs, idx, static_cast<cudaStream_t>(stream)));
    h->stats.kernel_launches += 1;
    return B2P_OK;
}

// K1 of the predict chain.  Crop-only batches of a ResNet-style program are written straight into the stem's
// space-to-depth slot (one pass instead of K1 + pack: bit-identical values); *first_op says where K2 then starts.
static int chain_preprocess(b2p_handle* h, const uint8_t* slab, const b2p_image* imgs, int32_t n, void* stream,
                            size_t* first_op) {
    *first_op = 0;
    if (h->fused_input_slot < 0 || !imgs || n <= 0 || n > h->max_batch || !preprocess_all_identity(imgs, n))
        return b2p_preprocess(h, slab, imgs, n, h->x_dev, 0, stream);
    if (!slab) return B2P_E_ARG;
    B2P_CUDA(h, cudaSetDevice(h->device));
    int rc = check_images(h, imgs, n);
    if (rc != B2P_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    rc = stage_images(h, imgs, n, st);
    if (rc != B2P_OK) return rc;
    B2P_CUDA(h, launch_preprocess_s2d(slab, h->cur_imgs_dev, n, h->slots[h->fused_input_slot], h->norm_lut, st));
    h->stats.kernel_launches += 1;
    *first_op = 1;
    return B2P_OK;
}

int32_t b2p_predict(b2p_handle* h, const uint8_t* slab, const b2p_image* imgs, int32_t n, int32_t k,
                    float* probs, int32_t* idx, void* stream) {
    if (!h || h->classes <= 0) { if (h) set_error(h, "program has no classifier output"); return B2P_E_ARG; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    B2P_CUDA(h, cudaSetDevice(h->device));
    cudaEvent_t* sev = h->sev[h->sev_pos];
    const bool sample = (h->predict_calls++ % h->stage_every) == 0;
    if (sample) B2P_CUDA(h, cudaEventRecord(sev[0], st));
    size_t first_op = 0;
    int rc = chain_preprocess(h, slab, imgs, n, stream, &first_op);
    if (rc != B2P_OK) return rc;
    if (sample) B2P_CUDA(h, cudaEventRecord(sev[1], st));
    rc = run_program(h, h->x_dev, n, h->logits_dev, st, first_op);
    if (rc != B2P_OK) return rc;
    if (sample) B2P_CUDA(h, cudaEventRecord(sev[2], st));
    rc = b2p_softmax_topk(h, h->logits_dev, n, h->classes, k, probs, idx, stream);
    if (rc != B2P_OK) return rc;
    if (sample) {
        B2P_CUDA(h, cudaEventRecord(sev[3], st));
        h->sev_pos = (h->sev_pos + 1) % b2p_handle::SRING;
        if (h->sev_unread < b2p_handle::SRING) ++h->sev_unread;
    }
    return B2P_OK;
}

int32_t b2p_predict_host(b2p_handle* h, const uint8_t* slab_host, int64_t slab_bytes, const b2p_image* imgs,
                         int32_t n, int32_t k, float* probs_host, int32_t* idx_host, void* stream) {
    if (!h || !slab_host || slab_bytes <= 0 || !imgs || !probs_host || !idx_host || n <= 0) return B2P_E_ARG;
    if (h->classes <= 0) { set_error(h, "program has no classifier output"); return B2P_E_ARG; }
    if (n > h->max_batch) { set_error(h, "n=%d exceeds max_batch=%d", n, h->max_batch); return B2P_E_LIMIT; }
    if (k <= 0 || k > B2P_MAX_TOPK) { set_error(h, "k=%d out of range", k); return B2P_E_LIMIT; }
    for (int i = 0; i < n; ++i) {
        // every descriptor must lie inside the slab that is about to be copied (K1 reads h*w*3 bytes at offset)
        if (imgs[i].offset < 0 || imgs[i].h <= 0 || imgs[i].w <= 0 ||
            imgs[i].offset + static_cast<int64_t>(imgs[i].h) * imgs[i].w * 3 > slab_bytes) {
            set_error(h, "image %d: [%lld, +%d x %d x 3) is outside the %lld-byte slab", i, (long long)imgs[i].offset,
                      imgs[i].h, imgs[i].w, (long long)slab_bytes);
            return B2P_E_ARG;
        }
    }
    B2P_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int si = h->slab_idx;
    h->slab_idx ^= 1;
    if (slab_bytes > h->slab_cap[si]) {
        // grow-only device slab; the first calls (or a larger batch) pay for the allocation
        B2P_CUDA(h, cudaStreamSynchronize(st));
        B2P_CUDA(h, cudaStreamSynchronize(h->copy_stream));
        if (h->slab_dev[si]) cudaFree(h->slab_dev[si]);
        h->slab_dev[si] = nullptr;
        h->slab_cap[si] = 0;
        h->slab_used[si] = false;
        const int64_t cap = ((slab_bytes + (1 << 20) - 1) >> 20) << 20;
        if (cudaMalloc(&h->slab_dev[si], static_cast<size_t>(cap)) != cudaSuccess) {
            set_error(h, "cudaMalloc(%lld) for the image slab failed", (long long)cap);
            return B2P_E_NOMEM;
        }
        h->slab_cap[si] = cap;
    }
    // copy stream: wait until the kernels of two calls ago have consumed this slab, then copy
    if (h->slab_used[si]) B2P_CUDA(h, cudaStreamWaitEvent(h->copy_stream, h->ev_consumed[si], 0));
    B2P_CUDA(h, cudaEventRecord(h->ev[0], h->copy_stream));
    B2P_CUDA(h, cudaMemcpyAsync(h->slab_dev[si], slab_host, static_cast<size_t>(slab_bytes), cudaMemcpyHostToDevice,
                                h->copy_stream));
    B2P_CUDA(h, cudaEventRecord(h->ev[1], h->copy_stream));
    B2P_CUDA(h, cudaEventRecord(h->ev_copied[si], h->copy_stream));
    B2P_CUDA(h, cudaStreamWaitEvent(st, h->ev_copied[si], 0));
    cudaEvent_t* sev = h->sev[h->sev_pos];
    const bool sample = (h->predict_calls++ % h->stage_every) == 0;
    if (sample) B2P_CUDA(h, cudaEventRecord(sev[0], st));
    size_t first_op = 0;
    int rc = chain_preprocess(h, h->slab_dev[si], imgs, n, stream, &first_op);
    if (rc != B2P_OK) return rc;
    if (sample) B2P_CUDA(h, cudaEventRecord(sev[1], st));
    // "K1 has read slab si": the copy of two calls later waits for this.  (Signalling it after the whole chain instead,
    // to keep K1 -> stem launch overlap, cost 12 % of the end-to-end rate: the copies then run one step ahead, not two.)
    B2P_CUDA(h, cudaEventRecord(h->ev_consumed[si], st));
    h->slab_used[si] = true;
    rc = run_program(h, h->x_dev, n, h->logits_dev, st, first_op);
    if (rc != B2P_OK) return rc;
    if (sample) B2P_CUDA(h, cudaEventRecord(sev[2], st));
    rc = b2p_softmax_topk(h, h->logits_dev, n, h->classes, k, h->probs_dev, h->idx_dev, stream);
    if (rc != B2P_OK) return rc;
    if (sample) {
        B2P_CUDA(h, cudaEventRecord(sev[3], st));
        h->sev_pos = (h->sev_pos + 1) % b2p_handle::SRING;
        if (h->sev_unread < b2p_handle::SRING) ++h->sev_unread;
    }
    h->sev_last = 0;
    B2P_CUDA(h, cudaEventRecord(h->ev[5], st));
    B2P_CUDA(h, cudaMemcpyAsync(probs_host, h->probs_dev, sizeof(float) * static_cast<size_t>(n) * k,
                                cudaMemcpyDeviceToHost, st));
    B2P_CUDA(h, cudaMemcpyAsync(idx_host, h->idx_dev, sizeof(int32_t) * static_cast<size_t>(n) * k,
                                cudaMemcpyDeviceToHost, st));
    B2P_CUDA(h, cudaEventRecord(h->ev[6], st));
    h->stats.calls += 1;
    return B2P_OK;
}

const float* b2p_last_logits(b2p_handle* h) { return h ? h->logits_dev : nullptr; }

}  // extern "C"

