// Baseline-JPEG decode on the GPU: the step immediately before the predict path.
//
// Replaces, for the files it covers, the pixels `Image.open(image_file.file)` (reference src/model/model.py:23)
// would hand to the preprocess stage: uint8 RGB, HWC.  The arithmetic is libjpeg-turbo's default decode pipeline
// (what Pillow runs), restated in oracle/jpeg_np.py and checked there bit-for-bit against Pillow:
//
//   J1 huffman_kernel   entropy decoding (T.81 F.2.2)            one thread per restart segment -> int16 blocks
//   J2 idct_kernel      dequantise + accurate integer IDCT       one thread per 8x8 block       -> uint8 planes
//   J3 colour_kernel    fancy chroma upsampling + YCbCr -> RGB   one thread per 4 output pixels -> uint8 HWC
//
// Scope: SOF0/SOF1, 8-bit, one interleaved scan, 1 or 3 components, luma 1x1 / 2x1 / 2x2 with 1x1 chroma, no Adobe
// marker.  b2p_jpeg_inspect() says B2P_E_UNSUPPORTED for everything else and the caller keeps those files on the
// host decoder (row D of SURVEY.md §8(a)); nothing here falls back to the CPU.
//
// The marker parser runs on the calling host thread (it is a few hundred bytes of header per file); everything
// that touches pixel or coefficient data runs on the device.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <array>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <new>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/b200predict.h"

namespace {

// ------------------------------------------------------------------------------------------- shared layouts
constexpr int LOOK_BITS = 11;

struct HuffDev {                  // one Huffman table in decode form
    uint16_t look[1 << LOOK_BITS];  // (code length << 8) | symbol for codes of <= LOOK_BITS bits, 0 = longer
    int32_t maxcode[18];            // largest code of length l (l = 1..16), -1 if none; [17] = sentinel
    int32_t valoff[18];             // index of the first symbol of length l minus its smallest code
    uint8_t vals[256];
};

struct ImageDev {
    int32_t width, height, ncomp;
    int32_t hs, vs;               // luma sampling factors (chroma is 1x1)
    int32_t mcux, mcuy;
    int32_t dc_tab[3], ac_tab[3]; // indices into the call's HuffDev array
    int32_t blocks_w[3], blocks_h[3];
    int64_t coef_off[3];          // int16 elements from the coefficient buffer base
    int64_t plane_off[3];         // bytes from the plane buffer base; plane pitch = blocks_w * 8
    int64_t out_off;              // bytes from the RGB slab base
    int64_t block_base;           // index of this image's first block in the call-wide block numbering
    uint16_t quant[3][64];        // natural order
};

struct SegmentDev {
    int32_t image;
    int32_t first_mcu, n_mcu;
    int32_t pad;
    int64_t begin;                // byte offset of the un-stuffed segment inside the scan buffer (4-byte aligned)
    int64_t nbits;                // its length in bits
};

__constant__ uint8_t c_zigzag[80] = {
    0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
    6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31,
    39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
    63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};   // guard entries for corrupt run lengths

const uint8_t h_zigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                              41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                              30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// ------------------------------------------------------------------------------------------- J1: entropy decode
// A Huffman-coded segment has no index: the position of symbol i is only known once symbols 0..i-1 are decoded.
// One CTA per restart segment breaks that chain with the self-synchronising property of Huffman streams:
//   1. the segment's bits are cut into one chunk per thread; thread 0 starts in the true state, every other thread
//      GUESSES (first bit of its chunk, first block of an MCU, DC position) and decodes its chunk without output,
//      recording the state it reaches at the chunk end and the number of blocks it completed;
//   2. every thread whose predecessor ended in a different state than the one it started from re-decodes from the
//      predecessor's end state; repeat until no thread changes.  Thread 0 is right from the start, so after j
//      rounds threads 0..j are provably right and the loop ends in at most blockDim rounds — measured 6-38 rounds
//      of 512-bit chunks on 224x224 files (it depends on the content: blocks that end without an end-of-block
//      code and repetitive content delay the moment a decoder started in the wrong state falls into step);
//   3. a prefix sum of the block counts gives every thread the index of its first block; it decodes its chunk once
//      more, now writing coefficients (DC as the difference it reads);
//   4. a prefix sum over each component's DC differences in decode order yields the DC values (T.81 F.2.1.3.1).
// The result is exactly what a sequential decoder produces; a stream that never synchronises only costs rounds.
//
// The host hands over each segment un-stuffed (FF 00 -> FF), 4-byte aligned and zero padded, so a bit position is a
// plain offset and the reader can take 32 bits at a time.
// Index of staged word j in shared memory: one word of skew per 16, so that lanes whose chunks start a multiple of
// 16 words apart (512-bit chunks) do not all hit the same bank.
__device__ __forceinline__ uint32_t stage_index(uint32_t j) { return j + (j >> 4); }

struct BitReader {
    const uint32_t* __restrict__ words;   // big-endian bit order inside each byte-swapped word
    uint32_t nwords;
    uint32_t next;                        // index of the next word to take
    uint64_t win;                         // bit 63 = next unread bit
    int have;                             // valid bits in win
    // The CTA's part of the stream staged in shared memory, already byte-swapped (stage_count == 0: not staged).
    // A refill is a dependent load on the critical path of every ~32 bits decoded: from shared memory it costs
    // ~30 cycles, from global memory (L2) several hundred.
    const uint32_t* stage;
    uint32_t stage_w0, stage_count;

    __device__ __forceinline__ uint32_t word(uint32_t i) const {
        const uint32_t j = i - stage_w0;
        if (j < stage_count) return stage[stage_index(j)];
        return i < nwords ? __byte_perm(__ldg(words + i), 0, 0x0123) : 0u;
    }
    __device__ __forceinline__ void seek(uint32_t bitpos) {
        next = bitpos >> 5;
        const int off = static_cast<int>(bitpos & 31u);
        win = static_cast<uint64_t>(word(next++)) << (32 + off);
        have = 32 - off;
    }
    __device__ __forceinline__ void refill() {           // afterwards at least 33 bits are valid
        if (have <= 32) {
            win |= static_cast<uint64_t>(word(next++)) << (32 - have);
            have += 32;
        }
    }
    __device__ __forceinline__ uint32_t position() const { return next * 32u - static_cast<uint32_t>(have); }
    __device__ __forceinline__ uint32_t peek(int bits) const { return static_cast<uint32_t>(win >> (64 - bits)); }
    __device__ __forceinline__ void skip(int bits) { win <<= bits; have -= bits; }
};

// The rare long code (more than LOOK_BITS bits): canonical length walk over the table in global memory
// (jdhuff.c: jpeg_huff_decode).  Returns (length << 8) | symbol.
__device__ __noinline__ uint32_t huff_long_code(uint64_t win, const HuffDev* __restrict__ t, int& bad) {
    int len = LOOK_BITS + 1;
    int code = static_cast<int>(win >> (64 - len));
    while (len <= 16 && code > __ldg(&t->maxcode[len])) {
        ++len;
        code = static_cast<int>(win >> (64 - len));
    }
    if (len > 16) {
        bad = 1;
        return 16u << 8;
    }
    return (static_cast<uint32_t>(len) << 8) | __ldg(&t->vals[(code + __ldg(&t->valoff[len])) & 255]);
}

struct DecodeState {
    uint32_t pos;     // bit position of the next symbol
    int blk;          // block inside the MCU
    int k;            // 0 = the DC symbol is next, else the zig-zag index of the next AC coefficient
};

struct SegmentCtx {
    const ImageDev* __restrict__ im;
    const HuffDev* __restrict__ tables;
    int bpm;          // blocks per MCU
    int luma;         // luma blocks per MCU (hs * vs)
    int hs;
    int first_mcu;
    int total_blocks;
    uint32_t nbits;
    const uint32_t* __restrict__ words;
    uint32_t nwords;
    const uint16_t* look;       // shared memory: [component][0 = DC, 1 = AC][512] lookahead entries
    const uint8_t* zigzag;      // shared memory copy of c_zigzag (divergent indices serialise in the constant cache)
    const uint32_t* stage;      // shared memory: this CTA's words of the stream (see BitReader), or nullptr
    uint32_t stage_w0, stage_count;
};

// Address of block `b` (segment-relative, decode order) in the coefficient planes.
__device__ __forceinline__ int16_t* block_ptr(const SegmentCtx& c, int16_t* __restrict__ coef, int b, int& comp) {
    const int mcu = c.first_mcu + b / c.bpm;
    const int j = b - (b / c.bpm) * c.bpm;
    const int my = mcu / c.im->mcux, mx = mcu - my * c.im->mcux;
    int by, bx;
    if (j < c.luma) {
        comp = 0;
        const int sy = j / c.hs, sx = j - sy * c.hs;
        const int vs = c.luma / c.hs;
        by = my * vs + sy;
        bx = mx * c.hs + sx;
    } else {
        comp = j - c.luma + 1;
        by = my;
        bx = mx;
    }
    return coef + c.im->coef_off[comp] + (static_cast<int64_t>(by) * c.im->blocks_w[comp] + bx) * 64;
}

// Decodes the symbols that START in [st.pos, end_bit).  WRITE = false: only tracks the state and counts completed
// blocks.  WRITE = true: stores coefficients for blocks first_block, first_block + 1, ... (DC as a difference) and
// stops after the segment's last block.
template <bool WRITE>
__device__ __forceinline__ void decode_range(const SegmentCtx& c, DecodeState& st, uint32_t end_bit, int& blocks,
                                             int first_block, int16_t* __restrict__ coef, int& bad, int& overrun) {
    BitReader br;
    br.words = c.words;
    br.nwords = c.nwords;
    br.stage = c.stage;
    br.stage_w0 = c.stage_w0;
    br.stage_count = c.stage_count;
    br.seek(st.pos);
    int blk = st.blk, k = st.k;
    int done = 0;
    int b = first_block;
    int comp = blk < c.luma ? 0 : blk - c.luma + 1;
    int16_t* __restrict__ out = nullptr;
    if (WRITE) {
        if (b >= c.total_blocks) return;
        out = block_ptr(c, coef, b, comp);
    }
    uint32_t pos = st.pos;
    // The 32 lanes of a warp walk 32 different pieces of the stream, so the loop body is written with selects
    // instead of branches wherever the two sides are short: DC and AC symbols take the same path, and the code and
    // its extra bits leave the window in one shift.
    while (pos < end_bit) {
        br.refill();
        const bool is_dc = k == 0;
        const uint32_t hi = static_cast<uint32_t>(br.win >> 32), lo = static_cast<uint32_t>(br.win);
        uint32_t e = c.look[(comp * 2 + (is_dc ? 0 : 1)) * (1 << LOOK_BITS) + (hi >> (32 - LOOK_BITS))];
        if (e == 0u) {
            const HuffDev* __restrict__ t = c.tables + (is_dc ? c.im->dc_tab[comp] : c.im->ac_tab[comp]);
            e = huff_long_code(br.win, t, bad);
        }
        const int len = static_cast<int>(e >> 8);                  // 1..16
        const int s = static_cast<int>(e & 15u);
        const int r = is_dc ? 0 : static_cast<int>((e >> 4) & 15u);
        const uint32_t after = __funnelshift_l(lo, hi, len);       // the 32 bits that follow the code
        const uint32_t raw = s ? after >> (32 - s) : 0u;
        const int v = raw < (1u << ((s - 1) & 31)) ? static_cast<int>(raw) - (1 << s) + 1 : static_cast<int>(raw);
        br.skip(len + s);                                          // at most 31 bits
        pos += static_cast<uint32_t>(len + s);
        const int kk = k + r;                                      // zig-zag position of this coefficient
        if (WRITE && s != 0) out[c.zigzag[kk]] = static_cast<int16_t>(v);
        k = s != 0 ? kk + 1 : (is_dc ? 1 : (r == 15 ? k + 16 : 64));
        if (k >= 64) {                       // block complete
            k = 0;
            ++done;
            if (++blk == c.bpm) blk = 0;
            comp = blk < c.luma ? 0 : blk - c.luma + 1;
            if (WRITE) {
                if (++b >= c.total_blocks) {
                    if (pos > c.nbits) overrun = 1;      // the last block needed bits the file does not have
                    break;
                }
                out = block_ptr(c, coef, b, comp);
            }
        }
    }
    st.pos = pos;
    st.blk = blk;
    st.k = k;
    blocks = done;
}

constexpr int HUFF_MAX_THREADS = 1024;
constexpr int HUFF_MIN_CHUNK_BITS = 512;      // default (measured best of 256..8192); B2P_JPEG_CHUNK_BITS overrides it

// Exclusive prefix sum of one int per thread over the CTA; `total` receives the sum of all.
__device__ __forceinline__ int cta_exclusive_scan(int v, int* s_warp, int& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    __syncthreads();                       // s_warp may still be read from a previous scan
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int base = 0, sum = 0;
    for (int w = 0; w < nwarps; ++w) {
        const int x = s_warp[w];
        if (w < warp) base += x;
        sum += x;
    }
    total = sum;
    return base + inc - v;
}

// ---- thread-block cluster helpers: a large segment is spread over up to 8 CTAs (8 SMs) that exchange chunk states
// through distributed shared memory instead of living on one SM's issue slots
__device__ __forceinline__ uint32_t cl_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cl_size() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cl_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// value of a 32-bit shared-memory variable (given by its address in THIS CTA) as CTA `rank` of the cluster holds it
__device__ __forceinline__ uint32_t cl_load(const void* local, uint32_t rank) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(local));
    uint32_t remote, v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(a), "r"(rank));
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(remote) : "memory");
    return v;
}

// Exclusive prefix sum over all threads of the cluster (CTA rank major); `slot` must not be reused within a launch.
template <bool CLUSTER>
__device__ __forceinline__ int group_exclusive_scan(int v, int* s_warp, int* s_tot, int slot, int& total) {
    int local_total = 0;
    const int base_local = cta_exclusive_scan(v, s_warp, local_total);
    if (!CLUSTER) {
        total = local_total;
        return base_local;
    }
    if (threadIdx.x == 0) s_tot[slot] = local_total;
    cl_sync();
    const uint32_t me = cl_rank(), n = cl_size();
    int base = 0, sum = 0;
    for (uint32_t r = 0; r < n; ++r) {
        const int x = static_cast<int>(cl_load(&s_tot[slot], r));
        if (r < me) base += x;
        sum += x;
    }
    total = sum;
    return base + base_local;
}

template <bool CLUSTER>
__global__ void __launch_bounds__(HUFF_MAX_THREADS)
huffman_kernel(const uint8_t* __restrict__ scan, const ImageDev* __restrict__ images,
               const SegmentDev* __restrict__ segments, const HuffDev* __restrict__ tables,
               int16_t* __restrict__ coef, int32_t* __restrict__ status, int min_chunk_bits, int debug_rounds,
               int stage_words) {
    extern __shared__ uint32_t s_stream[];          // stage_words (+ skew) words: this CTA's part of the stream
    __shared__ uint32_t s_pos[HUFF_MAX_THREADS];
    __shared__ int s_state[HUFF_MAX_THREADS];        // blk << 8 | k
    __shared__ int s_warp[HUFF_MAX_THREADS / 32];
    __shared__ uint16_t s_look[3 * 2 << LOOK_BITS];
    __shared__ uint8_t s_zigzag[80];
    __shared__ int s_tot[4];
    __shared__ int s_flag[2];
    const int C = CLUSTER ? static_cast<int>(cl_size()) : 1;           // CTAs that share this segment
    const int rank = CLUSTER ? static_cast<int>(cl_rank()) : 0;
    const SegmentDev sg = segments[CLUSTER ? blockIdx.x / C : blockIdx.x];
    const int t = threadIdx.x, P = blockDim.x;
    const int gt = rank * P + t, PT = C * P;                           // thread index / count over the whole segment
    SegmentCtx c;
    c.im = images + sg.image;
    c.tables = tables;
    c.look = s_look;
    c.zigzag = s_zigzag;
    for (int i = t; i < c.im->ncomp * 2 * (1 << LOOK_BITS) / 2; i += P) {       // 32-bit copies, two entries each
        const int tab = i / (1 << (LOOK_BITS - 1));                               // component * 2 + (0 DC, 1 AC)
        const int w = i - tab * (1 << (LOOK_BITS - 1));
        const HuffDev* src = tables + ((tab & 1) ? c.im->ac_tab[tab >> 1] : c.im->dc_tab[tab >> 1]);
        reinterpret_cast<uint32_t*>(s_look)[i] = __ldg(reinterpret_cast<const uint32_t*>(src->look) + w);
    }
    for (int i = t; i < 80; i += P) s_zigzag[i] = c_zigzag[i];
    __syncthreads();
    c.hs = c.im->hs;
    c.luma = c.im->hs * c.im->vs;
    c.bpm = c.im->ncomp == 1 ? 1 : c.luma + 2;
    c.first_mcu = sg.first_mcu;
    c.total_blocks = sg.n_mcu * c.bpm;
    c.nbits = static_cast<uint32_t>(sg.nbits);
    c.words = reinterpret_cast<const uint32_t*>(scan + sg.begin);
    c.nwords = (c.nbits + 31u) >> 5;
    // one chunk per thread: a multiple of 32 bits, at least HUFF_MIN_CHUNK_BITS (trailing threads may get nothing)
    uint32_t chunk = (c.nbits + static_cast<uint32_t>(PT) - 1u) / static_cast<uint32_t>(PT);
    chunk = max((chunk + 31u) & ~31u, static_cast<uint32_t>(min_chunk_bits));
    const uint32_t my_begin = static_cast<uint32_t>(min(static_cast<uint64_t>(gt) * chunk, static_cast<uint64_t>(c.nbits)));
    const uint32_t my_end = min(my_begin + chunk, c.nbits);
    // stage the words this CTA's threads will read (their chunks + the few words a symbol may reach past a chunk end)
    c.stage = nullptr;
    c.stage_w0 = 0;
    c.stage_count = 0;
    if (stage_words > 0) {
        const uint64_t cta_begin = min(static_cast<uint64_t>(rank) * P * chunk, static_cast<uint64_t>(c.nbits));
        const uint64_t cta_end = min(cta_begin + static_cast<uint64_t>(P) * chunk, static_cast<uint64_t>(c.nbits));
        const uint32_t w0 = static_cast<uint32_t>(cta_begin >> 5);
        const uint32_t w1 = min(c.nwords, static_cast<uint32_t>((cta_end + 31u) >> 5) + 4u);
        const uint32_t count = min(w1 > w0 ? w1 - w0 : 0u, static_cast<uint32_t>(stage_words));
        for (uint32_t i = t; i < count; i += P) s_stream[stage_index(i)] = __byte_perm(__ldg(c.words + w0 + i), 0, 0x0123);
        c.stage = s_stream;
        c.stage_w0 = w0;
        c.stage_count = count;
        __syncthreads();
    }

    // ---- 1 + 2: find every chunk's true entry state
    DecodeState in = {my_begin, 0, 0};
    DecodeState out = in;
    int blocks = 0, bad = 0, overrun = 0;
    bool dirty = true;
    int round = 0;
    for (; round <= PT; ++round) {
        if (dirty) {
            out = in;
            blocks = 0;
            int ignore = 0;
            decode_range<false>(c, out, my_end, blocks, 0, nullptr, ignore, ignore);
        }
        s_pos[t] = out.pos;
        s_state[t] = (out.blk << 8) | out.k;
        if (CLUSTER) cl_sync(); else __syncthreads();
        DecodeState want = in;
        if (t > 0) {
            want.pos = s_pos[t - 1];
            want.blk = s_state[t - 1] >> 8;
            want.k = s_state[t - 1] & 255;
        } else if (CLUSTER && rank > 0) {                   // the chunk before this CTA's first one lives in the previous CTA
            want.pos = cl_load(&s_pos[P - 1], static_cast<uint32_t>(rank - 1));
            const int st = static_cast<int>(cl_load(&s_state[P - 1], static_cast<uint32_t>(rank - 1)));
            want.blk = st >> 8;
            want.k = st & 255;
        }
        dirty = want.pos != in.pos || want.blk != in.blk || want.k != in.k;
        in = want;
        int any = __syncthreads_or(dirty ? 1 : 0);
        if (CLUSTER) {
            // (flag slots alternate: a slot is rewritten two rounds later, after every reader has passed two barriers)
            if (t == 0) s_flag[round & 1] = any;
            cl_sync();
            any = 0;
            for (int r = 0; r < C; ++r) any |= static_cast<int>(cl_load(&s_flag[round & 1], static_cast<uint32_t>(r)));
        }
        if (!any) break;
    }

    if (debug_rounds && status && gt == 0) atomicMax(status + sg.image, (round + 1) << 8);   // measurement aid only

    // ---- 3: index of each thread's first block, then the decode that writes
    int total = 0;
    const int first_block = group_exclusive_scan<CLUSTER>(blocks, s_warp, s_tot, 0, total);
    if (total < c.total_blocks) overrun = 1;            // the data ran out before the last block was complete
    {
        DecodeState st = in;
        int n = 0;
        decode_range<true>(c, st, my_end, n, first_block, coef, bad, overrun);
    }
    if ((overrun | bad) && status) atomicOr(status + sg.image, overrun | (bad << 1));
    if (CLUSTER) cl_sync(); else __syncthreads();       // every coefficient of this segment is in place

    // ---- 4: DC differences -> DC values, per component, in decode order
    for (int comp = 0; comp < c.im->ncomp; ++comp) {
        const int per_mcu = comp == 0 ? c.luma : 1;
        const int nblk = sg.n_mcu * per_mcu;
        const int per_thread = (nblk + PT - 1) / PT;
        const int i0 = static_cast<int>(min(static_cast<int64_t>(gt) * per_thread, static_cast<int64_t>(nblk)));
        const int i1 = min(i0 + per_thread, nblk);
        auto dc_ptr = [&](int i) -> int16_t* {
            const int m = i / per_mcu, j = i - m * per_mcu;
            int cc;
            return block_ptr(c, coef, m * c.bpm + (comp == 0 ? j : c.luma + comp - 1), cc);
        };
        int sum = 0;
        for (int i = i0; i < i1; ++i) sum += *dc_ptr(i);
        int all = 0;
        int run = group_exclusive_scan<CLUSTER>(sum, s_warp, s_tot, 1 + comp, all);
        for (int i = i0; i < i1; ++i) {
            int16_t* p = dc_ptr(i);
            run += *p;
            *p = static_cast<int16_t>(run);
        }
    }
    if (CLUSTER) cl_sync();        // nobody leaves while a neighbour may still read this CTA's shared memory
}

// ------------------------------------------------------------------------------------------- J2: dequantise + IDCT
// libjpeg's accurate integer IDCT (jidctint.c: jpeg_idct_islow): 13-bit constants, columns first with 2 extra
// fraction bits kept, then rows, then +128 through the range-limit table (index taken modulo 1024, then clamped).
__device__ __forceinline__ void idct_1d(int x0, int x1, int x2, int x3, int x4, int x5, int x6, int x7, int (&o)[8]) {
    int z1 = (x2 + x6) * 4433;
    const int tmp2 = z1 + x6 * (-15137);
    const int tmp3 = z1 + x2 * 6270;
    const int tmp0 = (x0 + x4) << 13;
    const int tmp1 = (x0 - x4) << 13;
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    int t0 = x7, t1 = x5, t2 = x3, t3 = x1;
    z1 = t0 + t3;
    int z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
    const int z5 = (z3 + z4) * 9633;
    t0 *= 2446;
    t1 *= 16819;
    t2 *= 25172;
    t3 *= 12299;
    z1 *= -7373;
    z2 *= -20995;
    z3 = z3 * (-16069) + z5;
    z4 = z4 * (-3196) + z5;
    t0 += z1 + z3;
    t1 += z2 + z4;
    t2 += z2 + z3;
    t3 += z1 + z4;
    o[0] = tmp10 + t3; o[7] = tmp10 - t3;
    o[1] = tmp11 + t2; o[6] = tmp11 - t2;
    o[2] = tmp12 + t1; o[5] = tmp12 - t1;
    o[3] = tmp13 + t0; o[4] = tmp13 - t0;
}

__device__ __forceinline__ uint32_t range_limit(int v) {      // v = descaled IDCT output, before the +128
    int s = v & 1023;
    s = s < 512 ? s : s - 1024;
    return static_cast<uint32_t>(min(max(s + 128, 0), 255));
}

// One thread per 8x8 block: grid (blocks of the largest image / 128, images).
__global__ void __launch_bounds__(128)
idct_kernel(const ImageDev* __restrict__ images, const int16_t* __restrict__ coef, uint8_t* __restrict__ planes) {
    const ImageDev* __restrict__ im = images + blockIdx.y;
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    int ci = 0;
    for (; ci < im->ncomp; ++ci) {
        const int nb = im->blocks_w[ci] * im->blocks_h[ci];
        if (b < nb) break;
        b -= nb;
    }
    if (ci >= im->ncomp) return;
    const int bw = im->blocks_w[ci];
    const int by = b / bw, bx = b - by * bw;
    const uint4* __restrict__ src = reinterpret_cast<const uint4*>(coef + im->coef_off[ci] + static_cast<int64_t>(b) * 64);
    const uint4* __restrict__ q4 = reinterpret_cast<const uint4*>(im->quant[ci]);
    int x[64];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const uint4 c = src[r];
        const uint4 q = __ldg(q4 + r);
        const uint32_t cw[4] = {c.x, c.y, c.z, c.w};
        const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            x[r * 8 + 2 * j] = static_cast<int>(static_cast<int16_t>(cw[j] & 0xFFFFu)) * static_cast<int>(qw[j] & 0xFFFFu);
            x[r * 8 + 2 * j + 1] = static_cast<int>(static_cast<int16_t>(cw[j] >> 16)) * static_cast<int>(qw[j] >> 16);
        }
    }
    // pass 1: columns
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        int o[8];
        idct_1d(x[c], x[8 + c], x[16 + c], x[24 + c], x[32 + c], x[40 + c], x[48 + c], x[56 + c], o);
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r * 8 + c] = (o[r] + (1 << 10)) >> 11;
    }
    // pass 2: rows
    uint8_t* __restrict__ dst = planes + im->plane_off[ci] + static_cast<int64_t>(by * 8) * (bw * 8) + bx * 8;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int o[8];
        idct_1d(x[r * 8], x[r * 8 + 1], x[r * 8 + 2], x[r * 8 + 3], x[r * 8 + 4], x[r * 8 + 5], x[r * 8 + 6], x[r * 8 + 7], o);
        uint32_t p[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) p[c] = range_limit((o[c] + (1 << 17)) >> 18);
        uint2 w;
        w.x = p[0] | (p[1] << 8) | (p[2] << 16) | (p[3] << 24);
        w.y = p[4] | (p[5] << 8) | (p[6] << 16) | (p[7] << 24);
        *reinterpret_cast<uint2*>(dst + static_cast<int64_t>(r) * (bw * 8)) = w;
    }
}

// ------------------------------------------------------------------------------------------- J3: upsample + colour
// Fancy (triangle) chroma upsampling exactly as jdsample.c does it, then jdcolor.c's fixed-point YCbCr -> RGB.
// One thread produces four horizontally adjacent pixels: grid (ceil(W/4 * H / 256) of the largest image, images).
__device__ __forceinline__ int clamp255(int v) { return min(max(v, 0), 255); }

__device__ __forceinline__ void chroma4(const uint8_t* __restrict__ plane, int pitch, int hs, int vs, int y, int x0, int w_ds,
                                        int h_ds, int (&out)[4]) {
    if (hs == 1) {                                   // 4:4:4
        const uint8_t* row = plane + static_cast<int64_t>(y) * pitch + x0;
#pragma unroll
        for (int j = 0; j < 4; ++j) out[j] = row[j];
        return;
    }
    const int c0 = x0 >> 1;                          // x0 is a multiple of 4: chroma columns c0, c0 + 1
    const int cm = max(c0 - 1, 0), c1 = min(c0 + 1, w_ds - 1), c2 = min(c0 + 2, w_ds - 1);
    if (vs == 1) {                                   // 4:2:2, h2v1_fancy_upsample
        const uint8_t* row = plane + static_cast<int64_t>(y) * pitch;
        const int am = row[cm], a0 = row[c0], a1 = row[c1], a2 = row[c2];
        out[0] = c0 == 0 ? a0 : (3 * a0 + am + 1) >> 2;
        out[1] = c0 == w_ds - 1 ? a0 : (3 * a0 + a1 + 2) >> 2;
        out[2] = (3 * a1 + a0 + 1) >> 2;
        out[3] = c0 + 1 >= w_ds - 1 ? a1 : (3 * a1 + a2 + 2) >> 2;
        return;
    }
    // 4:2:0, h2v2_fancy_upsample: blend the nearer chroma row 3:1 with the farther one (edge rows duplicated)
    const int cy = y >> 1;
    const int oy = (y & 1) ? min(cy + 1, h_ds - 1) : max(cy - 1, 0);
    const uint8_t* near = plane + static_cast<int64_t>(cy) * pitch;
    const uint8_t* far = plane + static_cast<int64_t>(oy) * pitch;
    const int sm = 3 * near[cm] + far[cm], s0 = 3 * near[c0] + far[c0], s1 = 3 * near[c1] + far[c1],
              s2 = 3 * near[c2] + far[c2];
    out[0] = c0 == 0 ? (4 * s0 + 8) >> 4 : (3 * s0 + sm + 8) >> 4;
    out[1] = c0 == w_ds - 1 ? (4 * s0 + 7) >> 4 : (3 * s0 + s1 + 7) >> 4;
    out[2] = (3 * s1 + s0 + 8) >> 4;
    out[3] = c0 + 1 >= w_ds - 1 ? (4 * s1 + 7) >> 4 : (3 * s1 + s2 + 7) >> 4;
}

__global__ void __launch_bounds__(256)
colour_kernel(const ImageDev* __restrict__ images, const uint8_t* __restrict__ planes, uint8_t* __restrict__ slab) {
    const ImageDev* __restrict__ im = images + blockIdx.y;
    const int W = im->width, H = im->height;
    const int quads = (W + 3) >> 2;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= quads * H) return;
    const int y = t / quads, x0 = (t - y * quads) * 4;
    const int pitch_y = im->blocks_w[0] * 8;
    const uint8_t* yrow = planes + im->plane_off[0] + static_cast<int64_t>(y) * pitch_y + x0;
    int yy[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) yy[j] = yrow[j];       // the plane is padded to whole blocks: x0 + 3 is inside
    uint8_t rgb[12];
    if (im->ncomp == 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) rgb[3 * j] = rgb[3 * j + 1] = rgb[3 * j + 2] = static_cast<uint8_t>(yy[j]);
    } else {
        const int hs = im->hs, vs = im->vs;
        const int w_ds = (W + hs - 1) / hs, h_ds = (H + vs - 1) / vs;
        const int pitch_c = im->blocks_w[1] * 8;
        int cb[4], cr[4];
        chroma4(planes + im->plane_off[1], pitch_c, hs, vs, y, x0, w_ds, h_ds, cb);
        chroma4(planes + im->plane_off[2], pitch_c, hs, vs, y, x0, w_ds, h_ds, cr);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int b = cb[j] - 128, r = cr[j] - 128;
            rgb[3 * j + 0] = static_cast<uint8_t>(clamp255(yy[j] + ((91881 * r + 32768) >> 16)));
            rgb[3 * j + 1] = static_cast<uint8_t>(clamp255(yy[j] + ((-22554 * b + 32768 - 46802 * r) >> 16)));
            rgb[3 * j + 2] = static_cast<uint8_t>(clamp255(yy[j] + ((116130 * b + 32768) >> 16)));
        }
    }
    uint8_t* dst = slab + im->out_off + (static_cast<int64_t>(y) * W + x0) * 3;
    if (x0 + 4 <= W && (reinterpret_cast<uintptr_t>(dst) & 3u) == 0) {
        uint32_t* d4 = reinterpret_cast<uint32_t*>(dst);
        d4[0] = rgb[0] | (rgb[1] << 8) | (rgb[2] << 16) | (static_cast<uint32_t>(rgb[3]) << 24);
        d4[1] = rgb[4] | (rgb[5] << 8) | (rgb[6] << 16) | (static_cast<uint32_t>(rgb[7]) << 24);
        d4[2] = rgb[8] | (rgb[9] << 8) | (rgb[10] << 16) | (static_cast<uint32_t>(rgb[11]) << 24);
    } else {
        const int npix = min(4, W - x0);
        for (int j = 0; j < 3 * npix; ++j) dst[j] = rgb[j];
    }
}

// ------------------------------------------------------------------------------------------- host: marker parser
static_assert(offsetof(ImageDev, quant) % 16 == 0 && sizeof(ImageDev) % 16 == 0, "quant tables are read as uint4");
static_assert(sizeof(HuffDev) % 4 == 0 && sizeof(SegmentDev) % 8 == 0, "blob sections stay aligned");

struct HuffHost {
    bool present = false;
    uint8_t counts[16] = {};
    uint8_t vals[256] = {};
    int nvals = 0;
};

struct Parsed {
    int width = 0, height = 0, ncomp = 0;
    int ident[3] = {}, hs[3] = {}, vs[3] = {}, tq[3] = {}, td[3] = {}, ta[3] = {};
    bool qt_present[4] = {};
    uint16_t qt[4][64] = {};            // natural order
    HuffHost dc[4], ac[4];
    int restart_interval = 0;
    int64_t scan_start = 0, scan_end = 0;
    int mcux = 0, mcuy = 0;
    std::vector<int64_t> seg_begin, seg_end;   // absolute byte positions in the file
};

void why(char* dst, size_t n, const char* fmt, ...) {
    if (!dst || !n) return;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, n, fmt, ap);
    va_end(ap);
}

// Returns B2P_OK, B2P_E_UNSUPPORTED (a fine file, but not for this decoder) or B2P_E_ARG (not a decodable JPEG).
int parse_jpeg(const uint8_t* d, int64_t len, Parsed& P, char* msg, size_t msg_len) {
    if (!d || len < 4 || d[0] != 0xFF || d[1] != 0xD8) { why(msg, msg_len, "not a JPEG (no SOI marker)"); return B2P_E_ARG; }
    int64_t i = 2;
    bool seen_sof = false;
    for (;;) {
        if (i + 4 > len) { why(msg, msg_len, "truncated before the scan header"); return B2P_E_ARG; }
        if (d[i] != 0xFF) { why(msg, msg_len, "marker expected at byte %lld", (long long)i); return B2P_E_ARG; }
        while (i + 1 < len && d[i + 1] == 0xFF) ++i;
        if (i + 4 > len) { why(msg, msg_len, "truncated before the scan header"); return B2P_E_ARG; }
        const int m = d[i + 1];
        i += 2;
        if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;
        const int64_t seg_len = (d[i] << 8) | d[i + 1];
        if (seg_len < 2 || i + seg_len > len) { why(msg, msg_len, "truncated marker segment"); return B2P_E_ARG; }
        const uint8_t* s = d + i + 2;
        const int64_t n = seg_len - 2;
        if (m == 0xDB) {
            int64_t p = 0;
            while (p < n) {
                const int pq = s[p] >> 4, tq = s[p] & 15;
                ++p;
                if (pq != 0) { why(msg, msg_len, "16-bit quantisation table"); return B2P_E_UNSUPPORTED; }
                if (tq > 3 || p + 64 > n) { why(msg, msg_len, "bad DQT segment"); return B2P_E_ARG; }
                for (int k = 0; k < 64; ++k) P.qt[tq][h_zigzag[k]] = s[p + k];
                P.qt_present[tq] = true;
                p += 64;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (seen_sof) { why(msg, msg_len, "more than one frame header"); return B2P_E_UNSUPPORTED; }
            seen_sof = true;
            if (n < 6) { why(msg, msg_len, "bad SOF segment"); return B2P_E_ARG; }
            if (s[0] != 8) { why(msg, msg_len, "sample precision %d", s[0]); return B2P_E_UNSUPPORTED; }
            P.height = (s[1] << 8) | s[2];
            P.width = (s[3] << 8) | s[4];
            P.ncomp = s[5];
            if (P.ncomp != 1 && P.ncomp != 3) { why(msg, msg_len, "%d components", P.ncomp); return B2P_E_UNSUPPORTED; }
            if (n < 6 + 3 * P.ncomp) { why(msg, msg_len, "bad SOF segment"); return B2P_E_ARG; }
            for (int c = 0; c < P.ncomp; ++c) {
                P.ident[c] = s[6 + 3 * c];
                P.hs[c] = s[7 + 3 * c] >> 4;
                P.vs[c] = s[7 + 3 * c] & 15;
                P.tq[c] = s[8 + 3 * c];
                if (P.tq[c] > 3 || P.hs[c] < 1 || P.vs[c] < 1) { why(msg, msg_len, "bad component spec"); return B2P_E_ARG; }
            }
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            why(msg, msg_len, "progressive, lossless or arithmetic-coded frame (SOF%d)", m - 0xC0);
            return B2P_E_UNSUPPORTED;
        } else if (m == 0xCC) {
            why(msg, msg_len, "arithmetic coding");
            return B2P_E_UNSUPPORTED;
        } else if (m == 0xC4) {
            int64_t p = 0;
            while (p < n) {
                if (p + 17 > n) { why(msg, msg_len, "bad DHT segment"); return B2P_E_ARG; }
                const int tc = s[p] >> 4, th = s[p] & 15;
                if (tc > 1 || th > 3) { why(msg, msg_len, "bad DHT table id"); return B2P_E_ARG; }
                HuffHost& t = tc == 0 ? P.dc[th] : P.ac[th];
                int total = 0;
                for (int k = 0; k < 16; ++k) { t.counts[k] = s[p + 1 + k]; total += t.counts[k]; }
                if (total > 256 || p + 17 + total > n) { why(msg, msg_len, "bad DHT segment"); return B2P_E_ARG; }
                memset(t.vals, 0, sizeof t.vals);
                memcpy(t.vals, s + p + 17, total);
                t.nvals = total;
                t.present = true;
                p += 17 + total;
            }
        } else if (m == 0xDD) {
            if (n < 2) { why(msg, msg_len, "bad DRI segment"); return B2P_E_ARG; }
            P.restart_interval = (s[0] << 8) | s[1];
        } else if (m == 0xEE && n >= 5 && memcmp(s, "Adobe", 5) == 0) {
            why(msg, msg_len, "Adobe colour-transform marker");
            return B2P_E_UNSUPPORTED;
        } else if (m == 0xDA) {
            if (!seen_sof) { why(msg, msg_len, "scan header before frame header"); return B2P_E_ARG; }
            if (n < 1 || s[0] != P.ncomp) { why(msg, msg_len, "non-interleaved or multi-scan file"); return B2P_E_UNSUPPORTED; }
            if (n < 4 + 2 * P.ncomp) { why(msg, msg_len, "bad SOS segment"); return B2P_E_ARG; }
            for (int c = 0; c < P.ncomp; ++c) {
                if (s[1 + 2 * c] != P.ident[c]) { why(msg, msg_len, "scan component order"); return B2P_E_UNSUPPORTED; }
                P.td[c] = s[2 + 2 * c] >> 4;
                P.ta[c] = s[2 + 2 * c] & 15;
                if (P.td[c] > 3 || P.ta[c] > 3) { why(msg, msg_len, "bad table selector"); return B2P_E_ARG; }
            }
            if (s[1 + 2 * P.ncomp] != 0 || s[2 + 2 * P.ncomp] != 63 || s[3 + 2 * P.ncomp] != 0) {
                why(msg, msg_len, "spectral selection / successive approximation");
                return B2P_E_UNSUPPORTED;
            }
            P.scan_start = i + seg_len;
            break;
        }
        i += seg_len;
    }
    if (P.width < 1 || P.height < 1) { why(msg, msg_len, "empty frame"); return B2P_E_ARG; }
    // A 600-byte file may claim 65535 x 65535 pixels; the workspace for that would be ~20 GB.  Leave anything above
    // 64 Mpixel to the host decoder and its own decompression-bomb policy.
    if (static_cast<int64_t>(P.width) * P.height > (64LL << 20)) {
        why(msg, msg_len, "%d x %d pixels is above the GPU decoder's 64 Mpixel limit", P.width, P.height);
        return B2P_E_UNSUPPORTED;
    }
    for (int c = 0; c < P.ncomp; ++c) {
        if (!P.qt_present[P.tq[c]]) { why(msg, msg_len, "missing quantisation table"); return B2P_E_ARG; }
        if (!P.dc[P.td[c]].present || !P.ac[P.ta[c]].present) { why(msg, msg_len, "missing Huffman table"); return B2P_E_ARG; }
    }
    if (P.ncomp == 1) {
        P.hs[0] = P.vs[0] = 1;          // a single-component scan is never interleaved (T.81 A.2.2)
    } else {
        if (P.ident[0] == 'R' && P.ident[1] == 'G' && P.ident[2] == 'B') { why(msg, msg_len, "RGB component ids"); return B2P_E_UNSUPPORTED; }
        const bool chroma_ok = P.hs[1] == 1 && P.vs[1] == 1 && P.hs[2] == 1 && P.vs[2] == 1;
        const bool luma_ok = (P.hs[0] == 1 && P.vs[0] == 1) || (P.hs[0] == 2 && P.vs[0] == 1) || (P.hs[0] == 2 && P.vs[0] == 2);
        if (!chroma_ok || !luma_ok) { why(msg, msg_len, "sampling factors %dx%d", P.hs[0], P.vs[0]); return B2P_E_UNSUPPORTED; }
        if (P.hs[0] == 2 && (P.width + 1) / 2 <= 2) { why(msg, msg_len, "chroma narrower than 3 samples"); return B2P_E_UNSUPPORTED; }
    }
    P.mcux = (P.width + 8 * P.hs[0] - 1) / (8 * P.hs[0]);
    P.mcuy = (P.height + 8 * P.vs[0] - 1) / (8 * P.vs[0]);
    // entropy-coded data: split at RSTn, ends at the first other marker
    int64_t seg0 = P.scan_start, j = P.scan_start;
    for (;;) {
        const uint8_t* f = j < len ? static_cast<const uint8_t*>(memchr(d + j, 0xFF, static_cast<size_t>(len - j))) : nullptr;
        if (!f || f + 1 >= d + len) { why(msg, msg_len, "entropy-coded data is not terminated by a marker"); return B2P_E_ARG; }
        j = f - d;
        const int nx = d[j + 1];
        if (nx == 0x00) { j += 2; continue; }
        if (nx == 0xFF) { j += 1; continue; }
        if (nx >= 0xD0 && nx <= 0xD7) {
            P.seg_begin.push_back(seg0);
            P.seg_end.push_back(j);
            seg0 = j + 2;
            j += 2;
            continue;
        }
        break;
    }
    P.seg_begin.push_back(seg0);
    P.seg_end.push_back(j);
    P.scan_end = j;
    const int64_t total_mcu = static_cast<int64_t>(P.mcux) * P.mcuy;
    const int64_t expect = P.restart_interval > 0 ? (total_mcu + P.restart_interval - 1) / P.restart_interval : 1;
    if (static_cast<int64_t>(P.seg_begin.size()) != expect) {
        why(msg, msg_len, "%zu restart segments found, %lld expected", P.seg_begin.size(), (long long)expect);
        return B2P_E_ARG;
    }
    return B2P_OK;
}

// Canonical Huffman table -> lookahead + length-walk arrays (T.81 Annex C; jdhuff.c: jpeg_make_d_derived_tbl).
bool build_huff(const HuffHost& h, HuffDev& out) {
    memset(&out, 0, sizeof out);
    memcpy(out.vals, h.vals, 256);
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
        const int n = h.counts[l - 1];
        out.valoff[l] = k - code;
        out.maxcode[l] = n ? code + n - 1 : -1;
        if (n && code + n > (1 << l)) return false;          // over-subscribed code space
        if (l <= LOOK_BITS) {
            for (int c = 0; c < n; ++c) {
                const int first = (code + c) << (LOOK_BITS - l);
                for (int f = 0; f < (1 << (LOOK_BITS - l)); ++f)
                    out.look[first + f] = static_cast<uint16_t>((l << 8) | h.vals[k + c]);
            }
        }
        code = (code + n) << 1;
        k += n;
    }
    out.maxcode[17] = 0x7FFFFFFF;
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------- C ABI
// A few persistent host threads for the per-file work of b2p_jpeg_decode (marker walk, un-stuffing copy): with
// megapixel files that work is tens of megabytes of memchr/memcpy per call and would otherwise serialise on the
// caller's thread.  run() hands out item indices through an atomic counter; the caller takes part as well.
class HostPool {
public:
    explicit HostPool(int helpers) {
        for (int i = 0; i < helpers; ++i) threads_.emplace_back([this] { loop(); });
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    int helpers() const { return static_cast<int>(threads_.size()); }
    void run(int items, const std::function<void(int)>& fn) {
        if (items <= 0) return;
        if (threads_.empty() || items == 1) {
            for (int i = 0; i < items; ++i) fn(i);
            return;
        }
        {
            std::lock_guard<std::mutex> lk(m_);
            fn_ = &fn;
            items_ = items;
            next_.store(0);
            pending_ = static_cast<int>(threads_.size());
            ++generation_;
        }
        cv_.notify_all();
        work();
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

private:
    void work() {
        for (;;) {
            const int i = next_.fetch_add(1);
            if (i >= items_) break;
            (*fn_)(i);
        }
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return stop_ || generation_ != seen; });
                if (stop_) return;
                seen = generation_;
            }
            work();
            {
                std::lock_guard<std::mutex> lk(m_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    std::vector<std::thread> threads_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int)>* fn_ = nullptr;
    std::atomic<int> next_{0};
    int items_ = 0, pending_ = 0;
    uint64_t generation_ = 0;
    bool stop_ = false;
};

struct b2p_jpeg {
    int device = -1;
    int num_sms = 0;
    char err[512] = "";
    // Per-call descriptors + entropy-coded bytes travel in ONE host->device copy from a pinned blob.  Two blobs
    // alternate so that filling the next call's blob never waits for the previous call's copy.
    static constexpr int RING = 2;
    uint8_t* pinned[RING] = {};
    uint8_t* blob_dev[RING] = {};
    int64_t blob_cap[RING] = {};
    cudaEvent_t copied[RING] = {};
    bool used[RING] = {};
    int pos = 0;
    int16_t* coef = nullptr;      // [blocks][64] int16, zeroed per call
    int64_t coef_cap = 0;         // bytes
    uint8_t* planes = nullptr;    // component sample planes
    int64_t planes_cap = 0;
    HostPool* pool = nullptr;     // helper threads for the host side of decode()
    cudaStream_t copy_stream = nullptr;   // the blob's host->device copy runs here, under the previous call's kernels
    cudaEvent_t kernels_done[RING] = {};  // the kernels that read blob_dev[i] have finished (caller's stream)
    cudaEvent_t ev_copy[2] = {};          // around the copy (copy stream)
    cudaEvent_t ev[5] = {};       // around each kernel of the most recent call ([0] unused, kept for layout)
    bool ev_valid = false;
    int64_t calls = 0;
};

namespace {

thread_local char g_jpeg_error[512] = "";   // per thread: inspect() is called from many request threads at once

void jerr(b2p_jpeg* d, const char* fmt, ...) {
    char* dst = d ? d->err : g_jpeg_error;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
}

#define JPEG_CUDA(d, call)                                                                            \
    do {                                                                                              \
        cudaError_t e__ = (call);                                                                     \
        if (e__ != cudaSuccess) {                                                                     \
            jerr(d, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__);     \
            return B2P_E_CUDA;                                                                        \
        }                                                                                             \
    } while (0)

int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" {

int32_t b2p_jpeg_inspect(const uint8_t* file, int64_t len, b2p_jpeg_info* info) {
    Parsed P;
    const int rc = parse_jpeg(file, len, P, g_jpeg_error, sizeof g_jpeg_error);
    if (rc != B2P_OK) return rc;
    if (info) {
        info->width = P.width;
        info->height = P.height;
        info->components = P.ncomp;
        info->h_samp = P.hs[0];
        info->v_samp = P.vs[0];
        info->restart_interval = P.restart_interval;
        info->segments = static_cast<int32_t>(P.seg_begin.size());
        info->reserved = 0;
        info->scan_bytes = P.scan_end - P.scan_start;
    }
    return B2P_OK;
}

int32_t b2p_jpeg_create(int32_t device, b2p_jpeg** out) {
    if (!out) return B2P_E_ARG;
    *out = nullptr;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        jerr(nullptr, "cudaGetDeviceProperties(%d) failed: no CUDA device (this library has no CPU path)", device);
        return B2P_E_DEVICE;
    }
    if (prop.major != 10) {
        jerr(nullptr, "device %d is sm_%d%d; libb200predict is built for sm_100a only", device, prop.major, prop.minor);
        return B2P_E_DEVICE;
    }
    b2p_jpeg* d = new (std::nothrow) b2p_jpeg();
    if (!d) return B2P_E_NOMEM;
    d->device = device;
    d->num_sms = prop.multiProcessorCount;
    if (cudaSetDevice(device) != cudaSuccess) { delete d; return B2P_E_CUDA; }
    for (int i = 0; i < b2p_jpeg::RING; ++i) {
        if (cudaEventCreateWithFlags(&d->copied[i], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&d->kernels_done[i], cudaEventDisableTiming) != cudaSuccess) {
            jerr(nullptr, "cudaEventCreate failed");
            delete d;
            return B2P_E_CUDA;
        }
    }
    if (cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&d->ev_copy[0]) != cudaSuccess || cudaEventCreate(&d->ev_copy[1]) != cudaSuccess) {
        jerr(nullptr, "could not create the copy stream");
        delete d;
        return B2P_E_CUDA;
    }
    for (auto& e : d->ev) {
        if (cudaEventCreate(&e) != cudaSuccess) {
            jerr(nullptr, "cudaEventCreate failed");
            delete d;
            return B2P_E_CUDA;
        }
    }
    {
        // helper threads: B2P_JPEG_THREADS (total, including the caller), default min(8, half the hardware threads)
        const char* e = getenv("B2P_JPEG_THREADS");
        int total = e ? atoi(e) : static_cast<int>(std::thread::hardware_concurrency() / 2);
        if (total > 8 && !e) total = 8;
        if (total < 1) total = 1;
        if (total > 64) total = 64;
        d->pool = new (std::nothrow) HostPool(total - 1);
    }
    *out = d;
    return B2P_OK;
}

int32_t b2p_jpeg_stats_get(b2p_jpeg* d, b2p_jpeg_stats* out) {
    if (!d || !out) return B2P_E_ARG;
    memset(out, 0, sizeof *out);
    out->calls = d->calls;
    if (!d->ev_valid) return B2P_OK;
    if (cudaEventSynchronize(d->ev[4]) != cudaSuccess) return B2P_E_CUDA;
    cudaEventElapsedTime(&out->ms_copy, d->ev_copy[0], d->ev_copy[1]);
    cudaEventElapsedTime(&out->ms_huffman, d->ev[1], d->ev[2]);
    cudaEventElapsedTime(&out->ms_idct, d->ev[2], d->ev[3]);
    cudaEventElapsedTime(&out->ms_colour, d->ev[3], d->ev[4]);
    return B2P_OK;
}

int32_t b2p_jpeg_destroy(b2p_jpeg* d) {
    if (!d) return B2P_OK;
    cudaSetDevice(d->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < b2p_jpeg::RING; ++i) {
        if (d->pinned[i]) cudaFreeHost(d->pinned[i]);
        if (d->blob_dev[i]) cudaFree(d->blob_dev[i]);
        if (d->copied[i]) cudaEventDestroy(d->copied[i]);
    }
    for (auto& e : d->ev)
        if (e) cudaEventDestroy(e);
    for (auto& e : d->ev_copy)
        if (e) cudaEventDestroy(e);
    for (auto& e : d->kernels_done)
        if (e) cudaEventDestroy(e);
    if (d->copy_stream) cudaStreamDestroy(d->copy_stream);
    if (d->coef) cudaFree(d->coef);
    if (d->planes) cudaFree(d->planes);
    delete d->pool;
    delete d;
    return B2P_OK;
}

const char* b2p_jpeg_last_error(const b2p_jpeg* d) { return d ? d->err : g_jpeg_error; }

int32_t b2p_jpeg_decode(b2p_jpeg* d, const uint8_t* const* files, const int64_t* lens, int32_t n, uint8_t* slab,
                        int64_t slab_bytes, const int64_t* offsets, int32_t* status, void* stream_) {
    if (!d) return B2P_E_ARG;
    if (!files || !lens || !slab || !offsets || n < 0) { jerr(d, "b2p_jpeg_decode: null argument"); return B2P_E_ARG; }
    if (n == 0) return B2P_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    JPEG_CUDA(d, cudaSetDevice(d->device));

    // ---- parse every header on this thread, de-duplicate the Huffman tables
    std::vector<Parsed> P(static_cast<size_t>(n));
    std::vector<HuffHost> uniq;
    auto table_id = [&](const HuffHost& h) -> int {
        for (size_t i = 0; i < uniq.size(); ++i)
            if (uniq[i].nvals == h.nvals && memcmp(uniq[i].counts, h.counts, 16) == 0 && memcmp(uniq[i].vals, h.vals, h.nvals) == 0)
                return static_cast<int>(i);
        uniq.push_back(h);
        return static_cast<int>(uniq.size() - 1);
    };
    std::vector<ImageDev> images(static_cast<size_t>(n));
    int64_t n_segments = 0, scan_total = 0, coef_elems = 0, plane_bytes = 0, block_total = 0;
    int max_blocks = 0;
    int64_t max_quads = 0;
    {
        std::vector<int> rcs(static_cast<size_t>(n), B2P_OK);
        std::vector<std::array<char, 200>> msgs(static_cast<size_t>(n));
        auto parse_one = [&](int i) {
            msgs[i][0] = 0;
            rcs[i] = parse_jpeg(files[i], lens[i], P[i], msgs[i].data(), msgs[i].size());
        };
        if (d->pool) d->pool->run(n, parse_one); else for (int i = 0; i < n; ++i) parse_one(i);
        for (int i = 0; i < n; ++i)
            if (rcs[i] != B2P_OK) { jerr(d, "file %d: %s", i, msgs[i].data()); return rcs[i]; }
    }
    for (int i = 0; i < n; ++i) {
        const Parsed& p = P[i];
        ImageDev& im = images[i];
        memset(&im, 0, sizeof im);
        im.width = p.width; im.height = p.height; im.ncomp = p.ncomp;
        im.hs = p.hs[0]; im.vs = p.vs[0]; im.mcux = p.mcux; im.mcuy = p.mcuy;
        im.block_base = block_total;
        int blocks = 0;
        for (int c = 0; c < p.ncomp; ++c) {
            im.dc_tab[c] = table_id(p.dc[p.td[c]]);
            im.ac_tab[c] = table_id(p.ac[p.ta[c]]);
            im.blocks_w[c] = p.mcux * (c == 0 ? p.hs[0] : 1);
            im.blocks_h[c] = p.mcuy * (c == 0 ? p.vs[0] : 1);
            const int64_t nb = static_cast<int64_t>(im.blocks_w[c]) * im.blocks_h[c];
            im.coef_off[c] = coef_elems;
            coef_elems += nb * 64;
            im.plane_off[c] = plane_bytes;
            plane_bytes += align_up(nb * 64, 256);
            blocks += static_cast<int>(nb);
            memcpy(im.quant[c], p.qt[p.tq[c]], 128);
        }
        block_total += blocks;
        if (blocks > max_blocks) max_blocks = blocks;
        const int64_t quads = static_cast<int64_t>((p.width + 3) / 4) * p.height;
        if (quads > max_quads) max_quads = quads;
        im.out_off = offsets[i];
        if (offsets[i] < 0 || offsets[i] + static_cast<int64_t>(p.width) * p.height * 3 > slab_bytes) {
            jerr(d, "file %d: %dx%d image at offset %lld does not fit the %lld-byte slab", i, p.width, p.height,
                 (long long)offsets[i], (long long)slab_bytes);
            return B2P_E_ARG;
        }
        n_segments += static_cast<int64_t>(p.seg_begin.size());
        scan_total += p.scan_end - p.scan_start;
    }
    if (n_segments > 0x7FFFFFFF / 64 || max_quads > 0x7FFFFFFFLL) { jerr(d, "batch too large"); return B2P_E_LIMIT; }
    std::vector<HuffDev> tables(uniq.size());
    for (size_t t = 0; t < uniq.size(); ++t) {
        if (!build_huff(uniq[t], tables[t])) { jerr(d, "a Huffman table over-subscribes its code space"); return B2P_E_ARG; }
    }

    // ---- blob layout: images | segments | tables | un-stuffed scan bytes (each segment 4-byte aligned)
    const int64_t off_img = 0;
    const int64_t off_seg = align_up(off_img + static_cast<int64_t>(n) * sizeof(ImageDev), 256);
    const int64_t off_tab = align_up(off_seg + n_segments * static_cast<int64_t>(sizeof(SegmentDev)), 256);
    const int64_t off_scan = align_up(off_tab + static_cast<int64_t>(tables.size() * sizeof(HuffDev)), 256);
    const int64_t blob_bytes = align_up(off_scan + scan_total + 8 * n_segments, 256);
    const int slot = d->pos;
    d->pos = (d->pos + 1) % b2p_jpeg::RING;
    if (d->used[slot]) JPEG_CUDA(d, cudaEventSynchronize(d->copied[slot]));     // the copy out of this blob finished
    if (blob_bytes > d->blob_cap[slot]) {
        JPEG_CUDA(d, cudaStreamSynchronize(stream));
        JPEG_CUDA(d, cudaStreamSynchronize(d->copy_stream));
        if (d->pinned[slot]) cudaFreeHost(d->pinned[slot]);
        if (d->blob_dev[slot]) cudaFree(d->blob_dev[slot]);
        d->pinned[slot] = nullptr; d->blob_dev[slot] = nullptr; d->blob_cap[slot] = 0;
        const int64_t cap = blob_bytes + blob_bytes / 2;
        if (cudaMallocHost(reinterpret_cast<void**>(&d->pinned[slot]), static_cast<size_t>(cap)) != cudaSuccess ||
            cudaMalloc(reinterpret_cast<void**>(&d->blob_dev[slot]), static_cast<size_t>(cap)) != cudaSuccess) {
            jerr(d, "could not allocate %lld bytes of staging", (long long)cap);
            return B2P_E_NOMEM;
        }
        d->blob_cap[slot] = cap;
    }
    const int64_t coef_bytes = coef_elems * 2;
    if (coef_bytes > d->coef_cap || plane_bytes > d->planes_cap) {
        JPEG_CUDA(d, cudaStreamSynchronize(stream));
        if (coef_bytes > d->coef_cap) {
            if (d->coef) cudaFree(d->coef);
            d->coef = nullptr; d->coef_cap = 0;
            const int64_t cap = coef_bytes + coef_bytes / 2;
            if (cudaMalloc(reinterpret_cast<void**>(&d->coef), static_cast<size_t>(cap)) != cudaSuccess) {
                jerr(d, "could not allocate %lld bytes for coefficients", (long long)cap);
                return B2P_E_NOMEM;
            }
            d->coef_cap = cap;
        }
        if (plane_bytes > d->planes_cap) {
            if (d->planes) cudaFree(d->planes);
            d->planes = nullptr; d->planes_cap = 0;
            const int64_t cap = plane_bytes + plane_bytes / 2;
            if (cudaMalloc(reinterpret_cast<void**>(&d->planes), static_cast<size_t>(cap)) != cudaSuccess) {
                jerr(d, "could not allocate %lld bytes for sample planes", (long long)cap);
                return B2P_E_NOMEM;
            }
            d->planes_cap = cap;
        }
    }

    // ---- fill the pinned blob; the entropy-coded bytes lose their stuffing (FF 00 -> FF) on the way
    uint8_t* blob = d->pinned[slot];
    memcpy(blob + off_img, images.data(), static_cast<size_t>(n) * sizeof(ImageDev));
    SegmentDev* segs = reinterpret_cast<SegmentDev*>(blob + off_seg);
    // every segment gets a slot sized for its raw (still stuffed) length, so the files can be copied in parallel
    std::vector<int64_t> first_seg(static_cast<size_t>(n) + 1, 0), first_byte(static_cast<size_t>(n) + 1, 0);
    for (int i = 0; i < n; ++i) {
        int64_t bytes = 0;
        for (size_t sg = 0; sg < P[i].seg_begin.size(); ++sg) bytes += align_up(P[i].seg_end[sg] - P[i].seg_begin[sg], 4) + 4;
        first_seg[i + 1] = first_seg[i] + static_cast<int64_t>(P[i].seg_begin.size());
        first_byte[i + 1] = first_byte[i] + bytes;
    }
    std::vector<int64_t> image_max_bits(static_cast<size_t>(n), 0);
    auto fill_one = [&](int i) {
        const Parsed& p = P[i];
        const int64_t total_mcu = static_cast<int64_t>(p.mcux) * p.mcuy;
        const int64_t ri = p.restart_interval > 0 ? p.restart_interval : total_mcu;
        int64_t scan_pos = first_byte[i], most = 0;
        for (size_t sg = 0; sg < p.seg_begin.size(); ++sg) {
            SegmentDev& g = segs[first_seg[i] + static_cast<int64_t>(sg)];
            g.image = i;
            g.first_mcu = static_cast<int32_t>(static_cast<int64_t>(sg) * ri);
            const int64_t left = total_mcu - g.first_mcu;
            g.n_mcu = static_cast<int32_t>(left < ri ? left : ri);
            g.pad = 0;
            g.begin = scan_pos;
            uint8_t* dst = blob + off_scan + scan_pos;
            const uint8_t* src = files[i] + p.seg_begin[sg];
            const uint8_t* end = files[i] + p.seg_end[sg];
            const int64_t slot_bytes = align_up(end - src, 4) + 4;
            int64_t w = 0;
            while (src < end) {
                const uint8_t* f = static_cast<const uint8_t*>(memchr(src, 0xFF, static_cast<size_t>(end - src)));
                const int64_t run = (f ? f + 1 : end) - src;          // up to and including the FF
                memcpy(dst + w, src, static_cast<size_t>(run));
                w += run;
                src += run;
                if (f && src < end && *src == 0x00) ++src;             // the stuffed zero
            }
            g.nbits = w * 8;
            if (g.nbits > most) most = g.nbits;
            memset(dst + w, 0, static_cast<size_t>(slot_bytes - w));
            scan_pos += slot_bytes;
        }
        image_max_bits[i] = most;
    };
    if (d->pool) d->pool->run(n, fill_one); else for (int i = 0; i < n; ++i) fill_one(i);
    int64_t max_seg_bits = 0;
    for (int i = 0; i < n; ++i) if (image_max_bits[i] > max_seg_bits) max_seg_bits = image_max_bits[i];
    if (max_seg_bits >= 0xFFFFFF00LL) { jerr(d, "an entropy-coded segment of 512 MiB or more"); return B2P_E_LIMIT; }
    memcpy(blob + off_tab, tables.data(), tables.size() * sizeof(HuffDev));

    // ---- device work.  The copy goes to a private stream so that it runs under the kernels of the previous call
    // (two blobs alternate); it only has to wait for the kernels that last read THIS device blob.  Everything else is
    // on the caller's stream, which waits for the copy.
    uint8_t* dev = d->blob_dev[slot];
    if (d->used[slot]) JPEG_CUDA(d, cudaStreamWaitEvent(d->copy_stream, d->kernels_done[slot], 0));
    JPEG_CUDA(d, cudaEventRecord(d->ev_copy[0], d->copy_stream));
    JPEG_CUDA(d, cudaMemcpyAsync(dev, blob, static_cast<size_t>(blob_bytes), cudaMemcpyHostToDevice, d->copy_stream));
    JPEG_CUDA(d, cudaEventRecord(d->ev_copy[1], d->copy_stream));
    JPEG_CUDA(d, cudaEventRecord(d->copied[slot], d->copy_stream));
    d->used[slot] = true;
    JPEG_CUDA(d, cudaMemsetAsync(d->coef, 0, static_cast<size_t>(coef_bytes), stream));
    if (status) JPEG_CUDA(d, cudaMemsetAsync(status, 0, static_cast<size_t>(n) * sizeof(int32_t), stream));
    JPEG_CUDA(d, cudaStreamWaitEvent(stream, d->copied[slot], 0));
    JPEG_CUDA(d, cudaEventRecord(d->ev[1], stream));
    const ImageDev* images_dev = reinterpret_cast<const ImageDev*>(dev + off_img);
    const SegmentDev* segs_dev = reinterpret_cast<const SegmentDev*>(dev + off_seg);
    const HuffDev* tabs_dev = reinterpret_cast<const HuffDev*>(dev + off_tab);
    {
        // one CTA per restart segment, one chunk of at least `min_chunk` bits per thread
        static const int min_chunk = [] {
            const char* e = getenv("B2P_JPEG_CHUNK_BITS");
            const int v = e ? atoi(e) : HUFF_MIN_CHUNK_BITS;
            return v >= 64 ? (v + 31) / 32 * 32 : HUFF_MIN_CHUNK_BITS;
        }();
        int64_t want = (max_seg_bits + min_chunk - 1) / min_chunk;
        want = (want + 31) / 32 * 32;
        static const int debug_rounds = getenv("B2P_JPEG_DEBUG_ROUNDS") != nullptr;   // read once
        // segments that want more than one CTA's worth of threads are spread over a cluster of CTAs (other SMs):
        // B2P_JPEG_CLUSTER caps the cluster size (1 = never; default 4: measured on 32 x 1080p files the Huffman
        // kernel takes 2.43 / 1.39 / 1.07 / 3.1 ms with 1 / 2 / 4 / 8 CTAs per file — 1024-thread CTAs are one per SM,
        // so 8-CTA clusters no longer fit the GPU in one wave)
        static const int max_cluster = [] {
            const char* e = getenv("B2P_JPEG_CLUSTER");
            const int v = e ? atoi(e) : 4;
            return v >= 8 ? 8 : (v >= 4 ? 4 : (v >= 2 ? 2 : 1));
        }();
        int cluster = 1;
        while (cluster < max_cluster && want > static_cast<int64_t>(cluster) * HUFF_MAX_THREADS) cluster *= 2;
        const int threads = cluster > 1 ? HUFF_MAX_THREADS
                                        : static_cast<int>(want < 32 ? 32 : (want > HUFF_MAX_THREADS ? HUFF_MAX_THREADS : want));
        // shared-memory staging of the stream: as many words as the largest CTA share needs, capped by what fits
        static const int stage_on = [] { const char* e = getenv("B2P_JPEG_STAGE"); return e ? atoi(e) : 1; }();
        int stage_words = 0;
        size_t stage_bytes = 0;
        if (stage_on) {
            const int64_t pt = static_cast<int64_t>(cluster) * threads;
            int64_t chunk = (max_seg_bits + pt - 1) / pt;
            chunk = (chunk + 31) / 32 * 32;
            if (chunk < min_chunk) chunk = min_chunk;
            int64_t words = static_cast<int64_t>(threads) * chunk / 32 + 8;
            const int64_t cap = (188 * 1024 / 4) * 16 / 17 - 32;          // dynamic shared memory left beside the tables
            if (words > cap) words = cap;
            stage_words = static_cast<int>(words);
            stage_bytes = static_cast<size_t>(words + words / 16 + 16) * 4;
            static std::atomic<bool> attr_set[64][2];
            int dev_id = 0;
            cudaGetDevice(&dev_id);
            const int which = cluster > 1 ? 1 : 0;
            if (dev_id >= 0 && dev_id < 64 && !attr_set[dev_id][which].load(std::memory_order_acquire)) {
                if (which) JPEG_CUDA(d, cudaFuncSetAttribute(huffman_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 190 * 1024));
                else JPEG_CUDA(d, cudaFuncSetAttribute(huffman_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 190 * 1024));
                attr_set[dev_id][which].store(true, std::memory_order_release);
            }
        }
        if (cluster == 1) {
            huffman_kernel<false><<<static_cast<unsigned>(n_segments), threads, stage_bytes, stream>>>(
                dev + off_scan, images_dev, segs_dev, tabs_dev, d->coef, status, min_chunk, debug_rounds, stage_words);
            JPEG_CUDA(d, cudaGetLastError());
        } else {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(static_cast<unsigned>(n_segments * cluster));
            cfg.blockDim = dim3(HUFF_MAX_THREADS);
            cfg.dynamicSmemBytes = stage_bytes;
            cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = static_cast<unsigned>(cluster);
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            const uint8_t* scan_dev = dev + off_scan;
            int16_t* coef_dev = d->coef;
            JPEG_CUDA(d, cudaLaunchKernelEx(&cfg, huffman_kernel<true>, scan_dev, images_dev, segs_dev, tabs_dev, coef_dev,
                                            status, min_chunk, debug_rounds, stage_words));
        }
        JPEG_CUDA(d, cudaEventRecord(d->ev[2], stream));
    }
    {
        dim3 grid(static_cast<unsigned>((max_blocks + 127) / 128), static_cast<unsigned>(n));
        idct_kernel<<<grid, 128, 0, stream>>>(images_dev, d->coef, d->planes);
        JPEG_CUDA(d, cudaGetLastError());
        JPEG_CUDA(d, cudaEventRecord(d->ev[3], stream));
    }
    {
        dim3 grid(static_cast<unsigned>((max_quads + 255) / 256), static_cast<unsigned>(n));
        colour_kernel<<<grid, 256, 0, stream>>>(images_dev, d->planes, slab);
        JPEG_CUDA(d, cudaGetLastError());
        JPEG_CUDA(d, cudaEventRecord(d->ev[4], stream));
        JPEG_CUDA(d, cudaEventRecord(d->kernels_done[slot], stream));
    }
    d->ev_valid = true;
    ++d->calls;
    return B2P_OK;
}

}  // extern "C"
