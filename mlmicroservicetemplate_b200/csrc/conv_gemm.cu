// K2 core: persistent, warp-specialised implicit-GEMM convolution / linear layer on the
// 5th-generation tensor cores.
//
//   D[M, N] = A[M, K] * W[N, K]^T  (+ bias[N]) (+ residual[M, N]) (ReLU)
//
//   M = output pixels (n * out_h * out_w, NHWC order),  N = output channels,
//   K = kh * kw * in_c (tap-major, channel fastest — the packed weight layout).
//
// A tiles come straight from the NHWC activation tensor: a 2-D tiled TMA box for 1x1/stride-1
// convs and linear layers, TMA *im2col* mode for everything else (3x3, strided 1x1, the
// space-to-depth stem), so no im2col matrix is ever materialised.  W tiles are 2-D tiled TMA.
// Both land in shared memory in the K-major swizzled layout tcgen05.mma consumes; the
// accumulator lives in TMEM (two stages, so the epilogue of tile i overlaps the MMAs of tile
// i+1), and two groups of four epilogue warps drain it in alternating 64-column chunks with tcgen05.ld, add the
// folded-BN bias and the residual (TMA-loaded 128 x 64 tiles), apply ReLU / GELU and hand bf16 tiles to TMA stores
// through swizzled staging buffers (LINEAR: fp32 logits stored directly).
//
// Warp roles (352 threads):  warp 0 = TMA producer (one elected lane), warp 1 = TMEM owner + MMA issuer (one
// elected lane), warp 2 = residual loader, warps 3..10 = two epilogue groups (TMEM lane quadrant = warp_idx % 4).
//
// Variants: a second A source whose K blocks follow the main operand's (folded projection shortcut); CTA pairs
// (CTA2: cta_group::2, 256-row MMAs, each CTA loads half of the weight tile); optional tail splitting (the tiles of
// a partly filled last wave handed out as 64-column pieces); forward / reverse tile order for L2 reuse.
//
// Oracle-side equivalent: nn.Conv2d + BatchNorm2d(eval) + ReLU as wired in
// SP/torchvision/models/resnet.py:89-105,143-163 (BN folded on the host, see program.py: fold_bn).
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {

namespace {

constexpr int BLOCK_M = 128;
constexpr int UMMA_K = 16;           // bf16: 32 bytes of K per tcgen05.mma
constexpr int NUM_THREADS = 352;     // 11 warps: TMA, MMA, residual loader, 2 epilogue groups x 4 warps
constexpr int EPI_THREADS = 128;     // per epilogue group
constexpr int ACC_STAGES = 2;
constexpr int SMEM_LIMIT = 227 * 1024;
constexpr int OUT_CHUNK_BYTES = BLOCK_M * 64 * 2;   // one 128 x 64 bf16 staging tile (128B swizzle)
constexpr int OUT_BUFS = 2;
constexpr int RES_STAGES = 4;          // residual kernels: four 16 KB tiles that are residual input AND output staging, two per group

struct GemmArgs {
    int m_rows;       // valid output rows
    int m_tiles;
    int n_tiles;
    int n_real;       // valid output columns
    int taps_h, taps_w, k_chunks;
    int k2_chunks;    // extra 64-wide K blocks that come from the second source (1x1 conv, stride2, no padding)
    int stride2;
    int out_h, out_w;
    int stride, pad_lo;
    int ldc;
    int relu;
    int reverse;      // walk the tiles from the last to the first (see GemmPlan::reverse)
    int split_tiles;  // the last `split_tiles` tiles of the walk are cut into `split_ways` 64-column pieces, so that
    int split_ways;   // a nearly empty last wave costs a fraction of a tile time instead of a whole one (0 = off)
    const float* bias;
    void* out;        // only used by the fp32 (linear) epilogue
    uint64_t pol_a, pol_b, pol_res, pol_out;   // L2 cache policies of the four TMA streams (ptx.cuh: L2_EVICT_*)
};

// GELU(x) = 0.5 x (1 + erf(x / sqrt 2)) with erf from Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, far
// below the bf16 rounding of the result): one reciprocal, one exp2, five FMAs instead of libdevice's erff.
__device__ __forceinline__ float gelu_erf(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));      // MUFU.RCP
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * z * z));   // MUFU.EX2
    float p = fmaf(1.061405429f, t, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    const float erf_abs = fmaf(-p * t, e, 1.0f);          // erf(|x| / sqrt 2)
    const float hx = 0.5f * x;
    return fmaf(fabsf(hx), erf_abs, hx);                  // 0.5 x (1 + sign(x) erf_abs)
}

template <int BLOCK_N, int BLOCK_K, bool OUT_F32, bool HAS_RES, bool CTA2>
struct Cfg {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_ROWS = CTA2 ? BLOCK_N / 2 : BLOCK_N;   // a CTA pair splits the weight rows
    static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    // without a residual: one staging tile per epilogue group.  With one: the residual tile is loaded INTO the staging
    // tile, updated in place and stored from there, which pays for two tiles per group in the same 64 KB, so that
    // the residual of a group's next chunk is always in flight while it works on the current one
    static constexpr int OUT_BYTES = (OUT_F32 || HAS_RES) ? 0 : OUT_BUFS * OUT_CHUNK_BYTES;
    static constexpr int RES_BYTES = HAS_RES ? RES_STAGES * OUT_CHUNK_BYTES : 0;
    static constexpr int BAR_BYTES = 256;
    static constexpr int FIXED = 1024 /*alignment slack*/ + BAR_BYTES + OUT_BYTES + RES_BYTES;
    static constexpr int STAGES_RAW = (SMEM_LIMIT - FIXED) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SWIZZLE_BYTES = BLOCK_K * 2;          // 128 or 32
    static constexpr int TMEM_COLS = ACC_STAGES * BLOCK_N;     // 128 / 256 / 512
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + FIXED;
    static_assert(BLOCK_K == 64 || BLOCK_K == 16, "BLOCK_K must give a 128B or 32B swizzle row");
    static_assert(TMEM_COLS == 128 || TMEM_COLS == 256 || TMEM_COLS == 512, "TMEM power of two");
    static_assert(A_BYTES % 1024 == 0 && B_BYTES % 1024 == 0, "stage bases stay 1024B aligned");
    static_assert(STAGES >= 3, "pipeline too shallow");
    static_assert((2 * STAGES + 2 * ACC_STAGES + 2 * RES_STAGES) * 8 + 8 <= BAR_BYTES, "barrier area");
};

template <int BLOCK_N, int BLOCK_K, bool IM2COL, bool OUT_F32, bool HAS_RES, bool CTA2>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                 const __grid_constant__ CUtensorMap tmap_a2, const __grid_constant__ CUtensorMap tmap_b64,
                 const GemmArgs args) {
    using C = Cfg<BLOCK_N, BLOCK_K, OUT_F32, HAS_RES, CTA2>;
    constexpr int STAGES = C::STAGES;
    static_assert(!(OUT_F32 && HAS_RES), "the fp32 epilogue has no residual input");
    static_assert(!(CTA2 && (OUT_F32 || BLOCK_K != 64 || BLOCK_N < 128)), "CTA pairs: bf16 output, 128B rows, N >= 128");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem_a + STAGES * C::A_BYTES;
    uint8_t* smem_out = smem_b + STAGES * C::B_BYTES;          // OUT_BUFS x 16 KB, 1024B aligned
    uint8_t* smem_res = smem_out + C::OUT_BYTES;               // RES_STAGES x 16 KB
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_res + C::RES_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + ACC_STAGES;
    uint64_t* res_full_bar = tmem_empty_bar + ACC_STAGES;
    uint64_t* res_empty_bar = res_full_bar + RES_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_empty_bar + RES_STAGES);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    // CTA2: the two CTAs of a cluster own the two 128-row halves of one 256-row tile; they walk the same
    // sequence of (row pair, column block) tiles.  An odd number of row tiles leaves one phantom half whose
    // loads are zero-filled and whose stores are clipped by the tensor maps.
    const uint32_t rank = CTA2 ? cluster_ctarank() : 0u;
    const int total_tiles = (CTA2 ? (args.m_tiles + 1) / 2 : args.m_tiles) * args.n_tiles;
    const int first_tile = CTA2 ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
    const int tile_step = CTA2 ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
    const int num_kb = args.taps_h * args.taps_w * args.k_chunks + args.k2_chunks;
    // Work items: the first `full_tiles` are whole 128 x BLOCK_N tiles; each of the remaining tiles is handed out
    // as `split_ways` separate 64-column pieces (sub >= 0), which other CTAs pick up in the same last wave.
    const int full_tiles = total_tiles - args.split_tiles;
    const int total_items = full_tiles + args.split_tiles * args.split_ways;
    auto tile_coords = [&](int item, int& m_blk, int& n_blk, int& sub) {
        int t = item;
        sub = -1;
        if (item >= full_tiles) {
            const int u = item - full_tiles;
            const int q = u / args.split_ways;
            t = full_tiles + q;
            sub = u - q * args.split_ways;
        }
        if (args.reverse) t = total_tiles - 1 - t;
        const int mp = t / args.n_tiles;
        n_blk = t - mp * args.n_tiles;
        m_blk = CTA2 ? 2 * mp + static_cast<int>(rank) : mp;
    };
    constexpr uint32_t SUB_B_BYTES = 64 * BLOCK_K * 2;          // weight rows of one 64-column piece

    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_a);
            tma_prefetch_desc(&tmap_b);
            if (!OUT_F32) tma_prefetch_desc(&tmap_out);
            if (HAS_RES) tma_prefetch_desc(&tmap_res);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int i = 0; i < STAGES; ++i) {
                mbar_init(&full_bar[i], 1);
                mbar_init(&empty_bar[i], 1);
            }
            for (int i = 0; i < ACC_STAGES; ++i) {
                mbar_init(&tmem_full_bar[i], 1);
                mbar_init(&tmem_empty_bar[i], CTA2 ? 16 : 8);   // one arrive per epilogue warp (of both CTAs)
            }
            for (int i = 0; i < RES_STAGES; ++i) {
                mbar_init(&res_full_bar[i], 1);
                mbar_init(&res_empty_bar[i], 1);      // the group's store leader, once the TMA store has read the tile
            }
            fence_mbar_init();
        }
        __syncwarp();
        if (CTA2) {
            tmem_alloc_pair(tmem_slot, C::TMEM_COLS);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(tmem_slot, C::TMEM_COLS);
            tmem_relinquish();
        }
    }
    pdl_launch_dependents();      // the next kernel may start its own prologue
    tc_fence_before();
    if (CTA2) cluster_sync_all(); else __syncthreads();     // the peer signals this CTA's barriers: both must be initialised
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                   // predecessors done: activations readable, output buffers writable

    if (warp_idx == 0) {
        // ===================================================================== TMA producer
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            const int pq = args.out_h * args.out_w;
            // CTA2: both producers credit the LEADER's full barrier; only the leader arms it (for both halves)
            const uint32_t full0 = CTA2 ? mapa_shared(smem_u32(&full_bar[0]), 0) : 0u;
            const bool arms = !CTA2 || rank == 0;
            constexpr uint32_t TX_BYTES = (CTA2 ? 2u : 1u) * C::STAGE_BYTES;
            constexpr uint32_t TX_BYTES_SUB = C::A_BYTES + SUB_B_BYTES;
            auto load_a = [&](const CUtensorMap* map, bool im2col, int stg, int c, int w, int h, int n, int s_, int r_) {
                uint8_t* dst = smem_a + stg * C::A_BYTES;
                if (CTA2) {
                    const uint32_t bar = full0 + static_cast<uint32_t>(stg * 8);
                    if (im2col) tma_load_im2col_4d_pair_hint(map, bar, dst, c, w, h, n, static_cast<uint16_t>(s_), static_cast<uint16_t>(r_), args.pol_a);
                    else tma_load_2d_pair_hint(map, bar, dst, c, w, args.pol_a);
                } else {
                    if (im2col) tma_load_im2col_4d_hint(map, &full_bar[stg], dst, c, w, h, n, static_cast<uint16_t>(s_), static_cast<uint16_t>(r_), args.pol_a);
                    else tma_load_2d_hint(map, &full_bar[stg], dst, c, w, args.pol_a);
                }
            };
            auto load_b = [&](int stg, int kcol, int nrow, bool piece) {
                uint8_t* dst = smem_b + stg * C::B_BYTES;
                if (CTA2) tma_load_2d_pair_hint(&tmap_b, full0 + static_cast<uint32_t>(stg * 8), dst, kcol, nrow, args.pol_b);
                else tma_load_2d_hint(piece ? &tmap_b64 : &tmap_b, &full_bar[stg], dst, kcol, nrow, args.pol_b);
            };
            for (int tile = first_tile; tile < total_items; tile += tile_step) {
                int m_blk, n_blk, sub;
                tile_coords(tile, m_blk, n_blk, sub);
                const bool piece = sub >= 0;
                const uint32_t tx = piece ? TX_BYTES_SUB : TX_BYTES;
                const int m0 = m_blk * BLOCK_M;
                const int nrow = n_blk * BLOCK_N + (piece ? sub * 64 : static_cast<int>(rank) * C::B_ROWS);
                int img = 0, base_w = 0, base_h = 0, base_w2 = 0, base_h2 = 0;
                if (IM2COL || args.k2_chunks > 0) {
                    img = m0 / pq;
                    const int rem = m0 - img * pq;
                    const int op = rem / args.out_w;
                    const int oq = rem - op * args.out_w;
                    base_w = oq * args.stride - args.pad_lo;
                    base_h = op * args.stride - args.pad_lo;
                    base_w2 = oq * args.stride2;
                    base_h2 = op * args.stride2;
                }
                int kcol = 0;
                for (int r = 0; r < args.taps_h; ++r) {
                    for (int s = 0; s < args.taps_w; ++s) {
                        for (int cc = 0; cc < args.k_chunks; ++cc, kcol += BLOCK_K) {
                            mbar_wait(&empty_bar[stage], phase ^ 1u);
                            if (arms) mbar_arrive_expect_tx(&full_bar[stage], tx);
                            if (IM2COL) load_a(&tmap_a, true, stage, cc * BLOCK_K, base_w, base_h, img, s, r);
                            else load_a(&tmap_a, false, stage, kcol, m0, 0, 0, 0, 0);
                            load_b(stage, kcol, nrow, piece);
                            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                        }
                    }
                }
                // second source: the projection shortcut's input pixels, K continues after the main operand
                for (int cc = 0; cc < args.k2_chunks; ++cc, kcol += BLOCK_K) {
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    if (arms) mbar_arrive_expect_tx(&full_bar[stage], tx);
                    load_a(&tmap_a2, true, stage, cc * BLOCK_K, base_w2, base_h2, img, 0, 0);
                    load_b(stage, kcol, nrow, piece);
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ======================================================================= MMA issuer
        if ((!CTA2 || rank == 0) && elect_one()) {
            constexpr uint32_t idesc_full = umma_idesc_bf16(CTA2 ? 2 * BLOCK_M : BLOCK_M, BLOCK_N);
            constexpr uint32_t idesc_piece = umma_idesc_bf16(BLOCK_M, 64);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = first_tile; tile < total_items; tile += tile_step) {
                const uint32_t idesc = tile >= full_tiles ? idesc_piece : idesc_full;
                mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BLOCK_N);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint64_t a_desc = umma_desc_kmajor(smem_u32(smem_a + stage * C::A_BYTES), C::SWIZZLE_BYTES);
                    const uint64_t b_desc = umma_desc_kmajor(smem_u32(smem_b + stage * C::B_BYTES), C::SWIZZLE_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        // advancing K inside a swizzle row: +32 bytes = +2 in the (addr >> 4) field
                        if (CTA2) {
                            umma_bf16_pair(d_tmem, a_desc + static_cast<uint64_t>(k * 2), b_desc + static_cast<uint64_t>(k * 2),
                                           idesc, static_cast<uint32_t>((kb | k) != 0));
                        } else {
                            umma_bf16(d_tmem, a_desc + static_cast<uint64_t>(k * 2), b_desc + static_cast<uint64_t>(k * 2),
                                      idesc, static_cast<uint32_t>((kb | k) != 0));
                        }
                    }
                    // smem slot is free once these MMAs retire (in both CTAs of a pair)
                    if (CTA2) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
                // accumulator complete -> epilogue (of both CTAs)
                if (CTA2) umma_commit_pair(&tmem_full_bar[acc]); else umma_commit(&tmem_full_bar[acc]);
                if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1u; }
            }
        }
    } else if (warp_idx == 2) {
        // ================================================================== residual loader
        if (HAS_RES && elect_one()) {
            // chunk q (numbered over the CTA's whole walk) belongs to group q & 1 and is that group's (q >> 1)-th
            // chunk: tile 2 * ((q >> 1) & 1) + (q & 1), i.e. the four tiles are simply used round-robin
            int rs = 0;
            uint32_t rphase = 0;
            for (int tile = first_tile; tile < total_items; tile += tile_step) {
                int m_blk, n_blk, sub;
                tile_coords(tile, m_blk, n_blk, sub);
                const int n_chunks = sub >= 0 ? 1 : BLOCK_N / 64;
                for (int c64 = 0; c64 < n_chunks; ++c64) {
                    mbar_wait(&res_empty_bar[rs], rphase ^ 1u);
                    mbar_arrive_expect_tx(&res_full_bar[rs], OUT_CHUNK_BYTES);
                    tma_load_2d_hint(&tmap_res, &res_full_bar[rs], smem_res + rs * OUT_CHUNK_BYTES,
                                     n_blk * BLOCK_N + (sub >= 0 ? sub : c64) * 64, m_blk * BLOCK_M, args.pol_res);
                    if (++rs == RES_STAGES) { rs = 0; rphase ^= 1u; }
                }
            }
        }
    } else {
        // ========================================================================= epilogue
        // Two groups of four warps drain the accumulator: the 64-column chunks are numbered globally
        // (tile after tile) and group g takes the chunks with sequence number % 2 == g.  That number is
        // also the staging buffer, so each group owns its buffers and the two never wait on one another.
        // Without a residual a group has ONE staging tile.  With a residual it has TWO tiles that are residual input
        // and output staging at once: the loader fills a tile with the residual, the group adds its accumulator in
        // place (a thread only touches its own row) and the tile is stored from where it lies; as soon as that store
        // has read the tile the loader may fetch the residual of the group's chunk after next into it.
        static_assert(RES_STAGES == 4 && OUT_BUFS == 2, "two in/out tiles per group with a residual, one staging tile without");
        const int quad = warp_idx & 3;                  // TMEM lanes [32*quad, 32*quad+32)
        const int group = (warp_idx - 3) >> 2;          // 0 or 1
        const int lane = static_cast<int>(lane_id());
        const int row_in_tile = quad * 32 + lane;
        const bool store_leader = ((warp_idx - 3) & 3) == 0 && (lane == 0);
        const uint32_t bar_id = 1u + static_cast<uint32_t>(group);
        int acc = 0;
        uint32_t acc_phase = 0;
        uint32_t nloc = 0;                              // this group's chunk counter (residual kernels)
        const int ob = group;
        uint32_t seq = 0;                               // global chunk sequence number at the start of the tile
        const uint32_t tmem_empty0 = CTA2 ? mapa_shared(smem_u32(&tmem_empty_bar[0]), 0) : 0u;   // the leader's
        for (int tile = first_tile; tile < total_items; tile += tile_step) {
            int m_blk, n_blk, sub;
            tile_coords(tile, m_blk, n_blk, sub);
            const int n_chunks = sub >= 0 ? 1 : BLOCK_N / 64;
            mbar_wait(&tmem_full_bar[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) +
                                    static_cast<uint32_t>(acc * BLOCK_N);
            if (OUT_F32 && group == 0) {
                // ---- fp32 logits: direct, column-masked stores (tiny: n x 1000), first group only
                const int row = m_blk * BLOCK_M + row_in_tile;
                const bool row_ok = row < args.m_rows;
#pragma unroll 1
                for (int chunk = 0; chunk < BLOCK_N / 32; ++chunk) {
                    uint32_t v[32];
                    tmem_ld_32x32(t_addr + static_cast<uint32_t>(chunk * 32), v);
                    tmem_ld_wait();
                    const int col0 = n_blk * BLOCK_N + chunk * 32;
                    if (row_ok) {
                        float* o = reinterpret_cast<float*>(args.out) + static_cast<size_t>(row) * args.ldc + col0;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (col0 + j < args.n_real) {
                                const float x = __uint_as_float(v[j]) + __ldg(args.bias + col0 + j);
                                o[j] = args.relu ? fmaxf(x, 0.f) : x;
                            }
                        }
                    }
                }
            } else if (!OUT_F32) {
                // ---- bf16 NHWC: TMEM -> regs -> (+bias, +residual, ReLU) -> swizzled smem -> TMA store
#pragma unroll 1
                for (int c64 = 0; c64 < n_chunks; ++c64) {
                    if (((seq + static_cast<uint32_t>(c64)) & 1u) != static_cast<uint32_t>(group)) continue;
                    const int col0 = n_blk * BLOCK_N + (sub >= 0 ? sub : c64) * 64;   // a piece sits in TMEM columns 0..63
                    uint32_t res_base = 0;
                    const int io = group + 2 * static_cast<int>(nloc & 1u);            // this chunk's in/out tile
                    if (HAS_RES) {
                        if (store_leader && nloc > 0) {
                            // the previous chunk's store has read its tile: hand the tile back to the loader now, so
                            // that the residual of the NEXT chunk travels while this one is being worked on
                            tma_store_wait_read<0>();
                            mbar_arrive(&res_empty_bar[group + 2 * static_cast<int>((nloc - 1) & 1u)]);
                        }
                        mbar_wait(&res_full_bar[io], (nloc >> 1) & 1u);
                        res_base = smem_u32(smem_res + io * OUT_CHUNK_BYTES) + static_cast<uint32_t>(row_in_tile * 128);
                    }
                    uint4 packed[8];
                    // both halves of the chunk are requested from TMEM before anything waits, and the bias / residual
                    // loads are issued under that latency
                    uint32_t v[64];
                    tmem_ld_32x32(t_addr + static_cast<uint32_t>(c64 * 64), *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                    tmem_ld_32x32(t_addr + static_cast<uint32_t>(c64 * 64 + 32), *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                    const float4* b4 = reinterpret_cast<const float4*>(args.bias + col0);
                    if (args.relu != 2) {
                        // packed path (no activation / ReLU): fp32 pairs, two results per instruction, same roundings
                        uint4 r[8];
                        if (HAS_RES) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                r[j] = ld_shared_v4(res_base + static_cast<uint32_t>((j ^ (row_in_tile & 7)) * 16));
                        }
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 8; ++j) {           // 8 channels per step
                            const float4 ba = __ldg(b4 + 2 * j), bb = __ldg(b4 + 2 * j + 1);
                            uint64_t a0 = f32x2_add(f32x2_pack(v[8 * j + 0], v[8 * j + 1]), f32x2_pack(ba.x, ba.y));
                            uint64_t a1 = f32x2_add(f32x2_pack(v[8 * j + 2], v[8 * j + 3]), f32x2_pack(ba.z, ba.w));
                            uint64_t a2 = f32x2_add(f32x2_pack(v[8 * j + 4], v[8 * j + 5]), f32x2_pack(bb.x, bb.y));
                            uint64_t a3 = f32x2_add(f32x2_pack(v[8 * j + 6], v[8 * j + 7]), f32x2_pack(bb.z, bb.w));
                            if (HAS_RES) {
                                a0 = f32x2_add(a0, f32x2_from_bf16x2(r[j].x));
                                a1 = f32x2_add(a1, f32x2_from_bf16x2(r[j].y));
                                a2 = f32x2_add(a2, f32x2_from_bf16x2(r[j].z));
                                a3 = f32x2_add(a3, f32x2_from_bf16x2(r[j].w));
                            }
                            uint4 o;
                            o.x = bf16x2_from_f32x2(a0); o.y = bf16x2_from_f32x2(a1);
                            o.z = bf16x2_from_f32x2(a2); o.w = bf16x2_from_f32x2(a3);
                            if (args.relu == 1) {
                                o.x = bf16x2_relu(o.x); o.y = bf16x2_relu(o.y); o.z = bf16x2_relu(o.z); o.w = bf16x2_relu(o.w);
                            }
                            packed[j] = o;
                        }
                    } else {
                        tmem_ld_wait();
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            float f[32];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4 b = __ldg(b4 + h * 8 + j);
                                f[4 * j + 0] = __uint_as_float(v[h * 32 + 4 * j + 0]) + b.x;
                                f[4 * j + 1] = __uint_as_float(v[h * 32 + 4 * j + 1]) + b.y;
                                f[4 * j + 2] = __uint_as_float(v[h * 32 + 4 * j + 2]) + b.z;
                                f[4 * j + 3] = __uint_as_float(v[h * 32 + 4 * j + 3]) + b.w;
                            }
                            if (HAS_RES) {
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    const int chunk16 = (h * 4 + j) ^ (row_in_tile & 7);     // 128B swizzle
                                    const uint4 r = ld_shared_v4(res_base + static_cast<uint32_t>(chunk16 * 16));
                                    f[8 * j + 0] += bf16_lo(r.x); f[8 * j + 1] += bf16_hi(r.x);
                                    f[8 * j + 2] += bf16_lo(r.y); f[8 * j + 3] += bf16_hi(r.y);
                                    f[8 * j + 4] += bf16_lo(r.z); f[8 * j + 5] += bf16_hi(r.z);
                                    f[8 * j + 6] += bf16_lo(r.w); f[8 * j + 7] += bf16_hi(r.w);
                                }
                            }
                            // exact GELU, nn.GELU() default: 0.5 x (1 + erf(x / sqrt(2)))
#pragma unroll
                            for (int j = 0; j < 32; ++j) f[j] = gelu_erf(f[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                packed[h * 4 + j].x = pack_bf16x2(f[8 * j + 0], f[8 * j + 1]);
                                packed[h * 4 + j].y = pack_bf16x2(f[8 * j + 2], f[8 * j + 3]);
                                packed[h * 4 + j].z = pack_bf16x2(f[8 * j + 4], f[8 * j + 5]);
                                packed[h * 4 + j].w = pack_bf16x2(f[8 * j + 6], f[8 * j + 7]);
                            }
                        }
                    }
                    uint8_t* stage_tile;
                    if (HAS_RES) {
                        stage_tile = smem_res + io * OUT_CHUNK_BYTES;       // in place: every thread rewrites its own row
                        ++nloc;
                    } else {
                        stage_tile = smem_out + ob * OUT_CHUNK_BYTES;
                        // this group's staging buffer was handed to the TMA engine one chunk ago: it must have been read
                        if (store_leader) tma_store_wait_read<0>();
                        named_bar_sync(bar_id, EPI_THREADS);
                    }
                    const uint32_t o_base = smem_u32(stage_tile) + static_cast<uint32_t>(row_in_tile * 128);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int chunk16 = j ^ (row_in_tile & 7);
                        st_shared_v4(o_base + static_cast<uint32_t>(chunk16 * 16), packed[j]);
                    }
                    fence_proxy_async_smem();
                    named_bar_sync(bar_id, EPI_THREADS);
                    if (store_leader) {
                        tma_store_2d_hint(&tmap_out, stage_tile, col0, m_blk * BLOCK_M, args.pol_out);
                        tma_store_commit();
                    }
                }
            }
            // every TMEM read this warp makes for the tile has been issued and waited for (the packed values
            // of its last chunk are in registers / shared memory): hand the accumulator back
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (CTA2) mbar_arrive_cluster(tmem_empty0 + static_cast<uint32_t>(acc * 8));
                else mbar_arrive(&tmem_empty_bar[acc]);
            }
            seq += static_cast<uint32_t>(n_chunks);
            if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1u; }
        }
        if (!OUT_F32 && store_leader) tma_store_wait_all<0>();
    }

    tc_fence_before();
    __syncwarp();
    // a pair must leave together: the leader's MMAs read the peer's shared memory and signal its barriers
    if (CTA2) cluster_sync_all(); else __syncthreads();
    if (warp_idx == 1) {
        tc_fence_after();
        if (CTA2) tmem_dealloc_pair(tmem_base, C::TMEM_COLS); else tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

// --------------------------------------------------------------------------- host side

PFN_cuTensorMapEncodeTiled_v12000 g_encode_tiled = nullptr;
PFN_cuTensorMapEncodeIm2col_v12000 g_encode_im2col = nullptr;

void set_err(char* err, size_t n, const char* msg) {
    if (err && n) snprintf(err, n, "%s", msg);
}

template <int BLOCK_N, int BLOCK_K, bool IM2COL, bool OUT_F32, bool HAS_RES, bool CTA2>
cudaError_t launch_variant(const GemmPlan& p, const GemmArgs& a, int grid, cudaStream_t stream) {
    using C = Cfg<BLOCK_N, BLOCK_K, OUT_F32, HAS_RES, CTA2>;
    auto kern = conv_gemm_kernel<BLOCK_N, BLOCK_K, IM2COL, OUT_F32, HAS_RES, CTA2>;
    static std::atomic<bool> attr_set[64];   // per instantiation, per device; zero-initialised, set from any thread
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             C::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_set[dev].store(true, std::memory_order_release);
    }
    if (!CTA2) {
        return launch_pdl(kern, dim3(grid), dim3(NUM_THREADS), C::SMEM_BYTES, stream, p.tmap_a, p.tmap_b, p.tmap_out,
                          p.tmap_res, p.tmap_a2, p.tmap_b64, a);
    }
    // CTA pairs: a 2-wide cluster (the two SMs of a TPC) on top of the programmatic-serialization attribute
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, kern, p.tmap_a, p.tmap_b, p.tmap_out, p.tmap_res, p.tmap_a2, p.tmap_b64, a);
}

}  // namespace

bool tma_init(char* err, size_t err_len) {
    if (g_encode_tiled && g_encode_im2col) return true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || fn == nullptr) {
        set_err(err, err_len, "cuTensorMapEncodeTiled not available from the driver");
        return false;
    }
    g_encode_tiled = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    fn = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || fn == nullptr) {
        set_err(err, err_len, "cuTensorMapEncodeIm2col not available from the driver");
        return false;
    }
    g_encode_im2col = reinterpret_cast<PFN_cuTensorMapEncodeIm2col_v12000>(fn);
    return true;
}

static bool encode_tiled_2d(CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows,
                            uint32_t box_inner, uint32_t box_rows, CUtensorMapSwizzle sw, char* err,
                            size_t err_len) {
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {inner * 2};
    cuuint32_t box[2] = {box_inner, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                                strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char b[160];
        snprintf(b, sizeof b, "cuTensorMapEncodeTiled failed (%d) inner=%llu rows=%llu box=%ux%u",
                 static_cast<int>(r), (unsigned long long)inner, (unsigned long long)rows, box_inner,
                 box_rows);
        set_err(err, err_len, b);
        return false;
    }
    return true;
}

bool gemm_plan_init(GemmPlan& p, const b2p_op& op, const void* weight_dev, int64_t weight_bytes,
                    const float* bias_dev, int rows_at_capacity, int num_sms, char* err,
                    size_t err_len) {
    const bool linear = op.kind == B2P_OP_LINEAR;
    p.in_h = linear ? 1 : op.in_h;
    p.in_w = linear ? 1 : op.in_w;
    p.in_c = op.in_c;
    p.out_h = linear ? 1 : op.out_h;
    p.out_w = linear ? 1 : op.out_w;
    p.taps_h = linear ? 1 : op.kh;
    p.taps_w = linear ? 1 : op.kw;
    p.stride = linear ? 1 : op.stride;
    p.pad_lo = linear ? 0 : op.pad_lo;
    p.pad_hi = linear ? 0 : op.pad_hi;
    p.relu = op.relu;
    p.out_f32 = linear ? 1 : 0;
    p.n_real = op.out_c;
    p.ldc = op.out_c;
    p.bias = bias_dev;
    p.weight = weight_dev;
    if (op.in_c % 64 == 0) {
        p.block_k = 64;
    } else if (op.in_c == 16) {
        p.block_k = 16;
    } else {
        set_err(err, err_len, "conv/linear in_c must be a multiple of 64 (or exactly 16 for the stem)");
        return false;
    }
    p.k_chunks = op.in_c / p.block_k;
    p.im2col = !(p.taps_h == 1 && p.taps_w == 1 && p.stride == 1 && p.pad_lo == 0 && p.pad_hi == 0);
    p.src2_c = (!linear && op.src2_c > 0) ? op.src2_c : 0;
    p.src2_stride = op.src2_stride > 0 ? op.src2_stride : 1;
    if (p.src2_c % 64 != 0 || (p.src2_c > 0 && p.block_k != 64)) {
        set_err(err, err_len, "second GEMM source needs channels that are a multiple of 64");
        return false;
    }
    const int64_t k_total = static_cast<int64_t>(p.taps_h) * p.taps_w * op.in_c + p.src2_c;
    if (weight_bytes % (k_total * 2) != 0) {
        set_err(err, err_len, "weight blob size is not a whole number of K-rows");
        return false;
    }
    p.n_pad = static_cast<int>(weight_bytes / (k_total * 2));
    if (p.n_pad < op.out_c || p.n_pad % 64 != 0) {
        set_err(err, err_len, "weight rows must be >= out_c and padded to a multiple of 64");
        return false;
    }
    if (!p.out_f32 && op.out_c % 64 != 0) {
        set_err(err, err_len, "conv out_c must be a multiple of 64");
        return false;
    }
    // Tile width: fewest (waves x tile width) at capacity; ties go to the wider tile, which
    // re-reads the activation tile fewer times.
    // CTA pairs (cta_group::2, 256-row MMAs) halve the weight traffic per SM.  Measured per layer on ResNet-50 b256
    // (profiles/r01g_cta2_ab.txt): 5-12 % faster for 256-wide tiles without a residual, 15-35 % SLOWER for
    // 128-wide tiles and slightly slower with a residual stream, so that is the policy.  B2P_CTA2=all / none
    // overrides it for A/B measurements.
    static const char* cta2_env = getenv("B2P_CTA2");
    const bool cta2_all = cta2_env && strcmp(cta2_env, "all") == 0;
    const bool cta2_none = cta2_env && strcmp(cta2_env, "none") == 0;
    const bool has_res_op = op.res != B2P_SLOT_NONE;
    const bool cta2_ok = !cta2_none && !p.out_f32 && p.block_k == 64 && num_sms >= 2;
    const int m_tiles = (rows_at_capacity + BLOCK_M - 1) / BLOCK_M;
    // Tail splitting is opt-in (B2P_TAIL_SPLIT=1, read per plan): A/B runs put it within the box-to-box noise on
    // ResNet-50 / ViT-B/16 / ResNet-18 (profiles/r01i_tail_split_ab.txt) — the GEMMs here are bound by operand
    // delivery, and a 64-column piece still loads the whole 128-row activation tile.
    p.tail_split = getenv("B2P_TAIL_SPLIT") != nullptr && !p.out_f32 && p.block_k == 64;
    int best = 0, best_cta2 = 0;
    double best_cost = 0;
    const int cands[3] = {256, 128, 64};
    for (int bn : cands) {
        if (p.force_block_n && bn != p.force_block_n) continue;
        if (p.n_pad % bn != 0) continue;
        if (p.block_k == 16 && bn != 64) continue;
        if (p.out_f32 && bn == 256) continue;      // fp32 logits: 128- or 64-wide tiles (the FC layer has 16 / 32 of them)
        const double tile_cost = bn + 32;           // +32: fixed per-tile overhead in "column" units
        // single CTAs; a nearly empty last wave is cut into 64-column pieces (see gemm_tail_split)
        {
            const long tiles = static_cast<long>(m_tiles) * (p.n_pad / bn);
            const long q = tiles / num_sms, r = tiles % num_sms;
            // (the tail split applied at launch is deliberately not part of this choice: letting it pick wider tiles
            // for the residual layers measured slower, profiles/r01i_tail_split_ab.txt)
            const double cost = q * tile_cost + (r ? tile_cost : 0.0);
            if (best == 0 || cost < best_cost) { best = bn; best_cost = cost; best_cta2 = 0; }
        }
        // CTA pairs: measured 7 % cheaper per tile for the layers the policy admits, no tail splitting
        static const bool cta2_res = getenv("B2P_CTA2_RES") != nullptr;     // experiment: pairs for residual layers too
        if (cta2_ok && (cta2_all ? bn >= 128 : (bn == 256 && (!has_res_op || cta2_res)))) {
            const long pairs = num_sms / 2;
            const long tiles = static_cast<long>((m_tiles + 1) / 2) * (p.n_pad / bn);
            // a pair tile pulls half of the weight bytes per CTA through the L2, which is what bounds these layers: worth
            // an extra partly filled wave (ResNet-50's 7x7 layers; A/B profiles/r02zb_l2_smem_bounds.txt, section 13)
            static const double pair_bias = getenv("B2P_PAIR_BIAS") ? atof(getenv("B2P_PAIR_BIAS")) : 0.8;
            const double cost = static_cast<double>((tiles + pairs - 1) / pairs) * tile_cost * pair_bias;
            if (best == 0 || cost <= best_cost) { best = bn; best_cost = cost; best_cta2 = 1; }
        }
    }
    if (best == 0) {
        set_err(err, err_len, "no tile width fits this layer");
        return false;
    }
    p.block_n = best;
    p.cta2 = best_cta2;
    const CUtensorMapSwizzle sw = p.block_k == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B;
    if (!encode_tiled_2d(&p.tmap_b, weight_dev, static_cast<uint64_t>(k_total),
                         static_cast<uint64_t>(p.n_pad), static_cast<uint32_t>(p.block_k),
                         static_cast<uint32_t>(p.cta2 ? p.block_n / 2 : p.block_n), sw, err, err_len))
        return false;
    // 64-row weight boxes for the pieces of a split last wave (also encoded when unused: the kernel takes the map)
    return encode_tiled_2d(&p.tmap_b64, weight_dev, static_cast<uint64_t>(k_total), static_cast<uint64_t>(p.n_pad),
                           static_cast<uint32_t>(p.block_k), 64, sw, err, err_len);
}

bool gemm_plan_set_src(GemmPlan& p, const void* src, int batch_cap, char* err, size_t err_len) {
    const CUtensorMapSwizzle sw = p.block_k == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B;
    if (!p.im2col) {
        const uint64_t rows = static_cast<uint64_t>(batch_cap) * p.in_h * p.in_w;
        return encode_tiled_2d(&p.tmap_a, src, static_cast<uint64_t>(p.in_c), rows,
                               static_cast<uint32_t>(p.block_k), BLOCK_M, sw, err, err_len);
    }
    // NHWC seen by TMA as {C, W, H, N}.  The corners bound where the *base* pixel of a filter
    // window may sit: lower = -pad_lo, upper = pad_hi - (filter - 1).
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(p.in_c), static_cast<cuuint64_t>(p.in_w),
                          static_cast<cuuint64_t>(p.in_h), static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(p.in_c) * 2,
                             static_cast<cuuint64_t>(p.in_c) * 2 * p.in_w,
                             static_cast<cuuint64_t>(p.in_c) * 2 * p.in_w * p.in_h};
    int lower[2] = {-p.pad_lo, -p.pad_lo};
    int upper[2] = {p.pad_hi - (p.taps_w - 1), p.pad_hi - (p.taps_h - 1)};
    cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(p.stride), static_cast<cuuint32_t>(p.stride), 1};
    CUresult r = g_encode_im2col(&p.tmap_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src),
                                 dims, strides, lower, upper, static_cast<cuuint32_t>(p.block_k),
                                 BLOCK_M, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char b[200];
        snprintf(b, sizeof b,
                 "cuTensorMapEncodeIm2col failed (%d) c=%d w=%d h=%d n=%d k=%dx%d stride=%d pad=%d/%d",
                 static_cast<int>(r), p.in_c, p.in_w, p.in_h, batch_cap, p.taps_h, p.taps_w, p.stride,
                 p.pad_lo, p.pad_hi);
        set_err(err, err_len, b);
        return false;
    }
    // Small-tensor erratum handled the same way CUTLASS does for drivers <= 13.1: clear bit 21 of
    // the descriptor's second word when the tensor is under 128 KiB.
    int drv = 0;
    cudaDriverGetVersion(&drv);
    const uint64_t bytes = static_cast<uint64_t>(batch_cap) * p.in_h * p.in_w * p.in_c * 2;
    if (drv <= 13010 && bytes < 131072) {
        reinterpret_cast<uint64_t*>(&p.tmap_a)[1] &= ~(1ull << 21);
    }
    return true;
}

bool gemm_plan_set_src2(GemmPlan& p, const void* src2, int batch_cap, char* err, size_t err_len) {
    if (p.src2_c <= 0) return true;
    const int h2 = p.out_h * p.src2_stride, w2 = p.out_w * p.src2_stride;
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(p.src2_c), static_cast<cuuint64_t>(w2), static_cast<cuuint64_t>(h2),
                          static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(p.src2_c) * 2, static_cast<cuuint64_t>(p.src2_c) * 2 * w2,
                             static_cast<cuuint64_t>(p.src2_c) * 2 * w2 * h2};
    int lower[2] = {0, 0}, upper[2] = {0, 0};
    cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(p.src2_stride), static_cast<cuuint32_t>(p.src2_stride), 1};
    CUresult r = g_encode_im2col(&p.tmap_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src2), dims, strides,
                                 lower, upper, 64, BLOCK_M, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_err(err, err_len, "cuTensorMapEncodeIm2col failed for the second GEMM source");
        return false;
    }
    int drv = 0;
    cudaDriverGetVersion(&drv);
    const uint64_t bytes = static_cast<uint64_t>(batch_cap) * h2 * w2 * p.src2_c * 2;
    if (drv <= 13010 && bytes < 131072) reinterpret_cast<uint64_t*>(&p.tmap_a2)[1] &= ~(1ull << 21);
    return true;
}

bool gemm_plan_set_dst(GemmPlan& p, const void* dst, int rows, char* err, size_t err_len) {
    if (p.out_f32) return true;     // the fp32 epilogue stores directly
    return encode_tiled_2d(&p.tmap_out, dst, static_cast<uint64_t>(p.ldc), static_cast<uint64_t>(rows), 64, BLOCK_M,
                           CU_TENSOR_MAP_SWIZZLE_128B, err, err_len);
}

bool gemm_plan_set_res(GemmPlan& p, const void* res, int rows, char* err, size_t err_len) {
    return encode_tiled_2d(&p.tmap_res, res, static_cast<uint64_t>(p.ldc), static_cast<uint64_t>(rows), 64, BLOCK_M,
                           CU_TENSOR_MAP_SWIZZLE_128B, err, err_len);
}

cudaError_t gemm_plan_launch(const GemmPlan& p, int m_rows, bool has_res, void* out_f32, int num_sms,
                             cudaStream_t stream) {
    GemmArgs a;
    a.m_rows = m_rows;
    a.m_tiles = (m_rows + BLOCK_M - 1) / BLOCK_M;
    a.n_tiles = p.n_pad / p.block_n;
    a.n_real = p.n_real;
    a.taps_h = p.taps_h;
    a.taps_w = p.taps_w;
    a.k_chunks = p.k_chunks;
    a.k2_chunks = p.src2_c / 64;
    a.stride2 = p.src2_stride;
    a.out_h = p.out_h;
    a.out_w = p.out_w;
    a.stride = p.stride;
    a.pad_lo = p.pad_lo;
    a.ldc = p.ldc;
    a.relu = p.relu;
    a.reverse = p.reverse;
    a.bias = p.bias;
    a.out = out_f32;
    a.pol_a = p.pol_a;
    a.pol_b = L2_EVICT_LAST;        // weights: re-read by every tile of every CTA
    a.pol_res = p.pol_res;
    a.pol_out = p.pol_out;
    int total = a.m_tiles * a.n_tiles;
    if (total <= 0) return cudaSuccess;
    if (p.cta2) total = 2 * ((a.m_tiles + 1) / 2) * a.n_tiles;      // CTAs, two per 256-row tile
    const int sms = p.cta2 ? (num_sms & ~1) : num_sms;
    a.split_tiles = 0;
    a.split_ways = 0;
    if (p.tail_split && !p.cta2 && p.block_n > 64 && total > sms) {
        // gemm_tail_split: the r tiles of the last, partly filled wave become r * (block_n / 64) pieces when that
        // makes the wave shorter (a piece costs about (64 + 32) / (block_n + 32) of a tile)
        const int r = total % sms, s = p.block_n / 64;
        if (r > 0 && static_cast<long>((static_cast<long>(r) * s + sms - 1) / sms) * (64 + 32) < p.block_n + 32) {
            a.split_tiles = r;
            a.split_ways = s;
            total += r * (s - 1);
        }
    }
    const int grid = total < sms ? total : sms;
#define B2P_LAUNCH(BN, BK, IM, F32, RES) return launch_variant<BN, BK, IM, F32, RES, false>(p, a, grid, stream)
#define B2P_LAUNCH2(BN, IM, RES) return launch_variant<BN, 64, IM, false, RES, true>(p, a, grid, stream)
    if (p.cta2) {
#define B2P_PICK2(BN)                                          \
    if (p.block_n == BN) {                                     \
        if (p.im2col && has_res) B2P_LAUNCH2(BN, true, true);  \
        if (p.im2col) B2P_LAUNCH2(BN, true, false);            \
        if (has_res) B2P_LAUNCH2(BN, false, true);             \
        B2P_LAUNCH2(BN, false, false);                         \
    }
        B2P_PICK2(128)
        B2P_PICK2(256)
#undef B2P_PICK2
        return cudaErrorInvalidConfiguration;
    }
    if (p.block_k == 16) {
        if (p.block_n == 64 && p.im2col && !p.out_f32 && !has_res) B2P_LAUNCH(64, 16, true, false, false);
        return cudaErrorInvalidConfiguration;
    }
    if (p.out_f32) {
        if (p.block_n == 128 && !p.im2col && !has_res) B2P_LAUNCH(128, 64, false, true, false);
        if (p.block_n == 64 && !p.im2col && !has_res) B2P_LAUNCH(64, 64, false, true, false);
        return cudaErrorInvalidConfiguration;
    }
#define B2P_PICK(BN)                                                   \
    if (p.block_n == BN) {                                             \
        if (p.im2col && has_res) B2P_LAUNCH(BN, 64, true, false, true); \
        if (p.im2col) B2P_LAUNCH(BN, 64, true, false, false);          \
        if (has_res) B2P_LAUNCH(BN, 64, false, false, true);           \
        B2P_LAUNCH(BN, 64, false, false, false);                       \
    }
    B2P_PICK(64)
    B2P_PICK(128)
    B2P_PICK(256)
#undef B2P_PICK
#undef B2P_LAUNCH
#undef B2P_LAUNCH2
    return cudaErrorInvalidConfiguration;
}

}  // namespace b2p
