// K2 fusion: a whole identity bottleneck block of the 56x56 stage as ONE persistent tcgen05 kernel
// (SP/torchvision/models/resnet.py:143-163):
//
//   t1  = relu(W1 x  + b1)        1x1, 256 -> 64
//   t2  = relu(W2 * t1 + b2)      3x3 / pad 1, 64 -> 64
//   out = relu(W3 t2 + b3 + x)    1x1, 64 -> 256, identity shortcut
//
// Layer by layer the two 64-channel intermediates make an HBM round trip each (4 x 103 MB per 256 images and block)
// and the 3x3 runs as a kernel of its own.  Here only the block input is read and the block output written.
//
// Work unit: one CTA takes two output rows of one image.
//   * the x patch - 4 input rows (2 output rows + halo) x 64 pixels (58 used) x 256 ch - streams through a ring of
//     [2 rows x 64 px x 64 ch] TMA boxes (128B swizzle, out-of-image pixels arrive as zeros), each with its 64-column
//     slice of W1; conv1 runs on all four rows (the halo rows are recomputed: 2x the conv1 math, no exchange between
//     CTAs).  The boxes of the tile after next are requested into the L2 ahead of time (TMA prefetch);
//   * epilogues A / C turn the conv1 accumulator into the bf16 t1 PATCH in shared memory (zero outside the image:
//     that is conv2's padding) in exactly the K-major swizzled layout a TMA load would have produced;
//   * conv2 = 9 taps x 4 K-steps of MMAs whose A descriptor is the patch start shifted by (r * 64 + s) pixels
//     (the row-shift trick of conv3x3_c64.cu), W2 (72 KB) resident; epilogue A writes t2 as the next A operand;
//   * conv3 (N = 256, W3 resident) accumulates into 256 TMEM columns; epilogue B adds bias and the residual (read from
//     global memory: the same pixels went through the x ring a moment ago, so it is an L2 hit), applies ReLU and
//     TMA-stores two 56-pixel rows per 64-channel chunk.
// TMEM (512 columns): conv1 2 x 64 | conv2 64 | - | conv3 256.
//
// Why not a CTA pair (cta_group::2) with the weights split across two SMs, which would leave room for deeper rings:
// measured, a cta_group::2 MMA with N = 64 takes as long as one with N = 256 (the kernel ran its tensor pipe
// back to back at 24 % of the peak op rate: profiles/r02k_bottleneck_pair_ncu.txt), so the 68 narrow MMAs of
// conv1 + conv2 cost four times their single-CTA price.
//
// Warps (512 threads): 0 = TMA producer (weights once, then x / W1 boxes), 1 = TMEM owner + MMA issuer, 2-3 idle,
// 4-7 = epilogue A (t1 patch rows 0-1, then t2), 12-15 = epilogue C (t1 patch rows 2-3), 8-11 = epilogue B (conv3
// accumulator + residual -> output).  The MMA order per tile i is conv2(i), conv1(i+1), conv3(i), so the tensor pipe
// works on the next tile's conv1 while epilogue A produces t2(i).
#include <atomic>
#include <cstdio>
#include <cudaTypedefs.h>
#include "kernels.h"
#include "ptx.cuh"

namespace b2p {
namespace {

constexpr int HW = 56;
constexpr int MID = 64;                     // bottleneck width of the 56x56 stage
constexpr int OUTC = 256;
constexpr int PITCH = 64;                   // patch row pitch in pixels
constexpr int TILE16K = 128 * 128;          // 128 rows x 64 ch bf16
constexpr int PATCH_BYTES = 4 * PITCH * 128 + 1024;   // 4 rows + the rows a shifted descriptor reads past the end
constexpr int X_STAGES = 3;
constexpr int BN_THREADS = 512;
constexpr uint32_t COL_ACC1 = 0, COL_ACC2 = 128, COL_ACC3 = 256;

template <int KC1, bool PROJ>
struct BnCfg {
    static constexpr int W1_BYTES = KC1 * 32 * 128;            // this CTA's 32 of the 64 conv1 output rows
    static constexpr int W2_BYTES = 9 * 32 * 128;
    static constexpr int W3_BYTES = (PROJ ? 2 : 1) * 128 * 128;  // this CTA's 128 of the 256 conv3 output rows
    static constexpr int IO_TILES = PROJ ? 2 : 3;               // staging tiles (identity blocks: residual in, result out)
    static constexpr int XC_BYTES = PROJ ? TILE16K : 0;         // the block input at the output pixels (shortcut operand)
    static constexpr int SMEM = 1024 + W1_BYTES + W2_BYTES + W3_BYTES + PATCH_BYTES + TILE16K + XC_BYTES +
                                X_STAGES * TILE16K + IO_TILES * TILE16K + 512;
    static_assert(SMEM <= 227 * 1024, "shared memory budget");
};

struct BnArgs {
    int n_imgs;
    const float* b1;
    const float* b2;
    const float* b3;
};

__device__ __forceinline__ void tma_load_4d_pair(const CUtensorMap* m, uint32_t bar_cluster_addr, void* dst, int32_t c0,
                                                 int32_t c1, int32_t c2, int32_t c3, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
        : "memory");
}
// Asks the L2 for a box ahead of time (no shared-memory destination, no barrier): the x ring is only three boxes deep,
// so a box that still has to come from HBM when it is requested (~2 us under load) would bound a CTA at ~25 GB/s.
__device__ __forceinline__ void tma_prefetch_4d(const CUtensorMap* m, int32_t c0, int32_t c1, int32_t c2, int32_t c3) {
    asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];" ::"l"(
                     reinterpret_cast<uint64_t>(m)),
                 "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
// Cluster-scope release / acquire on mbarriers: the data they order (the t1 patch, the t2 tile) is written by one
// CTA's threads and consumed by MMAs the OTHER CTA's thread issues.
__device__ __forceinline__ void mbar_arrive_remote_release(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// Waits with a watchdog: a protocol error becomes a trap (launch failure) after ~2 s instead of a hung GPU.
__device__ __forceinline__ void mbar_wait_wd(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    const long long t0 = clock64();
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (clock64() - t0 > 4000000000ll) {
            printf("bottleneck56: barrier %u (block %d thread %d) timed out, parity %u\n", addr, blockIdx.x, threadIdx.x, parity);
            __trap();
        }
    }
}

template <int KC1, bool PROJ>
__global__ void __launch_bounds__(BN_THREADS, 1)
bottleneck56_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w1,
                    const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_w3,
                    const __grid_constant__ CUtensorMap tmap_res, const __grid_constant__ CUtensorMap tmap_out,
                    const BnArgs args) {
    using C = BnCfg<KC1, PROJ>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* s_w1 = smem;
    uint8_t* s_w2 = s_w1 + C::W1_BYTES;
    uint8_t* s_w3 = s_w2 + C::W2_BYTES;
    uint8_t* s_patch = s_w3 + C::W3_BYTES;
    uint8_t* s_t2 = s_patch + PATCH_BYTES;
    uint8_t* s_xc = s_t2 + TILE16K;
    uint8_t* s_x = s_xc + C::XC_BYTES;
    uint8_t* s_io = s_x + X_STAGES * TILE16K;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_io + C::IO_TILES * TILE16K);
    uint64_t* x_full = bars;                    // [X_STAGES]  leader's copy counts the bytes of both CTAs
    uint64_t* x_empty = x_full + X_STAGES;      // [X_STAGES]  multicast commit
    uint64_t* acc1_full = x_empty + X_STAGES;   // commit
    uint64_t* acc1_empty = acc1_full + 1;       // leader's: 16 warp arrivals (the two conv1 epilogue groups of both CTAs)
    uint64_t* t1_ready = acc1_empty + 1;        // leader's: 16 warp arrivals
    uint64_t* t1_free = t1_ready + 1;           // commit
    uint64_t* acc2_full = t1_free + 1;          // commit
    uint64_t* acc2_empty = acc2_full + 1;       // leader's: 8
    uint64_t* t2_ready = acc2_empty + 1;        // leader's: 8
    uint64_t* t2_free = t2_ready + 1;           // commit
    uint64_t* acc3_full = t2_free + 1;          // commit
    uint64_t* acc3_empty = acc3_full + 1;       // leader's: 8 (epilogue B of both CTAs)
    uint64_t* xc_full = acc3_empty + 1;         // leader's, bytes of both CTAs (PROJ)
    uint64_t* xc_free = xc_full + 1;            // commit (PROJ)
    uint64_t* w_bar = xc_free + 1;              // leader's, bytes of both CTAs
    uint64_t* res_full = w_bar + 1;             // [3] local
    uint64_t* res_empty = res_full + 3;         // [3] local
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_empty + 3);

    const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const uint32_t rank = cluster_ctarank();
    const int pair_tiles = args.n_imgs * (HW / 4);
    const int first = static_cast<int>(blockIdx.x >> 1), step = static_cast<int>(gridDim.x >> 1);

    if (warp_idx == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_w1); tma_prefetch_desc(&tmap_w2);
            tma_prefetch_desc(&tmap_w3); tma_prefetch_desc(&tmap_out);
            if (!PROJ) tma_prefetch_desc(&tmap_res);
        }
    } else if (warp_idx == 1) {
        if (elect_one()) {
            for (int i = 0; i < X_STAGES; ++i) { mbar_init(&x_full[i], 1); mbar_init(&x_empty[i], 1); }
            mbar_init(acc1_full, 1); mbar_init(acc1_empty, 16);
            mbar_init(t1_ready, 16); mbar_init(t1_free, 1);
            mbar_init(acc2_full, 1); mbar_init(acc2_empty, 8);
            mbar_init(t2_ready, 8);  mbar_init(t2_free, 1);
            mbar_init(acc3_full, 1); mbar_init(acc3_empty, 8);
            mbar_init(xc_full, 1);   mbar_init(xc_free, 1);
            mbar_init(w_bar, 1);
            for (int i = 0; i < 3; ++i) { mbar_init(&res_full[i], 1); mbar_init(&res_empty[i], 1); }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_pair(tmem_slot, 512);
        tmem_relinquish_pair();
    }
    pdl_launch_dependents();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    // tile -> (image, first output row of THIS CTA)
    auto tile_pos = [&](int T, int& img, int& row0) {
        img = T / (HW / 4);
        row0 = (T - img * (HW / 4)) * 4 + 2 * static_cast<int>(rank);
    };

    if (warp_idx == 0) {
        // ================================================================ TMA producer: weights once, then x boxes
        if (elect_one()) {
            const uint32_t w_bar0 = mapa_shared(smem_u32(w_bar), 0);
            if (rank == 0) mbar_arrive_expect_tx(w_bar, 2u * (C::W1_BYTES + C::W2_BYTES + C::W3_BYTES));
            for (int kc = 0; kc < KC1; ++kc)
                tma_load_2d_pair_hint(&tmap_w1, w_bar0, s_w1 + kc * 4096, kc * 64, static_cast<int>(rank) * 32, L2_EVICT_LAST);
            for (int tap = 0; tap < 9; ++tap)
                tma_load_2d_pair_hint(&tmap_w2, w_bar0, s_w2 + tap * 4096, tap * 64, static_cast<int>(rank) * 32, L2_EVICT_LAST);
            for (int kc = 0; kc < (PROJ ? 2 : 1); ++kc)
                tma_load_2d_pair_hint(&tmap_w3, w_bar0, s_w3 + kc * TILE16K, kc * 64, static_cast<int>(rank) * 128, L2_EVICT_LAST);
            const uint32_t full0 = mapa_shared(smem_u32(&x_full[0]), 0);
            int stage = 0;
            uint32_t phase = 0;
            auto prefetch_tile = [&](int T) {
                if (T >= pair_tiles) return;
                int img, row0;
                tile_pos(T, img, row0);
                // this CTA's four input rows; rank 1's upper two are rank 0's lower two, so each CTA asks only for the
                // rows it reads FIRST (rank 0: rows 0-1, rank 1: rows 2-3 of its patch reach the L2 via the peer)
                for (int kc = 0; kc < KC1; ++kc) {
                    tma_prefetch_4d(&tmap_x, kc * 64, -1, row0 - 1, img);
                    tma_prefetch_4d(&tmap_x, kc * 64, -1, row0 + 1, img);
                }
            };
            prefetch_tile(first);
            prefetch_tile(first + step);
            for (int T = first; T < pair_tiles; T += step) {
                int img, row0;
                tile_pos(T, img, row0);
                prefetch_tile(T + 2 * step);
                for (int j = 0; j < 2; ++j) {
                    for (int kc = 0; kc < KC1; ++kc) {
                        mbar_wait_wd(&x_empty[stage], phase ^ 1u);
                        if (rank == 0) mbar_arrive_expect_tx(&x_full[stage], 2u * TILE16K);
                        // patch rows 2j, 2j+1 = image rows row0 - 1 + 2j ..; columns start at -1 (the left halo)
                        tma_load_4d_pair(&tmap_x, full0 + static_cast<uint32_t>(stage * 8), s_x + stage * TILE16K, kc * 64, -1,
                                         row0 - 1 + 2 * j, img, L2_EVICT_NORMAL);
                        if (++stage == X_STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ================================================================ MMA issuer (leader CTA)
        if (rank == 0 && elect_one()) {
            constexpr uint32_t idesc64 = umma_idesc_bf16(256, MID);
            constexpr uint32_t idesc256 = umma_idesc_bf16(256, OUTC);
            mbar_wait_wd(w_bar, 0);
            tc_fence_after();
            int stage = 0;
            uint32_t phase = 0;
            auto conv1 = [&](int it) {
                mbar_wait_wd(acc1_empty, (static_cast<uint32_t>(it) & 1u) ^ 1u);
                tc_fence_after();
                for (int j = 0; j < 2; ++j) {
                    const uint32_t d = tmem_base + COL_ACC1 + static_cast<uint32_t>(j * MID);
                    for (int kc = 0; kc < KC1; ++kc) {
                        mbar_wait_wd(&x_full[stage], phase);
                        tc_fence_after();
                        const uint64_t a = umma_desc_kmajor(smem_u32(s_x + stage * TILE16K), 128);
                        const uint64_t b = umma_desc_kmajor(smem_u32(s_w1 + kc * 4096), 128);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_pair(d, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc64,
                                           static_cast<uint32_t>((kc | k) != 0));
                        umma_commit_pair(&x_empty[stage]);
                        if (++stage == X_STAGES) { stage = 0; phase ^= 1u; }
                    }
                }
                umma_commit_pair(acc1_full);
            };
            int it = 0;
            if (first < pair_tiles) conv1(0);
            for (int T = first; T < pair_tiles; T += step, ++it) {
                const uint32_t par = static_cast<uint32_t>(it) & 1u;
                // ---- conv2(it): nine taps, each the t1 patch shifted by (r * 64 + s) pixels
                mbar_wait_wd(t1_ready, par);
                mbar_wait_wd(acc2_empty, par ^ 1u);
                tc_fence_after();
                {
                    const uint32_t patch = smem_u32(s_patch), w2 = smem_u32(s_w2);
                    const uint32_t d = tmem_base + COL_ACC2;
#pragma unroll
                    for (int r = 0; r < 3; ++r) {
#pragma unroll
                        for (int s = 0; s < 3; ++s) {
                            const uint64_t a = umma_desc_kmajor(patch + static_cast<uint32_t>((r * PITCH + s) * 128), 128);
                            const uint64_t b = umma_desc_kmajor(w2 + static_cast<uint32_t>((r * 3 + s) * 4096), 128);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_bf16_pair(d, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc64,
                                               static_cast<uint32_t>((r | s | k) != 0));
                        }
                    }
                }
                umma_commit_pair(t1_free);
                umma_commit_pair(acc2_full);
                // ---- conv1 of the NEXT tile keeps the tensor pipe busy while epilogue A produces t2(it)
                if (T + step < pair_tiles) conv1(it + 1);
                // ---- conv3(it)
                mbar_wait_wd(t2_ready, par);
                if (PROJ) mbar_wait_wd(xc_full, par);
                mbar_wait_wd(acc3_empty, par ^ 1u);
                tc_fence_after();
                {
                    const uint32_t d = tmem_base + COL_ACC3;
                    const uint64_t a = umma_desc_kmajor(smem_u32(s_t2), 128);
                    const uint64_t b = umma_desc_kmajor(smem_u32(s_w3), 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16_pair(d, a + static_cast<uint64_t>(2 * k), b + static_cast<uint64_t>(2 * k), idesc256,
                                       static_cast<uint32_t>(k != 0));
                    if (PROJ) {
                        const uint64_t a2 = umma_desc_kmajor(smem_u32(s_xc), 128);
                        const uint64_t b2 = umma_desc_kmajor(smem_u32(s_w3 + TILE16K), 128);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_pair(d, a2 + static_cast<uint64_t>(2 * k), b2 + static_cast<uint64_t>(2 * k), idesc256, 1u);
                    }
                }
                umma_commit_pair(t2_free);
                if (PROJ) umma_commit_pair(xc_free);
                umma_commit_pair(acc3_full);
            }
        }
    } else if (warp_idx == 2) {
        // ================================================================ loader: residual tiles / shortcut input
        if (elect_one()) {
            if (PROJ) {
                const uint32_t xc_full0 = mapa_shared(smem_u32(xc_full), 0);
                int it = 0;
                for (int T = first; T < pair_tiles; T += step, ++it) {
                    int img, row0;
                    tile_pos(T, img, row0);
                    mbar_wait_wd(xc_free, (static_cast<uint32_t>(it) & 1u) ^ 1u);
                    if (rank == 0) mbar_arrive_expect_tx(xc_full, 2u * TILE16K);
                    // the block input at this CTA's 2 x 56 output pixels, in the accumulator's row order (m = r*64 + c)
                    tma_load_4d_pair(&tmap_x, xc_full0, s_xc, 0, 0, row0, img, L2_EVICT_FIRST);
                }
            } else {
                uint32_t q = 0;
                for (int T = first; T < pair_tiles; T += step) {
                    int img, row0;
                    tile_pos(T, img, row0);
                    const int grow = (img * HW + row0) * HW;          // row of the [n*56*56, 256] view
                    for (int c = 0; c < 4; ++c, ++q) {
                        const uint32_t slot = q % 3u;
                        mbar_wait_wd(&res_empty[slot], ((q / 3u) & 1u) ^ 1u);
                        mbar_arrive_expect_tx(&res_full[slot], 2u * HW * 128u);
                        uint8_t* dst = s_io + slot * TILE16K;
                        tma_load_2d_hint(&tmap_res, &res_full[slot], dst, c * 64, grow, L2_EVICT_FIRST);
                        tma_load_2d_hint(&tmap_res, &res_full[slot], dst + 64 * 128, c * 64, grow + HW, L2_EVICT_FIRST);
                    }
                }
            }
        }
    } else if ((warp_idx >= 4 && warp_idx < 8) || warp_idx >= 12) {
        // ================================================================ epilogues A / C: acc1 -> t1 patch, acc2 -> t2
        // conv1 -> this epilogue -> conv2 is the serial chain of a tile, so the two 128-row halves of the patch are
        // converted by two warp groups at once: warps 4-7 take patch rows 0-1 (and then t2), warps 12-15 rows 2-3.
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int m = quad * 32 + lane;                              // accumulator row
        const int j = warp_idx >= 12 ? 1 : 0;                        // which half of the patch
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const uint32_t acc1_empty0 = mapa_shared(smem_u32(acc1_empty), 0), t1_ready0 = mapa_shared(smem_u32(t1_ready), 0);
        const uint32_t acc2_empty0 = mapa_shared(smem_u32(acc2_empty), 0), t2_ready0 = mapa_shared(smem_u32(t2_ready), 0);
        int it = 0;
        for (int T = first; T < pair_tiles; T += step, ++it) {
            const uint32_t par = static_cast<uint32_t>(it) & 1u;
            int img, row0;
            tile_pos(T, img, row0);
            // ---- t1 patch
            mbar_wait_wd(acc1_full, par);
            tc_fence_after();
            mbar_wait_wd(t1_free, par ^ 1u);                         // conv2 of the previous tile has read the patch
            {
                const int pr = j * 128 + m;                          // patch-linear pixel
                const int irow = row0 - 1 + (pr >> 6), icol = (pr & 63) - 1;
                const bool inside = irow >= 0 && irow < HW && icol >= 0 && icol < HW;
                uint4 packed[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + COL_ACC1 + static_cast<uint32_t>(j * MID + h * 32), v);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(args.b1 + h * 32);
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const float4 ba = __ldg(b4 + 2 * g), bb = __ldg(b4 + 2 * g + 1);
                        uint4 o;
                        o.x = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 0]) + ba.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 1]) + ba.y, 0.f));
                        o.y = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 2]) + ba.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 3]) + ba.w, 0.f));
                        o.z = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 4]) + bb.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 5]) + bb.y, 0.f));
                        o.w = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 6]) + bb.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 7]) + bb.w, 0.f));
                        if (!inside) o = make_uint4(0u, 0u, 0u, 0u);  // conv2's zero padding (relu(b1) != 0 there)
                        packed[h * 4 + g] = o;
                    }
                }
                const uint32_t row_addr = smem_u32(s_patch) + static_cast<uint32_t>(pr * 128);
                const uint32_t swz = static_cast<uint32_t>(pr & 7);
#pragma unroll
                for (int g = 0; g < 8; ++g) st_shared_v4(row_addr + ((static_cast<uint32_t>(g) ^ swz) << 4), packed[g]);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive_remote_release(t1_ready0);
                mbar_arrive_remote_release(acc1_empty0);
            }
            if (j != 0) continue;
            // ---- t2 tile (warps 4-7)
            mbar_wait_wd(acc2_full, par);
            tc_fence_after();
            {
                uint4 packed[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + COL_ACC2 + static_cast<uint32_t>(h * 32), v);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(args.b2 + h * 32);
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const float4 ba = __ldg(b4 + 2 * g), bb = __ldg(b4 + 2 * g + 1);
                        packed[h * 4 + g].x = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 0]) + ba.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 1]) + ba.y, 0.f));
                        packed[h * 4 + g].y = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 2]) + ba.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 3]) + ba.w, 0.f));
                        packed[h * 4 + g].z = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 4]) + bb.x, 0.f), fmaxf(__uint_as_float(v[8 * g + 5]) + bb.y, 0.f));
                        packed[h * 4 + g].w = pack_bf16x2(fmaxf(__uint_as_float(v[8 * g + 6]) + bb.z, 0.f), fmaxf(__uint_as_float(v[8 * g + 7]) + bb.w, 0.f));
                    }
                }
                mbar_wait_wd(t2_free, par ^ 1u);                     // conv3 of the previous tile has read t2
                const uint32_t row_addr = smem_u32(s_t2) + static_cast<uint32_t>(m * 128);
                const uint32_t swz = static_cast<uint32_t>(m & 7);
#pragma unroll
                for (int g = 0; g < 8; ++g) st_shared_v4(row_addr + ((static_cast<uint32_t>(g) ^ swz) << 4), packed[g]);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive_remote_release(t2_ready0);
                mbar_arrive_remote_release(acc2_empty0);
            }
        }
    } else if (warp_idx >= 8) {
        // ================================================================ epilogue B (warps 8-11): acc3 (+ residual) -> out
        const int quad = warp_idx & 3;
        const int lane = static_cast<int>(lane_id());
        const int m = quad * 32 + lane;
        const bool leader = warp_idx == 8 && lane == 0;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const uint32_t acc3_empty0 = mapa_shared(smem_u32(acc3_empty), 0);
        const uint32_t swz = static_cast<uint32_t>(m & 7);
        uint32_t q = 0;                                              // chunk sequence number over this CTA's walk
        int it = 0;
        for (int T = first; T < pair_tiles; T += step, ++it) {
            int img, row0;
            tile_pos(T, img, row0);
            const int grow = (img * HW + row0) * HW;
            mbar_wait_wd(acc3_full, static_cast<uint32_t>(it) & 1u);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 4; ++c, ++q) {
                // identity blocks: a ring of three tiles that are residual input and output staging at once (the loader
                // fills a tile, the accumulator is added in place, the tile is stored from where it lies).  Projection
                // block: two plain staging tiles used alternately.
                const uint32_t slot = PROJ ? (q & 1u) : q % 3u;
                uint8_t* tile = s_io + slot * TILE16K;
                const uint32_t row_addr = smem_u32(tile) + static_cast<uint32_t>(m * 128);
                if (!PROJ) {
                    if (leader && q > 0) {
                        // the previous chunk's store has read its tile: hand that tile back to the loader
                        tma_store_wait_read<0>();
                        mbar_arrive(&res_empty[(q - 1u) % 3u]);
                    }
                    mbar_wait_wd(&res_full[slot], (q / 3u) & 1u);
                }
                uint4 packed[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + COL_ACC3 + static_cast<uint32_t>(c * 64 + h * 32), v);
                    tmem_ld_wait();
                    const float4* b4 = reinterpret_cast<const float4*>(args.b3 + c * 64 + h * 32);
                    float f[32];
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        const float4 b = __ldg(b4 + g);
                        f[4 * g + 0] = __uint_as_float(v[4 * g + 0]) + b.x;
                        f[4 * g + 1] = __uint_as_float(v[4 * g + 1]) + b.y;
                        f[4 * g + 2] = __uint_as_float(v[4 * g + 2]) + b.z;
                        f[4 * g + 3] = __uint_as_float(v[4 * g + 3]) + b.w;
                    }
                    if (!PROJ) {
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const uint4 r = ld_shared_v4(row_addr + ((static_cast<uint32_t>(h * 4 + g) ^ swz) << 4));
                            f[8 * g + 0] += bf16_lo(r.x); f[8 * g + 1] += bf16_hi(r.x);
                            f[8 * g + 2] += bf16_lo(r.y); f[8 * g + 3] += bf16_hi(r.y);
                            f[8 * g + 4] += bf16_lo(r.z); f[8 * g + 5] += bf16_hi(r.z);
                            f[8 * g + 6] += bf16_lo(r.w); f[8 * g + 7] += bf16_hi(r.w);
                        }
                    }
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        packed[h * 4 + g].x = pack_bf16x2(fmaxf(f[8 * g + 0], 0.f), fmaxf(f[8 * g + 1], 0.f));
                        packed[h * 4 + g].y = pack_bf16x2(fmaxf(f[8 * g + 2], 0.f), fmaxf(f[8 * g + 3], 0.f));
                        packed[h * 4 + g].z = pack_bf16x2(fmaxf(f[8 * g + 4], 0.f), fmaxf(f[8 * g + 5], 0.f));
                        packed[h * 4 + g].w = pack_bf16x2(fmaxf(f[8 * g + 6], 0.f), fmaxf(f[8 * g + 7], 0.f));
                    }
                }
                if (c == 3) {
                    // last chunk of the tile: this warp's TMEM reads are done, hand the accumulator back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_remote_release(acc3_empty0);
                }
                if (PROJ) {
                    // the store issued from this tile two chunks ago must have been read (at most one may be pending)
                    if (leader) tma_store_wait_read<1>();
                    named_bar_sync(1, 128);
                }
#pragma unroll
                for (int g = 0; g < 8; ++g) st_shared_v4(row_addr + ((static_cast<uint32_t>(g) ^ swz) << 4), packed[g]);
                fence_proxy_async_smem();
                named_bar_sync(1, 128);
                if (leader) {
                    tma_store_2d(&tmap_out, tile, c * 64, grow);
                    tma_store_2d(&tmap_out, tile + 64 * 128, c * 64, grow + HW);
                    tma_store_commit();
                }
            }
        }
        if (leader) tma_store_wait_all<0>();
    }

    tc_fence_before();
    __syncwarp();
    cluster_sync_all();       // the leader's MMAs read the peer's shared memory and signal its barriers: leave together
    if (warp_idx == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 512);
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_bn_encode = nullptr;

bool bn_encode_init() {
    if (g_bn_encode) return true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn)
        return false;
    g_bn_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return true;
}

bool bn_encode_2d(CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows, uint32_t box_cols, uint32_t box_rows) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    return g_bn_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int KC1, bool PROJ>
cudaError_t bn_launch(const BneckPlan& p, const BnArgs& a, int grid, cudaStream_t s) {
    using C = BnCfg<KC1, PROJ>;
    auto kern = bottleneck56_kernel<KC1, PROJ>;
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) return e;
        attr_set[dev].store(true, std::memory_order_release);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(BN_THREADS);
    cfg.dynamicSmemBytes = C::SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, kern, p.tmap_x, p.tmap_w1, p.tmap_w2, p.tmap_w3, p.tmap_res, p.tmap_out, a);
}

}  // namespace

bool bneck56_supported(int in_h, int in_w, int cin, int mid, int cout, bool proj) {
    return in_h == HW && in_w == HW && mid == MID && cout == OUTC && ((proj && cin == 64) || (!proj && cin == OUTC));
}

bool bneck56_plan_init(BneckPlan& p, int cin, bool proj, const void* w1, const float* b1, const void* w2, const float* b2,
                       const void* w3, const float* b3, char* err, size_t err_len) {
    if (!bn_encode_init()) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled not available");
        return false;
    }
    p.cin = cin;
    p.proj = proj ? 1 : 0;
    p.b1 = b1; p.b2 = b2; p.b3 = b3;
    const uint64_t k3 = proj ? 128 : 64;
    if (!bn_encode_2d(&p.tmap_w1, w1, static_cast<uint64_t>(cin), MID, 64, 32) ||
        !bn_encode_2d(&p.tmap_w2, w2, 9 * MID, MID, 64, 32) || !bn_encode_2d(&p.tmap_w3, w3, k3, OUTC, 64, 128)) {
        snprintf(err, err_len, "bottleneck weight tensor maps failed");
        return false;
    }
    return true;
}

bool bneck56_plan_set(BneckPlan& p, const void* src, void* dst, int batch_cap, char* err, size_t err_len) {
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(p.cin), HW, HW, static_cast<cuuint64_t>(batch_cap)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(p.cin) * 2, static_cast<cuuint64_t>(p.cin) * 2 * HW,
                             static_cast<cuuint64_t>(p.cin) * 2 * HW * HW};
    cuuint32_t box[4] = {64, PITCH, 2, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (g_bn_encode(&p.tmap_x, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        snprintf(err, err_len, "bottleneck input tensor map failed");
        return false;
    }
    const uint64_t rows = static_cast<uint64_t>(batch_cap) * HW * HW;
    // identity blocks read the residual from the block input itself ([rows, 256] view of the same buffer)
    if (!bn_encode_2d(&p.tmap_out, dst, OUTC, rows, 64, HW) ||
        !bn_encode_2d(&p.tmap_res, p.proj ? dst : src, OUTC, rows, 64, HW)) {
        snprintf(err, err_len, "bottleneck output tensor maps failed");
        return false;
    }
    return true;
}

cudaError_t bneck56_launch(const BneckPlan& p, int n, int num_sms, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    BnArgs a;
    a.n_imgs = n;
    a.b1 = p.b1; a.b2 = p.b2; a.b3 = p.b3;
    const int pair_tiles = n * (HW / 4);
    const int pairs = num_sms / 2;
    const int grid = 2 * (pair_tiles < pairs ? pair_tiles : pairs);
    if (p.proj) return bn_launch<1, true>(p, a, grid, s);
    return bn_launch<4, false>(p, a, grid, s);
}

}  // namespace b2p
