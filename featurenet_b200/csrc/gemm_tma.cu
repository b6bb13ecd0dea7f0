// TMA-fed, warp-specialised, persistent tcgen05 GEMM for the layers whose operands are boxes: Dense and 1x1 stride-1
// convolutions (plain GEMMs: the "Reg_" channel aligners, the x8 / x0.25 1x1 expansions, Dense on rank-4 tensors)
// and k x k stride-1 convolutions (TgConv: one shifted 4-D box of the NHWC tensor per tap and 32-channel chunk).
// No gather instructions are involved: the TMA engine copies the boxes straight into the swizzled shared-memory
// layouts tcgen05 reads, and zero-fills whatever lies outside the tensor (ragged edges, convolution padding).
//
//   warp 0 (one lane)  producer: waits for a free stage, arms its mbarrier with the byte count and issues
//                      cp.async.bulk.tensor.{2d,4d} for A (128 x 32) and B (BN x 32);
//   warp 1 (one lane)  issues four tcgen05.mma.kind::tf32 per stage (twelve in TF32x3 mode) into one of two TMEM
//                      accumulators and releases the stage (tcgen05.commit -> mbarrier);
//   warps 2..9         epilogue (tma_gemm_kernel; 2..5 in the filter gradient): tcgen05.ld of their TMEM lane quadrant,
//                      bias + activation, stores -- while the other accumulator is already being filled for the next tile;
//   the next 8 warps   TF32x3 only: split every landed stage into tf32 heads (in place) and tails (twin buffer).
//
// One CTA per SM walks the output tiles (persistent), so barrier setup, TMEM allocation and descriptor
// prefetch are paid once per SM instead of once per tile, and the epilogue of tile i overlaps the main loop
// of tile i+1.  Forward: A = x [M][K] (K-major), B = w [K][N] (MN-major, SWIZZLE_128B_ATOM_32B boxes of
// 32 x 32).  Data gradient: A = dz [M][N] (K-major), B = w [C][N] (K-major).  The filter gradient has its own
// kernel below (both operands MN-major, reduction over the pixels, split over CTAs).  In plain TF32 mode the tensor
// core truncates fp32 to tf32.
#include <cuda.h>
#include "igemm_problems.cuh"

namespace fnet {

constexpr int TG_BM = 128, TG_BK = 32, TG_STAGES = 4, TG_THREADS = 192;
constexpr int TG_A_BYTES = TG_BM * TG_BK * 4;
// TF32x3 (fp32 accuracy): eight more warps rewrite every landed stage as hi = rna_tf32(v) in place and write the
// tail lo = v - hi to a second copy of the stage with the same layout, then hand the stage to the MMA
// warp, which issues lo*hi + hi*lo + hi*hi per k-step.  The layers this kernel serves are HBM-bound in plain tf32,
// so the two extra MMAs and the pass over shared memory hide behind the loads.
constexpr int TG_SPLITTERS = 256;                           // threads of the TF32x3 splitter warps (warps 6..13)
constexpr int TG_THREADS_X3 = TG_THREADS + TG_SPLITTERS;
// tma_gemm_kernel: eight epilogue warps (two per TMEM lane quadrant, alternating 32-column groups) -- its epilogue works from
// registers alone, and on wide outputs it is what the kernel waits for
constexpr int TG_GEMM_THREADS = 64 + 256, TG_GEMM_THREADS_X3 = TG_GEMM_THREADS + TG_SPLITTERS;
constexpr int TG_SMEM_LIMIT = 209 * 1024;     // + 17 KB of static shared memory = the 227 KB a CTA may have
constexpr int TG_SMEM_LIMIT_GEMM = 226 * 1024;       // tma_gemm_kernel keeps only its barriers in static shared memory
constexpr int TG_STAGES_MAX = 8;                       // weights-resident mode: the stages carry A alone and are small

// hi = x rounded to tf32 (nearest, ties away: what cvt.rna.tf32.f32 gives, in two integer instructions), lo = x - hi
// rounded the same way: left to the tensor core, which truncates, the tails would carry a one-sided error of 2^-22 of
// x each, and a 4800-term reduction (5x5 x 192 channels) came out at 3.8e-5 of the float64 result instead of < 2e-5.
__device__ __forceinline__ void tg_hi_lo(uint32_t x, uint32_t &hi, uint32_t &lo) {
    hi = (x + 0x1000u) & 0xffffe000u;
    lo = (__float_as_uint(__uint_as_float(x) - __uint_as_float(hi)) + 0x1000u) & 0xffffe000u;
}
// hi in place, lo to the twin buffer `twin_offset` bytes further, for the 16-byte chunks [first, first + count) of a
// stage; TG_SPLITTERS threads.  Four independent loads are in flight per thread before the first store (the loop is
// latency-bound otherwise).  ones_box >= 0 (filter gradient, MN-major A): the chunk that holds row `32 * ones_box +
// 4 * ones_sub` of every k becomes {1, 0, 0, 0} -- the all-ones row whose products are the column sums of dz, i.e.
// the bias gradient, in a row of the A tile that the layer does not use.
__device__ __forceinline__ void tg_split_chunks(uint8_t *stage, int first, int count, int twin_offset, int t, int ones_box,
                                                int ones_sub) {
    uint4 *hi = reinterpret_cast<uint4 *>(stage) + first;
    uint4 *lo = reinterpret_cast<uint4 *>(stage + twin_offset) + first;
    for (int c = t; c < count; c += 4 * TG_SPLITTERS) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c + u * TG_SPLITTERS < count) v[u] = hi[c + u * TG_SPLITTERS];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = c + u * TG_SPLITTERS;
            if (i < count) {
                // MN-major box of 32 x 32: k-row of 128 bytes = 8 chunks, 32-byte pairs XORed with k % 4
                if (ones_box >= 0 && (i >> 8) == ones_box && ((i & 7) ^ (((i >> 3) & 3) << 1)) == ones_sub)
                    v[u] = make_uint4(0x3f800000u, 0u, 0u, 0u);
                uint4 h, l;
                tg_hi_lo(v[u].x, h.x, l.x);
                tg_hi_lo(v[u].y, h.y, l.y);
                tg_hi_lo(v[u].z, h.z, l.z);
                tg_hi_lo(v[u].w, h.w, l.w);
                hi[i] = h;
                lo[i] = l;
            }
        }
    }
}
__device__ __forceinline__ void tg_split_done() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> the tensor core's reads
}

__device__ __forceinline__ uint32_t tg_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tg_bar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tg_smem(bar)), "r"(count));
}
__device__ __forceinline__ void tg_bar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t addr = tg_smem(bar), done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tg_bar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tg_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tg_bar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tg_smem(bar)) : "memory");
}
__device__ __forceinline__ void tg_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tg_smem(bar)) : "memory");
}
// one lane of a converged warp (the same one every time)
__device__ __forceinline__ bool tg_elect() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tg_tma_2d(uint32_t dst, const CUtensorMap *map, uint64_t *bar, int c_inner, int c_outer) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tg_smem(bar)), "r"(c_inner), "r"(c_outer)
                 : "memory");
}
__device__ __forceinline__ void tg_tma_4d(uint32_t dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(tg_smem(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}

// k x k stride-1 convolution as taps x ceil(C / 32) k-tiles of the same kernel: the A operand of k-tile (tap, chunk)
// is the box {32 channels, W, rows, images} of the NHWC activation tensor, shifted by the tap.  Whatever the shift
// moves outside the image the TMA unit zero-fills, which IS the convolution's padding; the box lands as 128
// consecutive pixels x 32 channels, the K-major tile tcgen05 reads.  taps == 0: plain GEMM (2-D boxes).
struct TgConv {
    int taps, kw, chunks;       // k-tiles = taps * chunks
    int tap_rows;               // rows of the weight matrix [taps * tap_rows][K] per tap (= C)
    int H, W;                   // the OUTPUT grid of this product: a tile is 128 consecutive pixels of it
    int off_h, off_w, step_h, step_w;   // input row of tap r at output row h: h + off_h + r * step_h
};
__device__ __forceinline__ void tg_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}

// K-major SWIZZLE_128B tile (rows of 128 bytes, 8-row atoms of 1024 bytes): what a {32, rows} TMA box produces
__device__ __forceinline__ uint64_t tg_desc_k(uint32_t saddr) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// MN-major SWIZZLE_128B_BASE32B tile made of {32 (mn), 32 (k)} TMA boxes of 4096 bytes: inside a box the 4-row
// atoms follow each other (stride byte offset 512), the next 32 columns are the next box (leading byte offset 4096)
__device__ __forceinline__ uint64_t tg_desc_mn(uint32_t saddr) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(4096 >> 4) << 16;
    d |= (uint64_t)(512 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;
    return d;
}

template <class P, bool X3>
__global__ void __launch_bounds__(X3 ? TG_GEMM_THREADS_X3 : TG_GEMM_THREADS, 1)
tma_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, P p, int M, int N, int Kg,
                int BN, int n_tiles, int total_tiles, int tmem_cols, int stages, TgConv cv, int resident) {
    pdl_prologue();
    constexpr bool B_MN = !P::kBContigK;
    extern __shared__ uint8_t tg_smem_raw[];
    __shared__ uint64_t full_bar[TG_STAGES_MAX], empty_bar[TG_STAGES_MAX], split_bar[TG_STAGES_MAX], acc_full[2], acc_empty[2],
        bres_full, bres_split;
    __shared__ uint32_t tmem_holder;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(tg_smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = tg_smem(smem);
    // Streaming mode: a stage = A tile + B tile (+ their tails).  Weights-resident mode (plain GEMMs whose whole B
    // column block fits: K <= 64 into 256 columns, 24 / 192 channels into anything): the CTA keeps ONE column block, loads
    // (and splits) its KT B tiles once, and the stages carry A alone -- a third of the bytes, so more of them are in
    // flight; the per-tile re-load and re-split of the same weights is gone.
    const int b_bytes = BN * TG_BK * 4;
    const int operand_bytes = resident ? TG_A_BYTES : TG_A_BYTES + b_bytes;        // what the TMA unit lands per stage
    const int stage_bytes = X3 ? 2 * operand_bytes : operand_bytes;                // + the tails, same layout
    const int KT = cv.taps ? cv.taps * cv.chunks : (Kg + TG_BK - 1) / TG_BK;
    const int bres_stage = X3 ? 2 * b_bytes : b_bytes;                              // resident B tile kt: heads, then tails
    uint8_t *stage0 = smem + (resident ? KT * bres_stage : 0);                      // the resident tiles come first
    const uint32_t stage0_addr = tg_smem(stage0);
    // this CTA's tiles: streaming = every gridDim-th (m, n) tile; resident = every (gridDim / n_tiles)-th row block of one column block
    const int tile_first = resident ? blockIdx.x / n_tiles : blockIdx.x, tile_step = resident ? gridDim.x / n_tiles : gridDim.x;
    const int tile_end = resident ? total_tiles / n_tiles : total_tiles, cta_n0 = (blockIdx.x % n_tiles) * BN;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < stages; ++s) {
                tg_bar_init(&full_bar[s], 1);
                tg_bar_init(&empty_bar[s], 1);
                tg_bar_init(&split_bar[s], TG_SPLITTERS / 32);
            }
            for (int a = 0; a < 2; ++a) { tg_bar_init(&acc_full[a], 1); tg_bar_init(&acc_empty[a], (TG_GEMM_THREADS - 64) / 32); }
            tg_bar_init(&bres_full, 1);
            tg_bar_init(&bres_split, TG_SPLITTERS / 32);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tg_smem(&tmem_holder)),
                     "r"((uint32_t)tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_holder;
    const uint32_t acc_stride = (uint32_t)tmem_cols / 2;      // column offset of the second accumulator

    if (warp == 0 && resident && tile_first < tile_end) {
        // the column block's weights, once: all k-tiles on one barrier
        if (lane == 0) tg_bar_expect_tx(&bres_full, (uint32_t)(KT * b_bytes));
        __syncwarp();
        for (int kt = 0; kt < KT; ++kt) {
            const uint32_t sb = smem_base + kt * bres_stage;
            if (B_MN) {
                if (lane >= 1 && lane <= BN / 32) tg_tma_2d(sb + (lane - 1) * 4096, &map_b, &bres_full, cta_n0 + 32 * (lane - 1), kt * TG_BK);
            } else if (lane == 1) {
                tg_tma_2d(sb, &map_b, &bres_full, kt * TG_BK, cta_n0);
            }
        }
        __syncwarp();
    }
    if (warp == 0) {
        // every lane issues at most one box per stage: a 4 KB box costs as much issue time as a 16 KB one, so the
        // MN-major operand (32 x 32 boxes) is spread over the lanes instead of being issued serially by one
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = tile_first; tile < tile_end; tile += tile_step) {
            const int m0 = (resident ? tile : tile / n_tiles) * TG_BM, n0 = resident ? cta_n0 : (tile % n_tiles) * BN;
            int img = 0, row0 = 0;                            // convolution: the tile's first pixel is (img, row0, 0)
            if (cv.taps) {
                img = m0 / (cv.H * cv.W);
                row0 = (m0 - img * cv.H * cv.W) / cv.W;
            }
            int tap = 0, chunk = 0, tr = 0, ts = 0;           // k-tile = (tap (tr, ts), 32-channel chunk)
            for (int kt = 0; kt < KT; ++kt) {
                tg_bar_wait(&empty_bar[stage], phase ^ 1);
                const uint32_t sa = stage0_addr + stage * stage_bytes, sb = sa + TG_A_BYTES;
                // first reduction index of this k-tile in the weight matrix / in the plain A matrix
                const int kb = cv.taps ? tap * cv.tap_rows + chunk * TG_BK : kt * TG_BK;
                if (lane == 0) {
                    tg_bar_expect_tx(&full_bar[stage], (uint32_t)operand_bytes);
                    if (cv.taps)
                        tg_tma_4d(sa, &map_a, &full_bar[stage], chunk * TG_BK, cv.off_w + ts * cv.step_w,
                                  row0 + cv.off_h + tr * cv.step_h, img);
                    else
                        tg_tma_2d(sa, &map_a, &full_bar[stage], kt * TG_BK, m0);
                }
                if (resident) {
                    // the weights are already here
                } else if (B_MN) {                            // forward: w [taps * C][K], boxes of 32 k-rows x 32 columns
                    if (lane >= 1 && lane <= BN / 32)
                        tg_tma_2d(sb + (lane - 1) * 4096, &map_b, &full_bar[stage], n0 + 32 * (lane - 1), kb);
                } else if (lane == 1) {                       // data gradient: w [taps * C][K] read as rows c, k along K
                    if (cv.taps) tg_tma_2d(sb, &map_b, &full_bar[stage], chunk * TG_BK, tap * cv.tap_rows + n0);
                    else tg_tma_2d(sb, &map_b, &full_bar[stage], kt * TG_BK, n0);
                }
                if (cv.taps && ++chunk == cv.chunks) {
                    chunk = 0;
                    ++tap;
                    if (++ts == cv.kw) { ts = 0; ++tr; }
                }
                __syncwarp();
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (X3 && warp >= TG_GEMM_THREADS / 32) {
        int stage = 0;
        uint32_t phase = 0;
        if (resident && tile_first < tile_end) {              // the resident weights: split once
            tg_bar_wait(&bres_full, 0);
            for (int kt = 0; kt < KT; ++kt)
                tg_split_chunks(smem + kt * bres_stage, 0, b_bytes / 16, b_bytes, threadIdx.x - TG_GEMM_THREADS, -1, 0);
            tg_split_done();
            __syncwarp();
            if (lane == 0) tg_bar_arrive(&bres_split);
        }
        for (int tile = tile_first; tile < tile_end; tile += tile_step) {
            for (int kt = 0; kt < KT; ++kt) {
                tg_bar_wait(&full_bar[stage], phase);
                tg_split_chunks(stage0 + stage * stage_bytes, 0, operand_bytes / 16, operand_bytes, threadIdx.x - TG_GEMM_THREADS, -1, 0);
                tg_split_done();
                __syncwarp();
                if (lane == 0) tg_bar_arrive(&split_bar[stage]);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // The whole warp walks the loop (warp-uniform control flow: addresses and descriptors stay in uniform registers)
        // and one elected lane issues the MMAs.  Issued from inside `if (lane == 0)`, every tcgen05.mma was wrapped in
        // a divergence-serialising loop with its descriptors rebuilt in vector registers and moved over (R2UR): about
        // twenty dependent instructions of a lone warp per MMA -- more than a 128 x 64 x 8 MMA takes to execute.
        // instruction descriptor: D fp32, A / B tf32, A K-major, B as the problem says, N = BN, M = 128
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (B_MN ? (1u << 16) : 0u) | ((uint32_t)(BN >> 3) << 17) |
                               ((uint32_t)(TG_BM >> 4) << 24);
        const bool leader = tg_elect();
        // descriptors of stage 0; a stage further is stage_bytes / 16 more in the start-address field, the tails operand_bytes / 16
        // resident mode: B tile kt lives at kt * bres_stage from the start of shared memory, its tails b_bytes further
        const uint32_t b_at = resident ? smem_base : stage0_addr + TG_A_BYTES;
        const uint64_t da0 = tg_desc_k(stage0_addr), db0 = B_MN ? tg_desc_mn(b_at) : tg_desc_k(b_at);
        const uint32_t stage_step = (uint32_t)stage_bytes >> 4, tail = (uint32_t)operand_bytes >> 4;
        const uint32_t b_step = resident ? (uint32_t)bres_stage >> 4 : stage_step, b_tail = resident ? (uint32_t)b_bytes >> 4 : tail;
        constexpr int badv = B_MN ? 64 : 2;                  // 16-byte units per 8 k: two 4-row atoms / 32 bytes along the row
        int stage = 0, it = 0;
        uint32_t phase = 0;
        if (resident && tile_first < tile_end) {
            tg_bar_wait(X3 ? &bres_split : &bres_full, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        for (int tile = tile_first; tile < tile_end; tile += tile_step, ++it) {
            const int acc = it & 1;
            tg_bar_wait(&acc_empty[acc], (uint32_t)(((it >> 1) & 1) ^ 1));      // the epilogue has drained this accumulator
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t d = tmem + acc * acc_stride;
            for (int kt = 0; kt < KT; ++kt) {
                tg_bar_wait(X3 ? &split_bar[stage] : &full_bar[stage], phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (leader) {
                    const uint64_t da = da0 + (uint64_t)((uint32_t)stage * stage_step);
                    const uint64_t db = db0 + (uint64_t)((uint32_t)(resident ? kt : stage) * b_step);
                    if (X3) {
#pragma unroll
                        for (int j = 0; j < TG_BK / 8; ++j) {
                            tg_mma(d, da + (uint64_t)(tail + 2 * j), db + (uint64_t)(badv * j), idesc, (kt > 0 || j > 0) ? 1u : 0u);
                            tg_mma(d, da + (uint64_t)(2 * j), db + (uint64_t)(b_tail + badv * j), idesc, 1u);
                            tg_mma(d, da + (uint64_t)(2 * j), db + (uint64_t)(badv * j), idesc, 1u);      // the large term last
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < TG_BK / 8; ++j)
                            tg_mma(d, da + (uint64_t)(2 * j), db + (uint64_t)(badv * j), idesc, (kt > 0 || j > 0) ? 1u : 0u);
                    }
                    tg_commit(&empty_bar[stage]);
                }
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
            if (leader) tg_commit(&acc_full[acc]);
        }
        __syncwarp();
    } else {
        const int quad = warp & 3;                            // the TMEM lane quadrant this warp may read
        int it = 0;
        for (int tile = tile_first; tile < tile_end; tile += tile_step, ++it) {
            const int acc = it & 1;
            const int m0 = (resident ? tile : tile / n_tiles) * TG_BM, n0 = resident ? cta_n0 : (tile % n_tiles) * BN;
            tg_bar_wait(&acc_full[acc], (uint32_t)((it >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // tcgen05.ld.16x256b: 16 lanes x 8 columns per repeat, laid out like an MMA accumulator fragment (thread t: rows
            // t / 4 and t / 4 + 8, columns 2 (t % 4) + {0, 1}).  The four lanes of a row then store one whole 32-byte sector
            // and a warp instruction eight of them, straight from the registers: the transpose through shared memory (a
            // quarter of the kernel's shared-memory traffic on wide outputs) is gone.
            const int m_first = m0 + quad * 32;
            for (int c0 = ((warp - 2) >> 2) * 32; c0 < BN; c0 += 64) {
                const int cols = N - (n0 + c0 + 2 * (lane & 3));          // columns that exist from this thread's first one on
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t v[16];
                    const uint32_t taddr = tmem + ((uint32_t)(quad * 32 + half * 16) << 16) + acc * acc_stride + (uint32_t)c0;
                    asm volatile(
                        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    const int m = m_first + half * 16 + (lane >> 2);
                    if (cols > 0 && m < M) p.store_frag(m, true, m + 8 < M, n0 + c0 + 2 * (lane & 3), cols, v);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) tg_bar_arrive(&acc_empty[acc]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)tmem_cols) : "memory");
}

// Filter gradient of a plain GEMM layer: dw[C][N] = x^T[C][pixels] * dz[pixels][N].  Both operands are MN-major
// (their contiguous axis is the tile's row axis), the reduction runs over the pixels and is long, the output is
// small: one (output tile, slice of the pixel axis) per CTA, partial tiles to the workspace, fixed-order reduce.
// The bias gradient is the column sum of dz, i.e. a second accumulator fed by an all-ones A tile (an extra MMA per
// k-step on a tensor pipe that is otherwise idle; the first row of that accumulator is the answer).
template <bool X3>
__global__ void __launch_bounds__(X3 ? TG_THREADS_X3 : TG_THREADS, 1)
tma_wgrad_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, WgradProblem p,
                 int M /* C (+1 bias row) */, int N, int pixels, int BN, int n_tiles, int tiles, int kt_per_split, int stages,
                 int tmem_cols, float *__restrict__ partial, TgConv cv) {
    pdl_prologue();
    extern __shared__ uint8_t tg_smem_raw[];
    __shared__ uint64_t full_bar[TG_STAGES], empty_bar[TG_STAGES], split_bar[TG_STAGES], acc_full;
    __shared__ uint32_t tmem_holder;
    __shared__ float xpose[4][32 * 33];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(tg_smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = tg_smem(smem);
    const int b_bytes = BN * TG_BK * 4, operand_bytes = TG_A_BYTES + b_bytes;
    const int stage_bytes = X3 ? 2 * operand_bytes : operand_bytes;
    const int tile = blockIdx.x % tiles, split = blockIdx.x / tiles;
    const int m0 = (tile / n_tiles) * TG_BM, n0 = (tile % n_tiles) * BN;
    const int KT_all = (pixels + TG_BK - 1) / TG_BK;
    const int kt_begin = split * kt_per_split, kt_end = min(KT_all, kt_begin + kt_per_split);
    // TF32x3: when the layer leaves a row of the A tile unused, the splitter makes it the all-ones row and the bias
    // gradient is row taps_c of the main accumulator; otherwise (and in plain tf32, which has no pass over the stage)
    // a second accumulator is fed by an all-ones tile
    // k x k convolution (cv.taps > 0): the rows of the A tile are (tap, 32-channel chunk, channel), each 32-row block
    // one shifted box of x; the reduction runs over 32-pixel boxes of the output grid.  Its bias gradient always
    // takes the all-ones tile (the padded row layout has no fixed free row).
    const int fold_row = (X3 && !cv.taps && p.dbias != nullptr && p.taps_c >= m0 && p.taps_c < m0 + TG_BM) ? p.taps_c - m0 : -1;
    const bool with_bias = p.dbias != nullptr && m0 == 0 && !(X3 && !cv.taps && p.taps_c % TG_BM != 0);
    float *ones = reinterpret_cast<float *>(smem + stages * stage_bytes);

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < stages; ++s) {
                tg_bar_init(&full_bar[s], 1);
                tg_bar_init(&empty_bar[s], 1);
                tg_bar_init(&split_bar[s], TG_SPLITTERS / 32);
            }
            tg_bar_init(&acc_full, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tg_smem(&tmem_holder)),
                     "r"((uint32_t)tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else if (warp >= 2 && warp < 6 && with_bias) {
        for (int i = threadIdx.x - 64; i < TG_A_BYTES / 4; i += 128) ones[i] = 1.f;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_holder;
    const uint32_t bias_col = (uint32_t)tmem_cols / 2;        // the column-sum accumulator
    const int m_rows = cv.taps ? cv.taps * cv.chunks * 32 : p.taps_c;               // rows of the (padded) A matrix
    const int rows_here = min(TG_BM, m_rows - m0) + (fold_row >= 0 ? 1 : 0);        // rows of this tile that are stored
    const int a_boxes = (rows_here + 31) / 32, b_boxes = (min(BN, N - n0) + 31) / 32;

    if (warp == 0) {
        int stage = 0;                                        // lanes 0..3: the A boxes, lanes 4..: the B boxes
        uint32_t phase = 0;
        for (int kt = kt_begin; kt < kt_end; ++kt) {
            tg_bar_wait(&empty_bar[stage], phase ^ 1);
            const uint32_t sa = smem_base + stage * stage_bytes, sb = sa + TG_A_BYTES;
            if (lane == 0) tg_bar_expect_tx(&full_bar[stage], (uint32_t)operand_bytes);
            if (lane < TG_BM / 32) {
                if (cv.taps) {
                    // row block -> (tap, chunk); k-tile -> the 32 output pixels starting at (img, row, col)
                    const int rb = m0 / 32 + lane, tap = rb / cv.chunks, chunk = rb - tap * cv.chunks;
                    const int tr = tap / cv.kw, ts = tap - tr * cv.kw;
                    const int pix = kt * TG_BK, img = pix / (cv.H * cv.W), rem = pix - img * cv.H * cv.W;
                    const int row = rem / cv.W, col = rem - row * cv.W;
                    // a row block beyond the last tap reads outside the tensor (channel coordinate past C): zeros
                    tg_tma_4d(sa + lane * 4096, &map_a, &full_bar[stage], tap < cv.taps ? chunk * 32 : (1 << 20),
                              col + cv.off_w + ts * cv.step_w, row + cv.off_h + tr * cv.step_h, img);
                } else {
                    tg_tma_2d(sa + lane * 4096, &map_a, &full_bar[stage], m0 + 32 * lane, kt * TG_BK);
                }
            } else if (lane - TG_BM / 32 < BN / 32)
                tg_tma_2d(sb + (lane - TG_BM / 32) * 4096, &map_b, &full_bar[stage], n0 + 32 * (lane - TG_BM / 32), kt * TG_BK);
            __syncwarp();
            if (++stage == stages) { stage = 0; phase ^= 1; }
        }
    } else if (X3 && warp >= 6) {
        int stage = 0;
        uint32_t phase = 0;
        for (int kt = kt_begin; kt < kt_end; ++kt) {
            tg_bar_wait(&full_bar[stage], phase);
            // only the boxes that hold rows / columns of the layer (the TMA unit zero-fills the rest; their tails are
            // never written, which can only reach output rows and columns that are not stored)
            uint8_t *st = smem + stage * stage_bytes;
            const int t = threadIdx.x - TG_THREADS;
            tg_split_chunks(st, 0, a_boxes * 256, operand_bytes, t, fold_row >= 0 ? fold_row >> 5 : -1, (fold_row & 31) >> 2);
            tg_split_chunks(st, TG_A_BYTES / 16, b_boxes * 256, operand_bytes, t, -1, 0);
            tg_split_done();
            __syncwarp();
            if (lane == 0) tg_bar_arrive(&split_bar[stage]);
            if (++stage == stages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        {
            // all lanes walk the loop, one elected lane issues (see tma_gemm_kernel)
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(BN >> 3) << 17) |
                                   ((uint32_t)(TG_BM >> 4) << 24);
            const uint32_t idesc_ones = idesc & ~(1u << 15);  // the all-ones tile is read K-major (any layout of ones is ones)
            const uint64_t d_ones = tg_desc_k(tg_smem(ones));
            const bool leader = tg_elect();
            const uint64_t da0 = tg_desc_mn(smem_base), db0 = tg_desc_mn(smem_base + TG_A_BYTES);
            const uint32_t stage_step = (uint32_t)stage_bytes >> 4, tail = (uint32_t)operand_bytes >> 4;
            int stage = 0;
            uint32_t phase = 0;
            for (int kt = kt_begin; kt < kt_end; ++kt) {
                tg_bar_wait(X3 ? &split_bar[stage] : &full_bar[stage], phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (leader) {
                    const uint64_t da = da0 + (uint64_t)((uint32_t)stage * stage_step), db = db0 + (uint64_t)((uint32_t)stage * stage_step);
#pragma unroll
                    for (int j = 0; j < TG_BK / 8; ++j) {
                        const uint32_t accumulate = (kt > kt_begin || j > 0) ? 1u : 0u;
                        if (X3) {
                            tg_mma(tmem, da + (uint64_t)(tail + 64 * j), db + (uint64_t)(64 * j), idesc, accumulate);
                            tg_mma(tmem, da + (uint64_t)(64 * j), db + (uint64_t)(tail + 64 * j), idesc, 1u);
                            tg_mma(tmem, da + (uint64_t)(64 * j), db + (uint64_t)(64 * j), idesc, 1u);
                            if (with_bias) {
                                tg_mma(tmem + bias_col, d_ones + (uint64_t)(2 * j), db + (uint64_t)(tail + 64 * j), idesc_ones, accumulate);
                                tg_mma(tmem + bias_col, d_ones + (uint64_t)(2 * j), db + (uint64_t)(64 * j), idesc_ones, 1u);
                            }
                        } else {
                            tg_mma(tmem, da + (uint64_t)(64 * j), db + (uint64_t)(64 * j), idesc, accumulate);
                            if (with_bias) tg_mma(tmem + bias_col, d_ones + (uint64_t)(2 * j), db + (uint64_t)(64 * j), idesc_ones, accumulate);
                        }
                    }
                    tg_commit(&empty_bar[stage]);
                }
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
            if (leader) tg_commit(&acc_full);
            __syncwarp();
        }
    } else {
        const int quad = warp & 3;
        float *xp = xpose[warp - 2];
        const bool any = kt_end > kt_begin;
        if (any) {
            tg_bar_wait(&acc_full, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        const int taps_c = p.taps_c;
        int m_first = m0 + quad * 32, count = min(32, m0 + rows_here - m_first);
        if (cv.taps) {                                        // padded row block -> rows tap * C + chunk * 32 ... of dw
            const int rb = m_first / 32, tap = rb / cv.chunks, chunk = rb - tap * cv.chunks;
            count = tap < cv.taps ? min(32, cv.tap_rows - chunk * 32) : 0;
            m_first = tap * cv.tap_rows + chunk * 32;
        }
        for (int pass = 0; pass < (with_bias && quad == 0 ? 2 : 1); ++pass) {
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t v[32];
                if (any) {
                    uint32_t taddr = tmem + ((uint32_t)(quad * 32) << 16) + (pass ? bias_col : 0u) + (uint32_t)c0;
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = 0u;
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) xp[lane * 33 + j] = __uint_as_float(v[j]);
                __syncwarp();
                const int n = n0 + c0 + lane;
                if (n < N) {
                    // pass 0: rows of dw; pass 1: row 0 of the column-sum accumulator is the bias gradient (row taps_c)
                    const int first = pass ? taps_c : m_first, rows = pass ? 1 : count;
                    for (int r = 0; r < rows; ++r) {
                        float val = xp[r * 33 + lane];
                        if (partial) partial[((long long)split * M + first + r) * N + n] = val;
                        else p.store(first + r, n, val);
                    }
                }
                __syncwarp();
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)tmem_cols) : "memory");
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn lookup_encode_tiled() {
    void *sym = nullptr;
    cudaDriverEntryPointQueryResult status;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &status) == cudaSuccess &&
        status == cudaDriverEntryPointSuccess)
        return reinterpret_cast<EncodeTiledFn>(sym);
    return nullptr;
}

// looked up once; a function-local static is initialised exactly once even when several host threads arrive together
// (a hand-rolled `tried` flag let a second thread see "looked up" before the pointer was stored: it then sized its
// workspace for the non-TMA kernels and failed at launch, when the TMA kernel was chosen after all)
static EncodeTiledFn encode_tiled() {
    static const EncodeTiledFn fn = lookup_encode_tiled();
    return fn;
}

// a row-major fp32 matrix [rows][cols] seen through boxes of box_rows x box_cols
static bool make_map(CUtensorMap *map, const float *base, long long rows, long long cols, int box_rows, int box_cols,
                     CUtensorMapSwizzle swizzle) {
    EncodeTiledFn fn = encode_tiled();
    if (!fn) return false;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t elem[2] = {1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, elem,
              CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

typedef EncodeTiledFn EncodeFn;

// an NHWC fp32 tensor [B][H][W][C] seen through boxes of {32 channels, W_box, H_box, B_box}
static bool make_map_nhwc(CUtensorMap *map, const float *base, long long B, long long H, long long W, long long C, int box_w,
                          int box_h, int box_b, bool mn_major = false) {
    EncodeFn fn = encode_tiled();
    if (!fn) return false;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
    cuuint32_t box[4] = {(cuuint32_t)TG_BK, (cuuint32_t)box_w, (cuuint32_t)box_h, (cuuint32_t)box_b};
    cuuint32_t elem[4] = {1, 1, 1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float *>(base), dims, strides, box, elem,
              CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// 128 consecutive pixels of an [H][W] grid as a box {W, rows, images}: W must divide 128 and the rows of a tile must
// either divide H (tiles inside one image) or be whole images
static bool tile_box(int H, int W, int &box_h, int &box_b) {
    if (W < 1 || W > TG_BM || TG_BM % W) return false;
    const int rows = TG_BM / W;
    if (rows <= H) {
        if (H % rows) return false;
        box_h = rows;
        box_b = 1;
    } else {
        if (rows % H) return false;
        box_h = H;
        box_b = rows / H;
    }
    return true;
}

static bool is_plain_gemm(const ConvGeom &g) {
    return g.kh == 1 && g.kw == 1 && g.sh == 1 && g.sw == 1 && g.pt == 0 && g.pl == 0 && g.Ho == g.H && g.Wo == g.W;
}

// width of the output tile; wide layers are cut into equal tiles (260 -> 2 x 160 rather than 256 + 32).  The TF32x3
// filter gradient stops at 128 columns: its stages are twice as large and three of them must fit (measured: 153 ->
// 137 us on 64 -> 512; the forward product is better off with 256 columns and two stages, 115 vs 126 us)
static int tg_bn(long long n, bool narrow = false) {
    const long long cap = narrow ? 128 : 256;
    const long long tiles = (n + cap - 1) / cap;
    long long bn = (((n + tiles - 1) / tiles + 31) / 32) * 32;
    return (int)(bn > cap ? cap : bn);
}

struct TgWgradPlan {
    int bn, n_tiles, tiles, splits, kt_per_split, stages, tmem_cols;
    size_t smem, ws_bytes;
};

// forward / data gradient: when the grid would leave SMs idle, narrower tiles make more of them (A is re-read from L2)
static int tg_bn_for(long long n, long long m_tiles) {
    int bn = tg_bn(n);
    while (2 * m_tiles * ceil_div<long long>(n, bn) <= kNumSMs && bn >= 64 && bn % 64 == 0) bn /= 2;    // < half a wave
    return bn;
}

// pipeline depth that fits: TF32x3 stages carry a second copy (the tails) of both operand tiles
static int tg_stages(int bn, bool x3, int most, size_t extra) {
    size_t stage = (size_t)(x3 ? 2 : 1) * (TG_A_BYTES + bn * TG_BK * 4);
    int n = (int)((TG_SMEM_LIMIT - 1024 - extra) / stage);
    return n > most ? most : n;
}

// 32 consecutive pixels of an [H][W] grid as a box {W_box, rows, images} (the k-tile of the filter gradient)
static bool pixel_box(int H, int W, int &box_w, int &box_h, int &box_b) {
    if (W >= TG_BK) {
        if (W % TG_BK) return false;
        box_w = TG_BK; box_h = 1; box_b = 1;
        return true;
    }
    if (W < 1 || TG_BK % W) return false;
    const int rows = TG_BK / W;
    box_w = W;
    if (rows <= H) {
        if (H % rows) return false;
        box_h = rows; box_b = 1;
    } else {
        if (rows % H) return false;
        box_h = H; box_b = rows / H;
    }
    return true;
}

static TgWgradPlan tg_wgrad_plan(const ConvGeom &g, bool x3) {
    TgWgradPlan t;
    const bool conv = !is_plain_gemm(g);
    const long long pixels = (long long)g.B * g.Ho * g.Wo, KT = ceil_div<long long>(pixels, TG_BK);
    // rows of the A matrix: C, or taps x (C rounded up to whole 32-channel chunks)
    const long long m_rows = conv ? (long long)g.kh * g.kw * ceil_div<long long>(g.C, 32) * 32 : g.C;
    t.bn = tg_bn(g.K, x3);
    t.n_tiles = (int)ceil_div<long long>(g.K, t.bn);
    t.tiles = (int)ceil_div<long long>(m_rows, TG_BM) * t.n_tiles;
    long long splits = kNumSMs / t.tiles, most = KT / 8;     // one CTA per SM, at least 8 k-tiles each
    if (splits > most) splits = most;
    if (splits < 1) splits = 1;
    t.kt_per_split = (int)ceil_div<long long>(KT, splits);
    t.splits = (int)ceil_div<long long>(KT, t.kt_per_split);
    t.stages = tg_stages(t.bn, x3, 3, TG_A_BYTES);
    t.tmem_cols = 64;
    while (t.tmem_cols < 2 * t.bn) t.tmem_cols *= 2;
    t.smem = (size_t)t.stages * (x3 ? 2 : 1) * (TG_A_BYTES + t.bn * TG_BK * 4) + TG_A_BYTES + 1024;
    t.ws_bytes = t.splits > 1 ? (size_t)t.splits * ((size_t)g.kh * g.kw * g.C + 1) * g.K * sizeof(float) : 0;
    return t;
}

// product: 0 forward (N = g.K, reduction g.C), 1 data gradient (N = g.C, reduction g.K), 2 filter gradient
bool tma_gemm_supported(const ConvGeom &g, int math, int product) {
    static const int mode = env_switch("FNET_TMA", 1);      // FNET_TMA=0 disables (A/B measurements)
    if (!mode || (math != FNET_MATH_TF32 && math != FNET_MATH_TF32X3) || product > 2 || !encode_tiled()) return false;
    if (!is_plain_gemm(g)) {
        // k x k convolutions, stride 1: forward and data gradient through shifted 4-D boxes (TgConv)
        static const int conv_mode = env_switch("FNET_TMA_CONV", 1);   // FNET_TMA_CONV=0 disables (A/B measurements)
        int bw, bh, bb;
        if (!conv_mode || g.sh != 1 || g.sw != 1 || (g.C & 3) || (g.K & 3)) return false;
        if (product == 2) {
            const long long pixels = (long long)g.B * g.Ho * g.Wo;
            return pixel_box(g.Ho, g.Wo, bw, bh, bb) && g.K >= 16 && pixels >= 8192 && pixels <= 0x7fffffffLL &&
                   (long long)g.kh * g.kw <= 64 && g.C <= 1024;
        }
        if (product == 0 ? !tile_box(g.Ho, g.Wo, bh, bb) : !tile_box(g.H, g.W, bh, bb)) return false;
        const long long Mc = product == 0 ? (long long)g.B * g.Ho * g.Wo : (long long)g.B * g.H * g.W;
        const long long Nc = product == 0 ? g.K : g.C;
        if (Nc < 16 || Mc > 0x7fffffffLL || (long long)g.kh * g.kw > 64) return false;
        static const long long min_tiles = [] {             // FNET_TMA_MIN_TILES: experiments with partial waves
            const char *e = getenv("FNET_TMA_MIN_TILES");
            return e ? atoll(e) : 32LL;
        }();
        const long long m_tiles = ceil_div<long long>(Mc, TG_BM);
        const long long tiles = m_tiles * ceil_div<long long>(Nc, tg_bn_for(Nc, m_tiles));
        const long long kt = (long long)g.kh * g.kw * ceil_div<long long>(product == 0 ? g.C : g.K, TG_BK);
        // a partial wave still wins when every tile has a long reduction to stream (no split reduction here): measured
        // 96 -> 192 5x5 on 8x8 images 184 / 220 -> 119 / 129 us, 128 -> 128 3x3 72 / 73 -> 53 / 43 us (fwd / dgrad)
        return tiles >= kNumSMs || (tiles >= min_tiles && kt >= 16);
    }
    long long M = (long long)g.B * g.H * g.W;
    if (product == 2)
        return !(g.C & 3) && !(g.K & 3) && g.C >= 16 && g.K >= 16 && M >= 8192 && M <= 0x7fffffffLL &&
               (long long)g.C * g.K * M >= (1LL << 24);
    long long N = product == 0 ? g.K : g.C, Kg = product == 0 ? g.C : g.K;
    if ((g.C & 3) || (g.K & 3) || N < 16 || Kg < 8 || M > 0x7fffffffLL) return false;      // 16-byte global strides
    const long long m_tiles = ceil_div<long long>(M, TG_BM);
    long long tiles = m_tiles * ceil_div<long long>(N, tg_bn_for(N, m_tiles));
    return tiles >= kNumSMs;                                 // no split reduction here: the cp.async kernels do that
}

// a: the A operand, [M][Kg] for a plain GEMM, or (cv.taps > 0) the NHWC tensor [src_b][src_h][src_w][src_c] the
// shifted boxes are cut from; b: the weights, [Kg][N] (b_mn) or [b_rows][Kg]
template <class P, bool X3>
static int tma_gemm_run(const P &p, const float *a, long long M, long long Kg, const float *b, long long N, bool b_mn,
                        cudaStream_t stream, const char *what, const TgConv &cv = TgConv{}, int src_b = 0, int src_h = 0,
                        int src_w = 0, int src_c = 0, long long b_rows = 0) {
    if ((reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15) ||
        ((reinterpret_cast<uintptr_t>(p.out_ptr()) | reinterpret_cast<uintptr_t>(p.out_aux_ptr())) & 7)) {
        set_error("%s: operands must be 16-byte aligned (results 8-byte) for the TMA path", what);
        return FNET_ERR_ARG;
    }
    const int BN = tg_bn_for(N, ceil_div<long long>(M, TG_BM));
    CUtensorMap map_a, map_b;
    bool ok;
    if (cv.taps) {
        int box_h = 0, box_b = 0;
        ok = tile_box(cv.H, cv.W, box_h, box_b) && make_map_nhwc(&map_a, a, src_b, src_h, src_w, src_c, cv.W, box_h, box_b);
    } else {
        ok = make_map(&map_a, a, M, Kg, TG_BM, TG_BK, CU_TENSOR_MAP_SWIZZLE_128B);
    }
    if (b_mn) ok = ok && make_map(&map_b, b, Kg, N, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);    // w [K][N]
    else ok = ok && make_map(&map_b, b, b_rows ? b_rows : N, Kg, BN, TG_BK, CU_TENSOR_MAP_SWIZZLE_128B);   // w [C][Kg]
    if (!ok) { set_error("%s: cuTensorMapEncodeTiled failed", what); return FNET_ERR_CUDA; }
    const int n_tiles = (int)ceil_div<long long>(N, BN);
    const long long m_tiles = ceil_div<long long>(M, TG_BM), total = m_tiles * n_tiles;
    int tmem_cols = 64;
    while (tmem_cols < 2 * BN) tmem_cols *= 2;
    // weights-resident mode (see the kernel): the whole column block of B stays in shared memory, the stages carry A alone
    static const int resident_mode = env_switch("FNET_TMA_RESIDENT", 1);       // FNET_TMA_RESIDENT=0: A/B measurements
    const long long KT = ceil_div<long long>(Kg, TG_BK), parts = X3 ? 2 : 1;
    const long long b_region = KT * parts * BN * TG_BK * 4, a_stage = parts * TG_A_BYTES;
    const bool resident = resident_mode && !cv.taps && n_tiles <= kNumSMs && m_tiles * n_tiles >= kNumSMs &&
                          b_region + 3 * a_stage + 1024 <= TG_SMEM_LIMIT_GEMM;
    int stages;
    size_t smem;
    unsigned grid;
    if (resident) {
        long long most = (TG_SMEM_LIMIT_GEMM - 1024 - b_region) / a_stage;
        stages = (int)(most > TG_STAGES_MAX ? TG_STAGES_MAX : most);
        smem = (size_t)(b_region + stages * a_stage + 1024);
        grid = (unsigned)((kNumSMs / n_tiles) * n_tiles);        // every CTA owns one column block
    } else {
        stages = tg_stages(BN, X3, TG_STAGES, 0);
        smem = (size_t)stages * (X3 ? 2 : 1) * (TG_A_BYTES + BN * TG_BK * 4) + 1024;
        grid = (unsigned)(total < kNumSMs ? total : kNumSMs);
    }
    static std::atomic<uint64_t> configured{0};              // one mask per instantiation, one bit per device
    {
        cudaError_t e = allow_dynamic_smem(tma_gemm_kernel<P, X3>, TG_SMEM_LIMIT_GEMM, configured);
        if (e != cudaSuccess) { set_error("%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e)); return FNET_ERR_CUDA; }
    }
    launch_pdl(tma_gemm_kernel<P, X3>, dim3(grid), dim3(X3 ? TG_GEMM_THREADS_X3 : TG_GEMM_THREADS), smem, stream, map_a, map_b, p, (int)M, (int)N, (int)Kg, BN,
                                                              n_tiles, (int)total, tmem_cols, stages, cv, resident ? 1 : 0);
    return check_launch(what);
}

int tma_gemm_fwd(const ConvGeom &g, int math, int act, const float *x, const float *w, const float *bias, float *y,
                 cudaStream_t stream) {
    FwdProblem p{g, x, w, bias, y, act == FNET_ACT_SOFTMAX ? FNET_ACT_NONE : act};
    const bool x3 = math == FNET_MATH_TF32X3;
    if (!is_plain_gemm(g)) {
        const long long M = (long long)g.B * g.Ho * g.Wo, Kg = (long long)g.kh * g.kw * g.C;
        TgConv cv{g.kh * g.kw, g.kw, (int)ceil_div<long long>(g.C, TG_BK), g.C, g.Ho, g.Wo, -g.pt, -g.pl, g.dh, g.dw};
        const char *what = x3 ? "conv2d_fwd[tf32x3,tma,taps]" : "conv2d_fwd[tf32,tma,taps]";
        if (x3) return tma_gemm_run<FwdProblem, true>(p, x, M, Kg, w, g.K, true, stream, what, cv, g.B, g.H, g.W, g.C);
        return tma_gemm_run<FwdProblem, false>(p, x, M, Kg, w, g.K, true, stream, what, cv, g.B, g.H, g.W, g.C);
    }
    const long long M = (long long)g.B * g.H * g.W;
    if (x3) return tma_gemm_run<FwdProblem, true>(p, x, M, g.C, w, g.K, true, stream, "conv2d_fwd[tf32x3,tma]");
    return tma_gemm_run<FwdProblem, false>(p, x, M, g.C, w, g.K, true, stream, "conv2d_fwd[tf32,tma]");
}

int tma_gemm_dgrad(const ConvGeom &g, int math, const float *dz, const float *w, float *dx, DxEpilogue ep, cudaStream_t stream) {
    DgradProblem p{g, dz, w, dx, ep.accumulate, ep.act, ep.y};
    if (!is_plain_gemm(g)) {
        // dx[h][w] = sum over taps of dz[h + pt - r*dh][w + pl - s*dw] . w[r][s]^T: the same shifted boxes, over dz
        const long long M = (long long)g.B * g.H * g.W, rows = (long long)g.kh * g.kw * g.C;
        TgConv cv{g.kh * g.kw, g.kw, (int)ceil_div<long long>(g.K, TG_BK), g.C, g.H, g.W, g.pt, g.pl, -g.dh, -g.dw};
        const bool x3 = math == FNET_MATH_TF32X3;
        const char *what = x3 ? "conv2d_bwd_data[tf32x3,tma,taps]" : "conv2d_bwd_data[tf32,tma,taps]";
        if (x3) return tma_gemm_run<DgradProblem, true>(p, dz, M, g.K, w, g.C, false, stream, what, cv, g.B, g.Ho, g.Wo, g.K, rows);
        return tma_gemm_run<DgradProblem, false>(p, dz, M, g.K, w, g.C, false, stream, what, cv, g.B, g.Ho, g.Wo, g.K, rows);
    }
    const long long M = (long long)g.B * g.H * g.W;
    if (math == FNET_MATH_TF32X3)
        return tma_gemm_run<DgradProblem, true>(p, dz, M, g.K, w, g.C, false, stream, "conv2d_bwd_data[tf32x3,tma]");
    return tma_gemm_run<DgradProblem, false>(p, dz, M, g.K, w, g.C, false, stream, "conv2d_bwd_data[tf32,tma]");
}

size_t tma_gemm_workspace(const ConvGeom &g, int math) {
    return tma_gemm_supported(g, math, 2) ? tg_wgrad_plan(g, math == FNET_MATH_TF32X3).ws_bytes : 0;
}

int tma_gemm_wgrad(const ConvGeom &g, int math, const float *x, const float *dz, float *dw, float *dbias, void *ws,
                   size_t ws_bytes, cudaStream_t stream) {
    const bool x3 = math == FNET_MATH_TF32X3;
    const char *what = x3 ? "conv2d_bwd_filter[tf32x3,tma]" : "conv2d_bwd_filter[tf32,tma]";
    if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(dz) & 15)) {
        set_error("%s: operands must be 16-byte aligned for the TMA path", what);
        return FNET_ERR_ARG;
    }
    TgWgradPlan t = tg_wgrad_plan(g, x3);
    float *partial = nullptr;
    if (t.splits > 1) {
        if (!ws || ws_bytes < t.ws_bytes) {
            set_error("%s: workspace of %zu bytes needed, %zu given", what, t.ws_bytes, ws_bytes);
            return FNET_ERR_WORKSPACE;
        }
        partial = reinterpret_cast<float *>(ws);
    }
    const bool conv = !is_plain_gemm(g);
    const long long pixels = (long long)g.B * g.Ho * g.Wo;
    CUtensorMap map_a, map_b;
    TgConv cv{};
    bool ok;
    if (conv) {
        int bw = 0, bh = 0, bb = 0;
        cv = TgConv{g.kh * g.kw, g.kw, (int)ceil_div<long long>(g.C, 32), g.C, g.Ho, g.Wo, -g.pt, -g.pl, g.dh, g.dw};
        ok = pixel_box(g.Ho, g.Wo, bw, bh, bb) && make_map_nhwc(&map_a, x, g.B, g.H, g.W, g.C, bw, bh, bb, true);
    } else {
        ok = make_map(&map_a, x, pixels, g.C, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    }
    if (!ok || !make_map(&map_b, dz, pixels, g.K, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) {
        set_error("%s: cuTensorMapEncodeTiled failed", what);
        return FNET_ERR_CUDA;
    }
    static std::atomic<uint64_t> configured_plain{0}, configured_x3{0};
    {
        cudaError_t e = allow_dynamic_smem(tma_wgrad_kernel<false>, TG_SMEM_LIMIT, configured_plain);
        if (e == cudaSuccess) e = allow_dynamic_smem(tma_wgrad_kernel<true>, TG_SMEM_LIMIT, configured_x3);
        if (e != cudaSuccess) { set_error("%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e)); return FNET_ERR_CUDA; }
    }
    WgradProblem p{g, x, dz, dw, dbias, g.kh * g.kw * g.C};
    const int M = g.kh * g.kw * g.C + (dbias ? 1 : 0);
    const dim3 grid((unsigned)(t.tiles * t.splits));
    if (x3)
        launch_pdl(tma_wgrad_kernel<true>, grid, dim3(TG_THREADS_X3), t.smem, stream, map_a, map_b, p, M, g.K, (int)pixels, t.bn, t.n_tiles, t.tiles,
                                                                   t.kt_per_split, t.stages, t.tmem_cols, partial, cv);
    else
        launch_pdl(tma_wgrad_kernel<false>, grid, dim3(TG_THREADS), t.smem, stream, map_a, map_b, p, M, g.K, (int)pixels, t.bn, t.n_tiles, t.tiles,
                                                                 t.kt_per_split, t.stages, t.tmem_cols, partial, cv);
    int rc = check_launch(what);
    if (rc == FNET_OK && t.splits > 1) rc = simt_reduce_wgrad_partials(g, dw, dbias, M, t.splits, partial, stream);
    return rc;
}

}  // namespace fnet
