// tcgen05 / TMEM implicit-GEMM kernels for the wide layers (FNET_MATH_TF32).
//
// D[128 x BN] (fp32, in tensor memory) += A[128 x 32] * B[BN x 32]^T per k-tile, issued by one
// thread as four tcgen05.mma.kind::tf32 (K = 8 each).  The operands are gathered by all 256
// threads straight from the NHWC tensors (implicit im2col, any channel count, any stride or
// padding -- FeatureNet's channel counts are rarely multiples of 4, which rules out TMA's
// 16-byte global strides) with cp.async into shared memory in the canonical layouts (K-major
// SWIZZLE_128B for im2col rows, MN-major SWIZZLE_128B_BASE32B for row-major weights / dz / x),
// three or four stages deep.  The tensor core truncates the fp32 bits to tf32.  Completion of the
// MMAs that read a stage is tracked with tcgen05.commit -> mbarrier, the accumulator never leaves
// TMEM until the epilogue (tcgen05.ld, bias + activation, store).
//
// The same "problem" types as the FFMA kernels describe forward, data gradient and filter
// gradient; split reductions write fp32 partial tiles to the caller's workspace and are summed
// in a fixed order (no atomics).
#include <stdlib.h>
#include "igemm_problems.cuh"

namespace fnet {

constexpr int TC_BM = 128;
constexpr int TC_BK = 32;        // fp32 elements per k-tile row: 128 bytes = one swizzle row
constexpr int TC_THREADS = 256;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t addr = smem_u32(bar), done = 0;
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

__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ float to_tf32(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}

// one lane of a converged warp (the same one every time)
__device__ __forceinline__ bool tc_elect() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// byte offset of element (row, kc) of a [rows x 32 fp32] K-major tile
__device__ __forceinline__ uint32_t tile_offset(int row, int kc, int layout) {
    if (layout == 0)   // SWIZZLE_128B: 8-row x 128-byte atoms, 16-byte chunks XOR-ed with the row
        return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((kc >> 2) ^ (row & 7)) << 4) | ((kc & 3) << 2)));
    // no swizzle: 8-row x 16-byte core matrices, K-adjacent core matrices 128 bytes apart
    return (uint32_t)((row >> 3) * 1024 + (kc >> 2) * 128 + (row & 7) * 16 + ((kc & 3) << 2));
}

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, int layout) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(layout == 0 ? 1 : (128 >> 4)) << 16;   // leading byte offset (ignored when swizzled)
    d |= (uint64_t)(1024 >> 4) << 32;                      // stride byte offset: next 8-row group
    d |= (uint64_t)1 << 46;                                // descriptor version for sm_100
    d |= (uint64_t)(layout == 0 ? 2 : 0) << 61;            // SWIZZLE_128B / SWIZZLE_NONE
    return d;
}

// MN-major operands (the contiguous global axis is the tile's row axis, not k: the weights [k][n] of the forward
// product, dz [pixel][n] and x [pixel][channel] of the filter gradient).  tcgen05 reads them in place when the
// instruction descriptor says so, which lets 16 contiguous bytes of global memory stay 16 contiguous bytes of
// shared memory instead of being scattered 4 bytes at a time into a K-major tile.  For 32-bit operands the only
// MN-major layout is SWIZZLE_128B with 32-byte atomicity: atoms of 4 k-rows x 128 bytes (32 elements along MN),
// the four 32-byte chunks of a row XOR-ed with k%4; atoms along MN are 512 bytes apart (leading byte offset),
// atoms along K `atoms`*512 bytes apart (stride byte offset); one K=8 MMA reads two atoms along K.
__device__ __forceinline__ uint32_t tile_offset_mn(int mn, int k, int atoms) {
    return (uint32_t)(((k >> 2) * atoms + (mn >> 5)) * 512 + (k & 3) * 128 + (((((mn >> 3) & 3) ^ (k & 3)) << 5) | ((mn & 7) << 2)));
}

__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t saddr, int atoms) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(512 >> 4) << 16;                       // leading byte offset: next atom along MN
    d |= (uint64_t)((atoms * 512) >> 4) << 32;             // stride byte offset: next atom along K
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;                                // SWIZZLE_128B_BASE32B
    return d;
}

// cp.async (LDGSTS): global -> shared without passing through registers; bytes beyond src_bytes are zero-filled,
// which is how padding, ragged tiles and out-of-range taps become zeros.
__device__ __forceinline__ void cp_async4(uint32_t dst, const float *src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const float *src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}

// X3 = 1: fp32 accuracy from tf32 MMAs.  Each landed stage is rewritten in place as hi = rna_tf32(v) and its
// tail lo = rna_tf32(v - hi) goes to one extra tile set behind the pipeline stages (same layout, so the same
// descriptors with another base address); a k-step is then three MMAs, hi*hi + lo*hi + hi*lo.  The dropped
// lo*lo term and the rounding of lo are ~2^-22 relative: the result sits with the FFMA kernels' against the
// float64 oracle.  The tensor pipe is idle most of the time in these gather-bound kernels, so the extra MMAs
// are free; the cost is one pass of the CTA over the stage in shared memory.
template <class P, int BN, int TC_STAGES, int X3>
__global__ void __launch_bounds__(TC_THREADS, 2) tc_igemm_kernel(P p, int M, int N, int Kg, int k_per_split,
                                                                 float *__restrict__ partial, int layout) {
    pdl_prologue();
    constexpr int B_BYTES = BN * TC_BK * 4;
    constexpr int STAGE_BYTES = TC_A_BYTES + B_BYTES;
    constexpr int B_PER = BN / 8;
    constexpr int TMEM_COLS = BN < 32 ? 32 : BN;          // BN is 32 / 64 / 128: a power of two
    // LAG = how many tiles ago the stage being refilled was multiplied: with 4 stages the refill waits on an
    // MMA issued two iterations earlier, so long reductions never stall on the tensor core's latency
    constexpr int LAG = TC_STAGES >= 4 ? 2 : 1;
    extern __shared__ uint8_t tc_smem_raw[];
    __shared__ uint64_t empty_bar[TC_STAGES];
    __shared__ uint64_t done_bar;
    __shared__ uint32_t tmem_holder;
    __shared__ typename P::Row rows[TC_BM];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const int m0 = blockIdx.x * TC_BM, n0 = blockIdx.y * BN;
    const int kbeg = blockIdx.z * k_per_split;
    const int kend = min(Kg, kbeg + k_per_split);
    const int n_tile = min(BN, ((N - n0) + 15) & ~15);     // UMMA N: multiple of 16

    for (int i = tid; i < TC_BM; i += TC_THREADS) p.make_row(m0 + i, M, rows[i]);
    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) mbar_init(&empty_bar[s], 1);
        mbar_init(&done_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_holder)),
                     "r"((uint32_t)TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_holder;

    // ---- operand loaders: every element (row, k) of a tile is copied global -> shared by cp.async straight
    // to tile_offset(row, k).  The lanes of a warp sit on the axis along which the operand is contiguous in
    // global memory (coalesced); where four consecutive k are four consecutive channels the A operand moves
    // 16 bytes per copy.  Up to TC_STAGES-1 tiles are in flight per CTA.
    // Tiles are issued in order, so the reduction-axis contexts are decoded once here and stepped by TC_BK per
    // tile: the loop below has no integer divisions (they, not the copies, used to bound the kernel).
    constexpr bool A_MN = !P::kAContigK, B_MN = !P::kBContigK;   // operand kept MN-major in shared memory
    constexpr int B_ATOMS = BN / 32;
    constexpr int B_CH = BN / 4;                            // 16-byte chunks per k-row of an MN-major B tile
    constexpr int B_KPW = 32 / B_CH;                        // k-rows one warp covers per pass (1, 2, 4)
    const bool vec4 = P::kAContigK && p.vec4_ok();
    const bool a_rows4 = A_MN && p.a_rows_vec4_ok();
    const bool b_rows4 = B_MN && p.b_rows_vec4_ok();
    constexpr int ACOLS = P::kAContigK ? 1 : 4;
    typename P::Col acol[ACOLS];
    const int a_koff = P::kAContigK ? (vec4 ? 4 * (lane & 7) : lane) : warp;
#pragma unroll
    for (int j = 0; j < ACOLS; ++j) p.make_col(kbeg + a_koff + 8 * j, kend, acol[j]);
    const int b_kl = warp * B_KPW + lane / B_CH, b_chunk = lane % B_CH;     // MN-major, 16-byte copies
    const bool b_k4 = P::kBContigK && p.b_k_vec4_ok();     // K-major B, 16 bytes = 4 consecutive k of one row
    typename P::BCursor bcur = P::kBContigK ? (b_k4 ? p.b_begin(kbeg + 4 * (lane & 7), n0 + warp * 4 + (lane >> 3))
                                                    : p.b_begin(kbeg + lane, n0 + warp))
                               : (b_rows4 ? p.b_begin(kbeg + b_kl, n0 + 4 * b_chunk) : p.b_begin(kbeg + warp, n0 + lane));
    const long long b_step_n = p.b_stride_n(), b_step_k = p.b_stride_k();
    int k_next = kbeg;
    // the row contexts a thread gathers for stay in registers (the copies below clobber memory, so shared-memory
    // reads would be repeated for every k-row, with bank conflicts on the strided ones)
    typename P::Row myrow[4];
    if (a_rows4) myrow[0] = rows[4 * lane];
    else if (A_MN) {
#pragma unroll
        for (int i = 0; i < 4; ++i) myrow[i] = rows[lane + 32 * i];
    } else if (vec4) {
#pragma unroll
        for (int i = 0; i < 4; ++i) myrow[i] = rows[warp * 4 + (lane >> 3) + 32 * i];
    }
    auto issue = [&](int stage) {
        const int k0 = k_next;
        const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + TC_A_BYTES;
        if (a_rows4) {                                      // MN-major A, lane = 4 consecutive rows (channels of one tap)
            const typename P::Row &row = myrow[0];
#pragma unroll
            for (int j = 0; j < ACOLS; ++j) {
                uint32_t dst = sa + tile_offset_mn(4 * lane, warp + 8 * j, 4);
                if (row.valid == 2) {                       // the all-ones row that yields the bias gradient
                    float one = acol[j].valid ? 1.f : 0.f;
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %2, %2};" ::"r"(dst), "f"(one), "f"(0.f) : "memory");
                } else {
                    const float *src = (acol[j].valid && row.valid) ? p.a_ptr(row, acol[j]) : nullptr;
                    cp_async16(dst, src ? src : p.any_ptr(), src ? 16u : 0u);
                }
            }
        } else if (A_MN) {
#pragma unroll
            for (int j = 0; j < ACOLS; ++j) {               // lanes along the rows: row = lane + 32*i, k = warp + 8*j
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    int r = lane + 32 * i;
                    uint32_t dst = sa + tile_offset_mn(r, warp + 8 * j, 4);
                    if (myrow[i].valid == 2) {
                        float one = acol[j].valid ? 1.f : 0.f;
                        asm volatile("st.shared.f32 [%0], %1;" ::"r"(dst), "f"(one) : "memory");
                    } else {
                        const float *src = (acol[j].valid && myrow[i].valid) ? p.a_ptr(myrow[i], acol[j]) : nullptr;
                        cp_async4(dst, src ? src : p.any_ptr(), src ? 4u : 0u);
                    }
                }
            }
        } else if (vec4) {                                  // lane = (row sub-index, 16-byte chunk); 4 passes of 32 rows
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int r = warp * 4 + (lane >> 3) + 32 * i;
                const float *src = (acol[0].valid && myrow[i].valid) ? p.a_ptr(myrow[i], acol[0]) : nullptr;
                cp_async16(sa + tile_offset(r, 4 * (lane & 7), layout), src ? src : p.any_ptr(), src ? 16u : 0u);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                int r = warp + 8 * i;
                const float *src = (acol[0].valid && rows[r].valid) ? p.a_ptr(rows[r], acol[0]) : nullptr;
                cp_async4(sa + tile_offset(r, lane, layout), src ? src : p.any_ptr(), src ? 4u : 0u);
            }
        }
        const float *bsrc = p.b_ptr(bcur);
        if (b_k4) {
            const bool k_ok = k0 + 4 * (lane & 7) < kend;
#pragma unroll
            for (int i = 0; i < BN / 32; ++i) {
                int nn = warp * 4 + (lane >> 3) + 32 * i;
                bool ok = k_ok && n0 + nn < N;
                cp_async16(sb + tile_offset(nn, 4 * (lane & 7), layout), ok ? bsrc + b_step_n * (32 * i) : p.any_ptr(), ok ? 16u : 0u);
            }
        } else if (P::kBContigK) {
            const bool k_ok = k0 + lane < kend;
#pragma unroll
            for (int i = 0; i < B_PER; ++i) {
                int nn = warp + 8 * i;
                bool ok = k_ok && n0 + nn < N;
                cp_async4(sb + tile_offset(nn, lane, layout), ok ? bsrc + b_step_n * (8 * i) : p.any_ptr(), ok ? 4u : 0u);
            }
        } else if (b_rows4) {                               // MN-major B: 16 bytes = 4 consecutive n of one k-row
            const bool n_ok = n0 + 4 * b_chunk < N;
#pragma unroll
            for (int ps = 0; ps < 4 / B_KPW; ++ps) {
                int kl = b_kl + ps * 8 * B_KPW;
                bool ok = n_ok && k0 + kl < kend;
                cp_async16(sb + tile_offset_mn(4 * b_chunk, kl, B_ATOMS), ok ? bsrc + b_step_k * (ps * 8 * B_KPW) : p.any_ptr(),
                           ok ? 16u : 0u);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int i = 0; i < BN / 32; ++i) {         // lanes along n: n = lane + 32*i, k = warp + 8*j
                    int nn = lane + 32 * i;
                    bool ok = k0 + warp + 8 * j < kend && n0 + nn < N;
                    cp_async4(sb + tile_offset_mn(nn, warp + 8 * j, B_ATOMS), ok ? bsrc + b_step_k * (8 * j) + 32 * i : p.any_ptr(),
                              ok ? 4u : 0u);
                }
        }
        k_next = k0 + TC_BK;
#pragma unroll
        for (int j = 0; j < ACOLS; ++j) p.next_col(k_next + a_koff + 8 * j, TC_BK, kend, acol[j]);
        p.b_next(bcur, TC_BK);
    };

    // instruction descriptor: D fp32, A/B tf32, both K-major, N = n_tile, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (A_MN ? (1u << 15) : 0u) | (B_MN ? (1u << 16) : 0u) |
                           ((uint32_t)(n_tile >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
    // descriptor start-address step per 8 k, in 16-byte units: 32 bytes along a swizzled K-major row, one atom
    // row (all MN atoms) of an MN-major tile
    const int kadv = layout == 0 ? 2 : 16;
    const int kadv_a = A_MN ? 4 * 64 : kadv, kadv_b = B_MN ? B_ATOMS * 64 : kadv;

    const int KT = (kend - kbeg + TC_BK - 1) / TC_BK;
#pragma unroll
    for (int t = 0; t < TC_STAGES - LAG; ++t) {            // prologue: tiles 0 .. STAGES-LAG-1 in flight
        if (t < KT) issue(t);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int kt = 0; kt < KT; ++kt) {
        const int s = kt % TC_STAGES;
        // refill the stage that tile kt-LAG used (its MMAs must have finished reading it) with tile kt+STAGES-LAG
        const int nt = kt + TC_STAGES - LAG;
        if (nt < KT) {
            const int ns = nt % TC_STAGES;
            if (kt >= LAG) mbar_wait(&empty_bar[ns], (uint32_t)(((kt - LAG) / TC_STAGES) & 1));
            issue(ns);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" ::"n"(TC_STAGES - LAG) : "memory"); // tile kt has landed (this thread's part)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");               // ... and is visible to the tensor core
        __syncthreads();
        if constexpr (X3) {
            // the MMAs of tile kt-1 read the (single) tail buffer: they must be done before it is rewritten
            if (kt >= 1) mbar_wait(&empty_bar[(kt - 1) % TC_STAGES], (uint32_t)(((kt - 1) / TC_STAGES) & 1));
            float4 *hi = reinterpret_cast<float4 *>(smem + s * STAGE_BYTES);
            float4 *lo = reinterpret_cast<float4 *>(smem + TC_STAGES * STAGE_BYTES);
#pragma unroll
            for (int c = tid; c < STAGE_BYTES / 16; c += TC_THREADS) {
                float4 v = hi[c], h, l;
                h.x = to_tf32(v.x); h.y = to_tf32(v.y); h.z = to_tf32(v.z); h.w = to_tf32(v.w);
                l.x = to_tf32(v.x - h.x); l.y = to_tf32(v.y - h.y); l.z = to_tf32(v.z - h.z); l.w = to_tf32(v.w - h.w);
                hi[c] = h;
                lo[c] = l;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
        }
        // one elected lane of the (converged) first warp issues: under `if (tid == 0)` the compiler wraps every
        // tcgen05.mma in a divergence-serialising loop and rebuilds its descriptors in vector registers (see gemm_tma.cu)
        if (warp == 0 && tc_elect()) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + TC_A_BYTES;
            uint64_t da = A_MN ? make_smem_desc_mn(sa, 4) : make_smem_desc(sa, layout);
            uint64_t db = B_MN ? make_smem_desc_mn(sb, B_ATOMS) : make_smem_desc(sb, layout);
            if constexpr (X3) {
                const uint32_t la = smem_base + TC_STAGES * STAGE_BYTES, lb = la + TC_A_BYTES;
                uint64_t dla = A_MN ? make_smem_desc_mn(la, 4) : make_smem_desc(la, layout);
                uint64_t dlb = B_MN ? make_smem_desc_mn(lb, B_ATOMS) : make_smem_desc(lb, layout);
#pragma unroll
                for (int j = 0; j < TC_BK / 8; ++j) {
                    tc_mma_tf32(tmem, dla + (uint64_t)(j * kadv_a), db + (uint64_t)(j * kadv_b), idesc, (kt > 0 || j > 0) ? 1u : 0u);
                    tc_mma_tf32(tmem, da + (uint64_t)(j * kadv_a), dlb + (uint64_t)(j * kadv_b), idesc, 1u);
                    tc_mma_tf32(tmem, da + (uint64_t)(j * kadv_a), db + (uint64_t)(j * kadv_b), idesc, 1u);   // the large term last
                }
            } else
#pragma unroll
            for (int j = 0; j < TC_BK / 8; ++j)
                tc_mma_tf32(tmem, da + (uint64_t)(j * kadv_a), db + (uint64_t)(j * kadv_b), idesc, (kt > 0 || j > 0) ? 1u : 0u);
            tc_commit(&empty_bar[s]);                                   // arrives when these MMAs have read the stage
            if (kt == KT - 1) tc_commit(&done_bar);
        }
    }

    // ---- epilogue.  Warp w owns TMEM lanes 32*(w%4)..+31 (its quadrant); 32-column chunks alternate
    // between the two warpgroups.  Each chunk is transposed through shared memory (the pipeline stages
    // are free by now) so that a warp stores 32 consecutive floats of one output row: 128-byte lines.
    if (KT > 0) {
        mbar_wait(&done_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    const int quad = warp & 3, half = warp >> 2;
    float *xpose = reinterpret_cast<float *>(smem) + warp * (32 * 33);
    for (int c0 = half * 32; c0 < n_tile; c0 += 64) {       // warp-uniform bounds
        uint32_t v[32];
        if (KT > 0) {
            uint32_t taddr = tmem + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0;
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
        for (int j = 0; j < 32; ++j) xpose[lane * 33 + j] = __uint_as_float(v[j]);   // thread = row
        __syncwarp();
        const int n = n0 + c0 + lane;                                                   // thread = column
        const int m_first = m0 + quad * 32, count = min(32, M - m_first);
        if (n < N && count > 0) {
            if (partial) {
                float *q = partial + ((long long)blockIdx.z * M + m_first) * N + n;
#pragma unroll 8
                for (int r = 0; r < count; ++r) q[(long long)r * N] = xpose[r * 33 + lane];
            } else {
                p.store_col(m_first, count, n, xpose + lane, 33);
            }
        }
        __syncwarp();
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)TMEM_COLS) : "memory");
}

struct TcPlan {
    int bn, splits, k_per_split, stages;
    size_t ws_bytes, smem_bytes;
};

static TcPlan tc_plan(long long M, long long N, long long Kg, bool x3 = false) {
    TcPlan t;
    t.bn = N <= 32 ? 32 : ((N <= 64 || x3) ? 64 : 128);     // x3 keeps a tail tile set: narrower tiles so two CTAs still fit
    long long tiles = ceil_div<long long>(M, TC_BM) * ceil_div<long long>(N, t.bn);
    // Two CTAs are resident per SM (shared memory: 3 stages of a 128-wide tile or 4 stages of a narrower one
    // are ~97 KB).  When the output has fewer tiles than that one wave, the reduction axis is split so that the
    // grid is exactly one wave: never a few CTAs more, which would cost a whole extra pass over the k-tiles.
    const long long slots = 2LL * kNumSMs;
    long long splits = 1;
    if (tiles < slots && Kg >= 16 * TC_BK) {
        splits = slots / tiles;
        long long most = Kg / (8 * TC_BK);        // at least 8 k-tiles per CTA
        if (splits > most) splits = most;
        if (splits < 1) splits = 1;
    }
    long long per = ceil_div<long long>(ceil_div<long long>(Kg, splits), TC_BK) * TC_BK;
    t.k_per_split = (int)per;
    t.splits = (int)ceil_div<long long>(Kg, per);
    t.ws_bytes = t.splits > 1 ? (size_t)t.splits * M * N * sizeof(float) : 0;
    t.stages = (per >= 8 * TC_BK && t.bn < 128 && !x3) ? 4 : 3;   // long reductions: deeper pipeline where it still fits twice
    t.smem_bytes = (size_t)(t.stages + (x3 ? 1 : 0)) * (TC_A_BYTES + t.bn * TC_BK * 4) + 1024;
    return t;
}

static int tc_layout_mode() {
    static const int mode = env_switch("FNET_TC_LAYOUT", 0);     // 0 = SWIZZLE_128B (default), 1 = no swizzle; for bring-up
    return mode;
}

template <class P, int BN, int STAGES, int X3>
static int tc_launch(const P &p, dim3 grid, size_t smem, int m, int n, int kg, int per, float *partial,
                     cudaStream_t stream, const char *what) {
    static std::atomic<uint64_t> configured{0};              // one mask per instantiation, one bit per device
    {
        cudaError_t e = allow_dynamic_smem(tc_igemm_kernel<P, BN, STAGES, X3>, (int)smem, configured);
        if (e != cudaSuccess) { set_error("%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e)); return FNET_ERR_CUDA; }
    }
    launch_pdl(tc_igemm_kernel<P, BN, STAGES, X3>, dim3(grid), dim3(TC_THREADS), smem, stream, p, m, n, kg, per, partial, tc_layout_mode());
    return check_launch(what);
}

template <class P>
static int tc_run(const P &p, int math, long long M, long long N, long long Kg, void *ws, size_t ws_bytes, cudaStream_t stream,
                  const char *what) {
    if (M <= 0 || N <= 0) return FNET_OK;
    const bool x3 = math == FNET_MATH_TF32X3;
    TcPlan plan = tc_plan(M, N, Kg, x3);
    float *partial = nullptr;
    if (plan.splits > 1) {
        if (!ws || ws_bytes < plan.ws_bytes) {
            set_error("%s: workspace of %zu bytes needed, %zu given", what, plan.ws_bytes, ws_bytes);
            return FNET_ERR_WORKSPACE;
        }
        partial = reinterpret_cast<float *>(ws);
    }
    dim3 grid((unsigned)ceil_div<long long>(M, TC_BM), (unsigned)ceil_div<long long>(N, plan.bn), plan.splits);
    int m = (int)M, n = (int)N, kg = (int)Kg, rc;
#define FNET_TC_LAUNCH(BN_, ST_, X3_) tc_launch<P, BN_, ST_, X3_>(p, grid, plan.smem_bytes, m, n, kg, plan.k_per_split, partial, stream, what)
    if (x3) rc = plan.bn == 32 ? FNET_TC_LAUNCH(32, 3, 1) : FNET_TC_LAUNCH(64, 3, 1);
    else if (plan.stages == 4) rc = plan.bn == 32 ? FNET_TC_LAUNCH(32, 4, 0) : FNET_TC_LAUNCH(64, 4, 0);
    else rc = plan.bn == 32 ? FNET_TC_LAUNCH(32, 3, 0) : (plan.bn == 64 ? FNET_TC_LAUNCH(64, 3, 0) : FNET_TC_LAUNCH(128, 3, 0));
#undef FNET_TC_LAUNCH
    if (rc != FNET_OK) return rc;
    if (plan.splits > 1) rc = launch_reduce_splits(p, m, n, plan.splits, partial, stream, what);
    return rc;
}

// The tensor-core tile is 128 rows x 32..256 columns x 32 deep; below these sizes it would be mostly
// padding and the FFMA kernels take over.
static bool tc_shape_ok(long long M, long long N, long long Kg) {
    return M >= 32 && N >= 1 && Kg >= 24 && M * N * Kg >= (1LL << 22) && M <= 0x7fffffffLL && Kg <= 0x7fffffffLL && N <= 65535LL * 64;
}

bool tc_supported(const ConvGeom &g, int math, int product) {
    if (math != FNET_MATH_TF32 && math != FNET_MATH_TF32X3) return false;   // bf16 operands: not built yet
    long long mo = (long long)g.B * g.Ho * g.Wo, mi = (long long)g.B * g.H * g.W, taps = (long long)g.kh * g.kw;
    if (product == 0) return tc_shape_ok(mo, g.K, taps * g.C);
    if (product == 1) return tc_shape_ok(mi, g.C, taps * g.K);
    return tc_shape_ok(taps * g.C + 1, g.K, mo);
}

size_t tc_conv_workspace(const ConvGeom &g, int math) {
    if (math != FNET_MATH_TF32 && math != FNET_MATH_TF32X3) return 0;
    long long mo = (long long)g.B * g.Ho * g.Wo, mi = (long long)g.B * g.H * g.W, taps = (long long)g.kh * g.kw;
    size_t need = 0;
    for (int x3 = 0; x3 < 2; ++x3) {                        // the caller's buffer serves either tensor-core mode
        auto upto = [&](size_t v) { need = v > need ? v : need; };
        if (tc_supported(g, math, 0)) upto(tc_plan(mo, g.K, taps * g.C, x3).ws_bytes);
        if (tc_supported(g, math, 1)) upto(tc_plan(mi, g.C, taps * g.K, x3).ws_bytes);
        if (tc_supported(g, math, 2)) {                     // with / without the bias row
            upto(tc_plan(taps * g.C + 1, g.K, mo, x3).ws_bytes);
            upto(tc_plan(taps * g.C, g.K, mo, x3).ws_bytes);
        }
    }
    return need;
}

int tc_conv_fwd(const ConvGeom &g, int math, int act, const float *x, const float *w, const float *bias, float *y,
                void *ws, size_t ws_bytes, cudaStream_t stream) {
    FwdProblem p{g, x, w, bias, y, act == FNET_ACT_SOFTMAX ? FNET_ACT_NONE : act};
    return tc_run(p, math, (long long)g.B * g.Ho * g.Wo, g.K, (long long)g.kh * g.kw * g.C, ws, ws_bytes, stream, "conv2d_fwd[tf32]");
}

int tc_conv_dgrad(const ConvGeom &g, int math, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws,
                  size_t ws_bytes, cudaStream_t stream) {
    DgradProblem p{g, dz, w, dx, ep.accumulate, ep.act, ep.y};
    return tc_run(p, math, (long long)g.B * g.H * g.W, g.C, (long long)g.kh * g.kw * g.K, ws, ws_bytes, stream, "conv2d_bwd_data[tf32]");
}

int tc_conv_wgrad(const ConvGeom &g, int math, const float *x, const float *dz, float *dw, float *dbias, void *ws,
                  size_t ws_bytes, cudaStream_t stream) {
    int taps_c = g.kh * g.kw * g.C;
    WgradProblem p{g, x, dz, dw, dbias, taps_c};
    return tc_run(p, math, (long long)taps_c + (dbias ? 1 : 0), g.K, (long long)g.B * g.Ho * g.Wo, ws, ws_bytes, stream,
                  "conv2d_bwd_filter[tf32]");
}

}  // namespace fnet
