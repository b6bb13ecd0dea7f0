// k x k stride-1 convolutions (forward and data gradient) on tcgen05 with the input tile RESIDENT in shared memory.
//
// An implicit GEMM that materialises one A tile per (tap, channel chunk) -- what gemm_tma.cu's shifted boxes do -- moves
// every input element from L2 into shared memory kh*kw times (and, in TF32x3 mode, splits it into head and tail kh*kw
// times).  Here a CTA loads the pixels its 128 outputs touch ONCE per channel chunk, as a "slab", and every tap reads
// the same slab through a descriptor whose start address is shifted by the tap:
//
//   * positions.  The output grid is embedded in a padded grid of Hs = Ho + kh - 1 rows and Wp = Wo + kw - 1 columns
//     per image, linearised over the whole batch: q = (img * Hs + hh) * Wp + ww.  Output (hh, ww) is real when
//     hh < Ho and ww < Wo; the input pixel AT position q is (hh - pad_top, ww - pad_left), zero when outside the
//     image.  With this embedding tap (r, s) of output q is the input at position q + r * Wp + s -- one constant shift
//     for the whole tile, rows, image borders and padding included.  A tile is 128 consecutive positions; the few
//     positions that are not real outputs (1 - Ho*Wo / (Hs*Wp) of them) are computed and dropped.
//   * slab layout.  No swizzle, K-major: [16-byte channel chunk][position][4 floats].  A core matrix (8 rows x 16
//     bytes) is 8 consecutive positions = 128 contiguous bytes, the next 8 rows follow at +128 (stride byte offset),
//     the next 4 channels at +pitch (leading byte offset).  Shifting the A operand by n positions is start + 16 n.
//   * TF32x3.  The loader warps split each value into head and tail as they store it (two slabs), so the split costs
//     no extra pass over shared memory, and it happens once per input element instead of once per tap.
//   * weights.  A small kernel rewrites w into the exact shared-memory image of every (chunk, tap) B tile (K-major,
//     no swizzle, heads then tails; transposed for the forward product, mirrored for the data gradient); the main
//     kernel fetches a tile with ONE cp.async.bulk.
//
//   warp 0        B producer (cp.async.bulk + mbarrier), one lane
//   warp 1        tcgen05.mma issue (one lane), TMEM allocation; two accumulators so the epilogue of tile i overlaps tile i+1
//   warps 2..9    epilogue: tcgen05.ld in the MMA fragment layout, bias / activation / dx epilogue, 32-byte-sector stores
//   warps 10..15  slab loaders: 16-byte global loads, (split,) conflict-free 16-byte shared stores
//
// The filter gradient keeps its own kernels (gemm_tma.cu / igemm_tc.cu): its reduction runs over the pixels.
#include "igemm.cuh"

namespace fnet {

constexpr int DC_BM = 128;
constexpr int DC_EPI_WARPS = 8, DC_LOAD_WARPS = 6;     // 16 warps: four per scheduler, 128 registers each
constexpr int DC_THREADS = 32 * (2 + DC_EPI_WARPS + DC_LOAD_WARPS);
constexpr int DC_LOADERS = 32 * DC_LOAD_WARPS;
constexpr int DC_SLAB_STAGES_MAX = 4, DC_B_STAGES_MAX = 4;
constexpr int DC_LOAD_DEPTH = 12;
constexpr int DC_SMEM_LIMIT = 224 * 1024;

// n / d for n < 2^31 without a division: q = (umulhi(n, mul) + n) >> shr
struct DcDiv {
    uint32_t mul, shr, d;
};
static DcDiv dc_div(uint32_t d) {
    DcDiv f;
    f.d = d;
    f.shr = 0;
    while ((1u << f.shr) < d) ++f.shr;
    f.mul = (uint32_t)((((uint64_t)1 << 32) * (((uint64_t)1 << f.shr) - d)) / d + 1);
    return f;
}
__device__ __forceinline__ uint32_t dc_quot(uint32_t n, const DcDiv &f) { return (__umulhi(n, f.mul) + n) >> f.shr; }

struct DcParams {
    const float *in;        // [B][Hi][Wi][Ci]
    const float *wprep;     // per (chunk, tap): the shared-memory image of the B tile
    float *out;             // [B][Ho][Wo][N]
    const float *bias;      // forward: [N] or null
    const float *yact;      // data gradient: output of the layer that produced x (act' is folded in), or null
    int act;                // forward: activation; data gradient: which derivative
    int accumulate;         // data gradient: dx += result
    int dgrad;
    int B, Hi, Wi, Ci, Ho, Wo, N;
    int pad_top, pad_left, kw, taps;
    int Hs, Wp;             // padded grid per image
    int chunks, CK, KC;     // channel chunks, channels per chunk (multiple of 8), 16-byte columns per chunk (CK / 4)
    int BN;                 // accumulator width (N rounded up to 16)
    int slab_positions, pitch;              // positions per slab, bytes between 16-byte columns (== 16 mod 128)
    int slab_stage_bytes, slab_stages, b_stage_bytes, b_part_bytes, b_stages;
    int total_positions, tiles, tmem_cols;
    DcDiv div_wp, div_hs;
    long long *trace;       // FNET_DC_TRACE: per-role clock64 stamps of CTA 0 (development aid), else null
};

__device__ __forceinline__ uint32_t dc_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void dc_bar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dc_smem(bar)), "r"(count));
}
__device__ __forceinline__ void dc_bar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t addr = dc_smem(bar), done = 0;
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
__device__ __forceinline__ void dc_bar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dc_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dc_bar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(dc_smem(bar)) : "memory");
}
__device__ __forceinline__ void dc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(dc_smem(bar)) : "memory");
}
__device__ __forceinline__ void dc_bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(dc_smem(bar))
                 : "memory");
}
__device__ __forceinline__ void dc_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
// one lane of a converged warp (the same one every time: the lowest)
__device__ __forceinline__ bool dc_elect() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
// K-major, no swizzle: rows 16 bytes apart inside a core matrix, 8-row groups `sbo` bytes apart, 16-byte columns `lbo` apart
__device__ __forceinline__ uint64_t dc_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(lbo >> 4) << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// head = x rounded to tf32 (nearest, ties away), tail = x - head rounded the same way (see gemm_tma.cu)
__device__ __forceinline__ void dc_hi_lo(float x, float &hi, float &lo) {
    const uint32_t h = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
    hi = __uint_as_float(h);
    lo = __uint_as_float((__float_as_uint(x - hi) + 0x1000u) & 0xffffe000u);
}

// development aid: CTA 0 writes (role, event, clock) triples; slot 0 of the buffer is the running count
__device__ __forceinline__ void dc_stamp(const DcParams &p, int role, int event) {
    if (p.trace && blockIdx.x == 0) {
        const unsigned long long i = atomicAdd(reinterpret_cast<unsigned long long *>(p.trace), 1ull);
        if (i < 4000) { p.trace[1 + 3 * i] = role; p.trace[2 + 3 * i] = event; p.trace[3 + 3 * i] = clock64(); }
    }
}

// position q of the padded grid -> (image, padded row, padded column)
__device__ __forceinline__ void dc_decode(const DcParams &p, uint32_t q, int &img, int &hh, int &ww) {
    const uint32_t g = dc_quot(q, p.div_wp);
    ww = (int)(q - g * (uint32_t)p.Wp);
    const uint32_t i = dc_quot(g, p.div_hs);
    hh = (int)(g - i * (uint32_t)p.Hs);
    img = (int)i;
}

// w [kh][kw][C][K] -> the B tiles.  Tile (chunk, tap) = heads [KC][BN][4] (then tails [KC][BN][4] when X3): element
// (kc, n, e) is the weight that multiplies reduction channel k = chunk * CK + 4 kc + e into output column n --
// forward: w[tap][k][n]; data gradient: w[taps - 1 - tap][n][k] (mirrored taps, reduction over the out channels).
template <bool X3>
__global__ void __launch_bounds__(256) dconv_prep_kernel(const float *__restrict__ w, float *__restrict__ wprep, int dgrad,
                                                         int C, int K, int taps, int chunks, int CK, int KC, int BN,
                                                         int Cred, int N, int b_stage_floats, int b_part_floats) {
    pdl_prologue();
    const int total = chunks * taps * KC * BN;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int n = i % BN;
        int t = i / BN;
        const int kc = t % KC;
        t /= KC;
        const int tap = t % taps, chunk = t / taps;
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int k = chunk * CK + kc * 4 + e;
            float x = 0.f;
            if (k < Cred && n < N && kc * 4 + e < CK)
                x = dgrad ? __ldg(w + ((size_t)(taps - 1 - tap) * C + n) * K + k) : __ldg(w + ((size_t)tap * C + k) * K + n);
            v[e] = x;
        }
        float *dst = wprep + (size_t)(chunk * taps + tap) * b_stage_floats + ((size_t)kc * BN + n) * 4;
        if (X3) {
            float4 hi, lo;
            dc_hi_lo(v[0], hi.x, lo.x);
            dc_hi_lo(v[1], hi.y, lo.y);
            dc_hi_lo(v[2], hi.z, lo.z);
            dc_hi_lo(v[3], hi.w, lo.w);
            *reinterpret_cast<float4 *>(dst) = hi;
            *reinterpret_cast<float4 *>(dst + b_part_floats) = lo;
        } else {
            *reinterpret_cast<float4 *>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        }
    }
}

// KSTEPS = CK / 8: MMAs (triples of them in TF32x3 mode) per tap and chunk; a template parameter because the lane that
// issues them is the kernel's critical path -- with the loop unrolled its instructions per tap are a wait, the MMAs and a commit
template <bool X3, int KSTEPS>
__global__ void __launch_bounds__(DC_THREADS, 1) dconv_kernel(const DcParams p) {
    pdl_prologue();
    extern __shared__ uint8_t dc_smem_raw[];
    __shared__ uint64_t slab_full[DC_SLAB_STAGES_MAX], slab_empty[DC_SLAB_STAGES_MAX], b_full[DC_B_STAGES_MAX], b_empty[DC_B_STAGES_MAX],
        acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_holder;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(dc_smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t *slabs = smem;
    uint8_t *bst = slabs + p.slab_stages * p.slab_stage_bytes;
    const int slab_part = p.KC * p.pitch;                     // bytes of the head slab (the tails follow)

    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < DC_SLAB_STAGES_MAX; ++s) { dc_bar_init(&slab_full[s], DC_LOAD_WARPS); dc_bar_init(&slab_empty[s], 1); }
            for (int s = 0; s < DC_B_STAGES_MAX; ++s) { dc_bar_init(&b_full[s], 1); dc_bar_init(&b_empty[s], 1); }
            for (int a = 0; a < 2; ++a) { dc_bar_init(&acc_full[a], 1); dc_bar_init(&acc_empty[a], DC_EPI_WARPS); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dc_smem(&tmem_holder)),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_holder;
    const uint32_t acc_stride = (uint32_t)p.tmem_cols / 2;

    if (warp == 0) {
        if (dc_elect()) {
            int bs = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x)
                for (int ct = 0; ct < p.chunks * p.taps; ++ct) {          // (chunk, tap) in the order the MMA warp consumes them
                    dc_bar_wait(&b_empty[bs], phase ^ 1);
                    dc_bar_expect_tx(&b_full[bs], (uint32_t)p.b_stage_bytes);
                    dc_bulk_load(dc_smem(bst + bs * p.b_stage_bytes), reinterpret_cast<const uint8_t *>(p.wprep) + (size_t)ct * p.b_stage_bytes,
                                 (uint32_t)p.b_stage_bytes, &b_full[bs]);
                    if (++bs == p.b_stages) { bs = 0; phase ^= 1; }
                }
        }
    } else if (warp == 1) {
        // The whole warp walks the loop (warp-uniform control flow and addresses: the compiler keeps them in uniform
        // registers) and ONE elected lane issues the MMAs.  Issued from inside `if (lane == 0)` instead, every
        // tcgen05.mma was wrapped in a divergence-serialising loop with its descriptors rebuilt in vector registers and
        // moved over (R2UR): ~21 dependent instructions of a lone warp per MMA, 230 cycles against the 32 the MMA takes.
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(DC_BM >> 4) << 24);
        const uint32_t b_lbo = (uint32_t)p.BN * 16u;
        const uint64_t desc_hi = (uint64_t)((128u >> 4) | (1u << 14)) << 32;     // stride byte offset 128, descriptor version 1
        const uint32_t a_step = (uint32_t)(2 * p.pitch) >> 4, b_step = (2u * b_lbo) >> 4;      // one k-step (8 channels), in 16-byte units
        const uint32_t a_tail = (uint32_t)slab_part >> 4, b_tail = (uint32_t)p.b_part_bytes >> 4;
        const uint32_t slab0_lo = (dc_smem(slabs) >> 4) | (((uint32_t)p.pitch >> 4) << 16), slab_stride = (uint32_t)p.slab_stage_bytes >> 4;
        const uint32_t b0_lo = (dc_smem(bst) >> 4) | ((b_lbo >> 4) << 16), b_stride = (uint32_t)p.b_stage_bytes >> 4;
        const bool leader = dc_elect();
        int ss = 0, bs = 0, it = 0;
        uint32_t sphase = 0, bphase = 0, b_lo = b0_lo;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int acc = it & 1;
            if (leader) dc_stamp(p, 1, 3);
            dc_bar_wait(&acc_empty[acc], (uint32_t)(((it >> 1) & 1) ^ 1));
            if (leader) dc_stamp(p, 1, 4);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t d = tmem + acc * acc_stride;
            uint32_t accumulate = 0;
            for (int chunk = 0; chunk < p.chunks; ++chunk) {
                if (leader) dc_stamp(p, 1, 0);
                dc_bar_wait(&slab_full[ss], sphase);
                if (leader) dc_stamp(p, 1, 1);
                uint32_t a_row = slab0_lo + (uint32_t)ss * slab_stride;         // + 16 bytes per position of the tap's shift
                int s = 0;
                for (int tap = 0; tap < p.taps; ++tap) {
                    dc_bar_wait(&b_full[bs], bphase);
                    if (leader) {
                        const uint64_t da = desc_hi | (a_row + (uint32_t)s), db = desc_hi | b_lo;
#pragma unroll
                        for (int j = 0; j < KSTEPS; ++j) {
                            if (X3) {
                                dc_mma(d, da + (a_tail + j * a_step), db + j * b_step, idesc, j ? 1u : accumulate);
                                dc_mma(d, da + j * a_step, db + (b_tail + j * b_step), idesc, 1u);
                                dc_mma(d, da + j * a_step, db + j * b_step, idesc, 1u);                  // the large term last
                            } else {
                                dc_mma(d, da + j * a_step, db + j * b_step, idesc, j ? 1u : accumulate);
                            }
                        }
                        dc_commit(&b_empty[bs]);
                    }
                    accumulate = 1u;
                    b_lo += b_stride;
                    if (++bs == p.b_stages) { bs = 0; bphase ^= 1; b_lo = b0_lo; }
                    if (++s == p.kw) { s = 0; a_row += (uint32_t)p.Wp; }
                }
                if (leader) { dc_commit(&slab_empty[ss]); dc_stamp(p, 1, 2); }
                if (++ss == p.slab_stages) { ss = 0; sphase ^= 1; }
            }
            if (leader) dc_commit(&acc_full[acc]);
        }
        __syncwarp();
    } else if (warp < 2 + DC_EPI_WARPS) {
        // tcgen05.ld.16x256b: the accumulator arrives in the MMA fragment layout -- thread t holds rows t / 4 and t / 4 + 8 of a
        // 16-lane half, columns 2 (t % 4) + {0, 1} of every 8-column block -- so the four lanes of a row store one 32-byte
        // sector and a warp instruction eight of them, straight from the registers (no transpose through shared memory).
        const int e = warp - 2, quad = warp & 3, group = e >> 2;         // TMEM lane quadrant = warp % 4; two warps share it
        int it = 0;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int acc = it & 1;
            long long pix[2][2];                                         // [16-lane half][row, row + 8]: element offset of the pixel, -1 = not an output
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    int img, hh, ww;
                    dc_decode(p, (uint32_t)tile * DC_BM + quad * 32 + h * 16 + r * 8 + (lane >> 2), img, hh, ww);
                    pix[h][r] = (img < p.B && hh < p.Ho && ww < p.Wo) ? (long long)((img * p.Ho + hh) * p.Wo + ww) * p.N : -1;
                }
            if (e == 0 && lane == 0) dc_stamp(p, 2, 0);
            dc_bar_wait(&acc_full[acc], (uint32_t)((it >> 1) & 1));
            if (e == 0 && lane == 0) dc_stamp(p, 2, 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int c0 = group * 32; c0 < p.BN; c0 += 64) {
                const int n = c0 + 2 * (lane & 3), cols = p.N - n;       // this thread's first column, and how many exist from it on
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t v[16];
                    const uint32_t taddr = tmem + ((uint32_t)(quad * 32 + h * 16) << 16) + acc * acc_stride + (uint32_t)c0;
                    asm volatile(
                        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (cols <= 0) continue;
                    const bool ra = pix[h][0] >= 0, rb = pix[h][1] >= 0;
                    float *qa = p.out + (ra ? pix[h][0] : 0) + n, *qb = p.out + (rb ? pix[h][1] : 0) + n;
                    if (!p.dgrad) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            if (8 * i < cols) {
                                const float bx = p.bias ? __ldg(p.bias + n + 8 * i) : 0.f, by = p.bias ? __ldg(p.bias + n + 8 * i + 1) : 0.f;
                                float2 a = make_float2(__uint_as_float(v[4 * i]) + bx, __uint_as_float(v[4 * i + 1]) + by);
                                float2 b = make_float2(__uint_as_float(v[4 * i + 2]) + bx, __uint_as_float(v[4 * i + 3]) + by);
                                if (p.act == FNET_ACT_RELU) {
                                    a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); b.x = fmaxf(b.x, 0.f); b.y = fmaxf(b.y, 0.f);
                                } else if (p.act != FNET_ACT_NONE) {
                                    a.x = act_apply(a.x, p.act); a.y = act_apply(a.y, p.act); b.x = act_apply(b.x, p.act); b.y = act_apply(b.y, p.act);
                                }
                                if (ra) *reinterpret_cast<float2 *>(qa + 8 * i) = a;
                                if (rb) *reinterpret_cast<float2 *>(qb + 8 * i) = b;
                            }
                        }
                    } else {
                        // multipliers act'(y) and old values first, all loads in flight at once
                        float2 ya[4], yb[4], oa[4], ob[4];
                        if (p.act) {
                            const float *pa = p.yact + (ra ? pix[h][0] : 0) + n, *pb = p.yact + (rb ? pix[h][1] : 0) + n;
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                ya[i] = (ra && 8 * i < cols) ? __ldg(reinterpret_cast<const float2 *>(pa + 8 * i)) : make_float2(0.f, 0.f);
                                yb[i] = (rb && 8 * i < cols) ? __ldg(reinterpret_cast<const float2 *>(pb + 8 * i)) : make_float2(0.f, 0.f);
                            }
                        }
                        if (p.accumulate) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                oa[i] = (ra && 8 * i < cols) ? *reinterpret_cast<const float2 *>(qa + 8 * i) : make_float2(0.f, 0.f);
                                ob[i] = (rb && 8 * i < cols) ? *reinterpret_cast<const float2 *>(qb + 8 * i) : make_float2(0.f, 0.f);
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            if (8 * i < cols) {
                                float2 a = make_float2(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]));
                                float2 b = make_float2(__uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                                if (p.accumulate) { a.x += oa[i].x; a.y += oa[i].y; b.x += ob[i].x; b.y += ob[i].y; }
                                if (p.act) {
                                    a.x *= act_grad(ya[i].x, p.act); a.y *= act_grad(ya[i].y, p.act);
                                    b.x *= act_grad(yb[i].x, p.act); b.y *= act_grad(yb[i].y, p.act);
                                }
                                if (ra) *reinterpret_cast<float2 *>(qa + 8 * i) = a;
                                if (rb) *reinterpret_cast<float2 *>(qb + 8 * i) = b;
                            }
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) dc_bar_arrive(&acc_empty[acc]);
            if (e == 0 && lane == 0) dc_stamp(p, 2, 2);
        }
    } else {
        // thread t owns one 16-byte channel column (kc = t % KC; KC is 2, 4, 6 or 8 and divides the 192 loaders) and the
        // positions t / KC, t / KC + step, ... of every slab: at most DC_LOAD_DEPTH of them (dc_plan checks), so a whole
        // slab fits in the loaders' registers.  That makes the loads independent of the slab buffers: the values of the
        // NEXT (tile, chunk) are requested right after the current slab has been handed over, and travel while the
        // loaders wait for a free buffer -- the round trip to L2 / HBM is off the critical path.
        const int t = threadIdx.x - 32 * (2 + DC_EPI_WARPS);
        const int kc = t % p.KC, j_first = t / p.KC, step = DC_LOADERS / p.KC;
        const int step_rows = step / p.Wp, step_cols = step - step_rows * p.Wp;
        float4 v[DC_LOAD_DEPTH];
        // The element offset of the pixel is carried along with (img, hh, ww): a step of `step` positions is a handful of
        // integer adds (no multiplication per load), a wrap of the padded column / row a constant correction.
        const int d_col = step_cols * p.Ci, d_row = p.Wi * p.Ci, d_wrap_w = -p.Wp * p.Ci, d_wrap_h = -p.Hs * p.Wi * p.Ci + p.Hi * p.Wi * p.Ci;
        auto request = [&](int tile, int chunk) {
            int img, hh, ww;
            dc_decode(p, (uint32_t)tile * DC_BM + (uint32_t)j_first, img, hh, ww);
            const int c = chunk * p.CK + kc * 4;
            const bool c_ok = c < p.Ci && tile < p.tiles;
            int off = ((img * p.Hi + (hh - p.pad_top)) * p.Wi + (ww - p.pad_left)) * p.Ci + c;   // < 2^31: dconv_supported
            int row = hh - p.pad_top, cl = ww - p.pad_left;
#pragma unroll
            for (int u = 0; u < DC_LOAD_DEPTH; ++u) {
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (c_ok && j_first + u * step < p.slab_positions && img < p.B && (unsigned)row < (unsigned)p.Hi && (unsigned)cl < (unsigned)p.Wi)
                    v[u] = __ldg(reinterpret_cast<const float4 *>(p.in + off));
                cl += step_cols;
                row += step_rows;
                off += d_col + step_rows * d_row;
                if (cl >= p.Wp - p.pad_left) { cl -= p.Wp; ++row; off += d_wrap_w + d_row; }
                while (row >= p.Hs - p.pad_top) { row -= p.Hs; ++img; off += d_wrap_h; }
            }
        };
        int ss = 0;
        uint32_t phase = 0;
        request(blockIdx.x, 0);
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
            for (int chunk = 0; chunk < p.chunks; ++chunk) {
                if (t == 0) dc_stamp(p, 3, 0);
                dc_bar_wait(&slab_empty[ss], phase ^ 1);
                if (t == 0) dc_stamp(p, 3, 1);
                uint8_t *col = slabs + ss * p.slab_stage_bytes + kc * p.pitch;
#pragma unroll
                for (int u = 0; u < DC_LOAD_DEPTH; ++u) {
                    const int j = j_first + u * step;
                    if (j < p.slab_positions) {
                        if (X3) {
                            float4 hi, lo;
                            dc_hi_lo(v[u].x, hi.x, lo.x);
                            dc_hi_lo(v[u].y, hi.y, lo.y);
                            dc_hi_lo(v[u].z, hi.z, lo.z);
                            dc_hi_lo(v[u].w, hi.w, lo.w);
                            *reinterpret_cast<float4 *>(col + j * 16) = hi;
                            *reinterpret_cast<float4 *>(col + slab_part + j * 16) = lo;
                        } else {
                            *reinterpret_cast<float4 *>(col + j * 16) = v[u];
                        }
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy stores -> the tensor core's reads
                __syncwarp();
                if (lane == 0) dc_bar_arrive(&slab_full[ss]);
                if (t == 0) dc_stamp(p, 3, 2);
                if (++ss == p.slab_stages) { ss = 0; phase ^= 1; }
                if (chunk + 1 < p.chunks) request(tile, chunk + 1);
                else request(tile + (int)gridDim.x, 0);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)p.tmem_cols) : "memory");
}

struct DcPlan {
    DcParams p;
    size_t smem, ws_bytes;
    bool ok;
};

// product 0: forward (reduce over C, N = K, output grid Ho x Wo); 1: data gradient (reduce over K, N = C, output grid
// H x W, input dz with the padding mirrored: kh - 1 - pt rows above, kw - 1 - pl columns to the left)
static DcPlan dc_plan(const ConvGeom &g, bool x3, int product) {
    DcPlan t{};
    DcParams &p = t.p;
    p.dgrad = product;
    p.B = g.B;
    if (product == 0) {
        p.Hi = g.H; p.Wi = g.W; p.Ci = g.C; p.Ho = g.Ho; p.Wo = g.Wo; p.N = g.K;
        p.pad_top = g.pt; p.pad_left = g.pl;
    } else {
        p.Hi = g.Ho; p.Wi = g.Wo; p.Ci = g.K; p.Ho = g.H; p.Wo = g.W; p.N = g.C;
        p.pad_top = g.kh - 1 - g.pt; p.pad_left = g.kw - 1 - g.pl;
    }
    p.kw = g.kw;
    p.taps = g.kh * g.kw;
    p.Hs = p.Ho + g.kh - 1;
    p.Wp = p.Wo + g.kw - 1;
    p.chunks = (p.Ci + 31) / 32;
    p.CK = (((p.Ci + p.chunks - 1) / p.chunks) + 7) / 8 * 8;
    p.KC = p.CK / 4;
    p.BN = (p.N + 15) / 16 * 16;
    p.slab_positions = DC_BM + (g.kh - 1) * p.Wp + g.kw - 1;
    p.pitch = (p.slab_positions * 16 + 127) / 128 * 128 + 16;       // == 16 (mod 128): the 8 stores of a quarter warp hit 8 bank groups
    const int parts = x3 ? 2 : 1;
    p.slab_stage_bytes = (parts * p.KC * p.pitch + 127) / 128 * 128;
    p.b_part_bytes = p.KC * p.BN * 16;
    p.b_stage_bytes = parts * p.b_part_bytes;
    const long long positions = (long long)p.B * p.Hs * p.Wp;
    if (positions + p.slab_positions + DC_BM >= 0x7fffffffLL || p.BN > 256) return t;
    if (((long long)p.B + 1) * p.Hs * p.Wp * p.Ci >= 0x7fffffffLL) return t;      // the loaders carry 32-bit element offsets (with slack for the halo)
    if (p.slab_positions > DC_LOAD_DEPTH * (DC_LOADERS / p.KC)) return t;       // a slab must fit in the loaders' registers
    p.total_positions = (int)positions;
    p.tiles = (int)((positions + DC_BM - 1) / DC_BM);
    // shared memory: two slab stages and two B stages at least; then B up to three stages, then slabs up to four
    // (a slab stage is refilled while the other chunks are multiplied: more stages = more loads in flight), then B up to four
    const long long work = (long long)p.chunks * p.taps, avail = DC_SMEM_LIMIT - 128;
    long long ns = 2, nb = work < 2 ? work : 2;
    if (ns * p.slab_stage_bytes + nb * p.b_stage_bytes > avail) return t;
    auto fits = [&](long long a, long long b) { return a * p.slab_stage_bytes + b * p.b_stage_bytes <= avail; };
    if (nb < work && nb < 3 && fits(ns, nb + 1)) ++nb;
    while (ns < DC_SLAB_STAGES_MAX && fits(ns + 1, nb)) ++ns;
    while (nb < work && nb < DC_B_STAGES_MAX && fits(ns, nb + 1)) ++nb;
    p.slab_stages = (int)ns;
    p.b_stages = (int)nb;
    const long long fixed = 128 + ns * p.slab_stage_bytes, b_stages = nb;
    p.tmem_cols = 64;
    while (p.tmem_cols < 2 * ((p.BN + 31) / 32 * 32)) p.tmem_cols *= 2;
    p.div_wp = dc_div((uint32_t)p.Wp);
    p.div_hs = dc_div((uint32_t)p.Hs);
    t.smem = (size_t)(fixed + b_stages * p.b_stage_bytes);
    t.ws_bytes = (size_t)p.chunks * p.taps * p.b_stage_bytes;
    t.ok = true;
    return t;
}

bool dconv_supported(const ConvGeom &g, int math, int product) {
    static const int mode = [] {                                // FNET_DCONV=0 disables, 2 = wherever supported (A/B measurements)
        const char *e = getenv("FNET_DCONV");
        return e && e[0] ? atoi(e) : 1;
    }();
    if (!mode || (math != FNET_MATH_TF32 && math != FNET_MATH_TF32X3) || product > 1) return false;
    const int taps = g.kh * g.kw;
    if (g.sh != 1 || g.sw != 1 || g.dh != 1 || g.dw != 1 || taps < 2 || taps > 64 || (g.C & 3) || (g.K & 3)) return false;
    if (g.pt > g.kh - 1 || g.pl > g.kw - 1) return false;
    const int red = product == 0 ? g.C : g.K, n = product == 0 ? g.K : g.C;
    if (red < 8 || n < 8 || n > 256) return false;
    // Measured against the shifted-box kernel (gemm_tma.cu) at batch 128, scripts/dconv_ab.sh.  TF32x3: 1.2x - 1.45x faster
    // (3x3 64 -> 64: 127 / 110 -> 85 / 82 us forward / data gradient; 24 -> 24: 60 / 53 -> 42 / 38; 5x5 12 -> 120 forward
    // 56 -> 43) except where a short reduction feeds a wide output (3x1 24 -> 192: 103 -> 123 us), which is bound by
    // the stores.  Plain TF32 (one MMA per k-step instead of three, no split) gains only on long reductions into
    // narrow outputs (64 -> 64: 60 / 57 -> 54 / 50 us) and loses on the rest.  FNET_DCONV=2 takes every supported layer.
    if (mode != 2) {
        const long long reduction = (long long)taps * red;
        if (math == FNET_MATH_TF32X3 ? (n > 128 && reduction < 256) : !(red >= 64 && n <= 64)) return false;
    }
    const DcPlan t = dc_plan(g, math == FNET_MATH_TF32X3, product);
    static const long long min_tiles = [] {
        const char *e = getenv("FNET_DCONV_MIN_TILES");
        return e ? atoll(e) : (long long)kNumSMs / 2;
    }();
    return t.ok && t.p.tiles >= min_tiles;
}

size_t dconv_workspace(const ConvGeom &g, int math) {
    size_t need = 0;
    for (int product = 0; product < 2; ++product)
        if (dconv_supported(g, math, product)) {
            const size_t b = dc_plan(g, math == FNET_MATH_TF32X3, product).ws_bytes;
            if (b > need) need = b;
        }
    return need;
}

template <bool X3>
static int dc_run(const ConvGeom &g, DcPlan &t, const float *w, void *ws, size_t ws_bytes, cudaStream_t stream, const char *what) {
    DcParams &p = t.p;
    if (!ws || ws_bytes < t.ws_bytes) {
        set_error("%s: workspace of %zu bytes needed, %zu given", what, t.ws_bytes, ws_bytes);
        return FNET_ERR_WORKSPACE;
    }
    if ((reinterpret_cast<uintptr_t>(p.in) & 15) || (reinterpret_cast<uintptr_t>(ws) & 15) ||
        ((reinterpret_cast<uintptr_t>(p.out) | reinterpret_cast<uintptr_t>(p.yact)) & 7)) {
        set_error("%s: operands must be 16-byte aligned (results 8-byte)", what);
        return FNET_ERR_ARG;
    }
    float *wprep = reinterpret_cast<float *>(ws);
    p.wprep = wprep;
    const int prep_items = p.chunks * p.taps * p.KC * p.BN;
    launch_pdl(dconv_prep_kernel<X3>, dim3((unsigned)stream_grid(prep_items, 256)), dim3(256), 0, stream, w, wprep, p.dgrad, g.C, g.K,
               p.taps, p.chunks, p.CK, p.KC, p.BN, p.Ci, p.N, p.b_stage_bytes / 4, p.b_part_bytes / 4);
    const unsigned grid = (unsigned)(p.tiles < kNumSMs ? p.tiles : kNumSMs);
    cudaError_t e = cudaSuccess;
    p.trace = nullptr;
    static const char *trace_path = getenv("FNET_DC_TRACE");   // development aid: dump CTA 0's role timeline of every launch
    if (trace_path) {
        cudaMalloc(&p.trace, (1 + 3 * 4000) * sizeof(long long));
        cudaMemsetAsync(p.trace, 0, (1 + 3 * 4000) * sizeof(long long), stream);
    }
#define FNET_DC_LAUNCH(KS)                                                                                        \
    case KS: {                                                                                                    \
        static std::atomic<uint64_t> configured{0}; /* one mask per instantiation, one bit per device */          \
        e = allow_dynamic_smem(dconv_kernel<X3, KS>, DC_SMEM_LIMIT, configured);                                  \
        if (e == cudaSuccess) launch_pdl(dconv_kernel<X3, KS>, dim3(grid), dim3(DC_THREADS), t.smem, stream, p);   \
        break;                                                                                                    \
    }
    switch (p.CK >> 3) { FNET_DC_LAUNCH(1) FNET_DC_LAUNCH(2) FNET_DC_LAUNCH(3) default: FNET_DC_LAUNCH(4) }
#undef FNET_DC_LAUNCH
    if (e != cudaSuccess) { set_error("%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e)); return FNET_ERR_CUDA; }
    if (trace_path) {
        static long long host[1 + 3 * 4000];
        cudaStreamSynchronize(stream);
        cudaMemcpy(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        if (FILE *f = fopen(trace_path, "w")) {
            fprintf(f, "# %s tiles %d chunks %d taps %d slab_stages %d b_stages %d\n", what, p.tiles, p.chunks, p.taps, p.slab_stages, p.b_stages);
            const long long n = host[0] < 4000 ? host[0] : 4000;
            for (long long i = 0; i < n; ++i) fprintf(f, "%lld %lld %lld\n", host[1 + 3 * i], host[2 + 3 * i], host[3 + 3 * i]);
            fclose(f);
        }
    }
    return check_launch(what);
}

int dconv_fwd(const ConvGeom &g, int math, int act, const float *x, const float *w, const float *bias, float *y, void *ws,
              size_t ws_bytes, cudaStream_t stream) {
    const bool x3 = math == FNET_MATH_TF32X3;
    DcPlan t = dc_plan(g, x3, 0);
    if (!t.ok) { set_error("conv2d_fwd[direct]: unsupported geometry"); return FNET_ERR_UNSUPPORTED; }
    t.p.in = x; t.p.out = y; t.p.bias = bias; t.p.yact = nullptr;
    t.p.act = act == FNET_ACT_SOFTMAX ? FNET_ACT_NONE : act;
    t.p.accumulate = 0;
    return x3 ? dc_run<true>(g, t, w, ws, ws_bytes, stream, "conv2d_fwd[tf32x3,direct]")
              : dc_run<false>(g, t, w, ws, ws_bytes, stream, "conv2d_fwd[tf32,direct]");
}

int dconv_dgrad(const ConvGeom &g, int math, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws, size_t ws_bytes,
                cudaStream_t stream) {
    const bool x3 = math == FNET_MATH_TF32X3;
    DcPlan t = dc_plan(g, x3, 1);
    if (!t.ok) { set_error("conv2d_bwd_data[direct]: unsupported geometry"); return FNET_ERR_UNSUPPORTED; }
    t.p.in = dz; t.p.out = dx; t.p.bias = nullptr; t.p.yact = ep.act ? ep.y : nullptr;
    t.p.act = ep.act;
    t.p.accumulate = ep.accumulate;
    return x3 ? dc_run<true>(g, t, w, ws, ws_bytes, stream, "conv2d_bwd_data[tf32x3,direct]")
              : dc_run<false>(g, t, w, ws, ws_bytes, stream, "conv2d_bwd_data[tf32,direct]");
}

}  // namespace fnet
