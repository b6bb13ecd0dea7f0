// fp32 implicit-GEMM convolution / dense kernels on the FFMA pipe (FNET_MATH_FP32).
//
// This is the parity path: fp32 operands, fp32 accumulation, a fixed reduction order, so the
// result is reproducible run to run and comparable with the float64 oracle at 1e-5.  The
// tensor-core path (igemm_tc.cu) shares the geometry code below.
//
// One kernel template serves the three products of a convolution; the "problem" type says
// how GEMM coordinates map to tensor elements:
//   forward        C[m, n]  m = (b, ho, wo)   n = out channel   k = (r, s, c)
//   backward data  C[m, n]  m = (b, h, w)     n = in channel    k = (r, s, out channel)
//   backward filter C[m, n] m = (r, s, c) (+1 row of ones = bias)  n = out channel  k = (b, ho, wo)
// A tile rows keep a decoded context in shared memory (decoded once per CTA); the column
// context is decoded once per thread per k-tile, so the inner gather has no divisions.
// When the output has too few tiles to fill 148 SMs the reduction axis is split across
// blockIdx.z; partial tiles go to the caller's workspace and are summed in a fixed order.
#include "igemm_problems.cuh"

namespace fnet {

// ------------------------------------------------------------------ the tile kernel
// 256 threads compute a BM x BN tile; the reduction axis is walked BK at a time.  The operands of
// tile k+1 are gathered into registers (all loads of a tile are issued back to back) while tile k
// is being multiplied out of shared memory, so the global-load latency of the small, latency-bound
// problems FeatureNet produces overlaps with the FFMA work.
template <class P, int BM, int BN, int TM, int TN, int BK>
__global__ void __launch_bounds__(256, 3) igemm_simt_kernel(P p, int M, int N, int Kg, int k_per_split,
                                                          float *__restrict__ partial) {
    pdl_prologue();
    static_assert((BM / TM) * (BN / TN) == 256, "256 threads per CTA");
    constexpr int RG = 256 / BK;                 // row groups of the A loader
    constexpr int A_PER = BM / RG;               // A elements gathered per thread per tile
    constexpr int B_PER = (BK * BN + 255) / 256; // B elements per thread per tile
    constexpr int PAD = TM == 1 ? 1 : (TM == 2 ? 2 : 4);
    static_assert(BM % RG == 0, "loader shape");
    // Lanes of a warp sit on the axis along which an operand is contiguous in global memory.  A: along k
    // (channels of an im2col row) for forward / data gradient, along the rows (channels of x) for the filter
    // gradient; B: along n for row-major weights / dz, along k for the data gradient's w[tap][c][ko].
    constexpr bool A_ROWS = !P::kAContigK;
    constexpr bool B_KLANES = P::kBContigK;
    constexpr int KG = BM >= 256 ? 1 : 256 / BM;  // k-groups of the row-wise A loader
    static_assert(!A_ROWS || (BK % KG == 0 && BK / KG == A_PER), "row-wise loader shape");
    __shared__ __align__(16) float As[BK][BM + PAD];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    __shared__ typename P::Row rows[A_ROWS ? 1 : BM];
    __shared__ typename P::Col cols[A_ROWS ? 2 : 1][A_ROWS ? BK : 1];   // per-tile k contexts, double-buffered

    const int tid = threadIdx.x;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int kbeg = blockIdx.z * k_per_split;
    const int kend = min(Kg, kbeg + k_per_split);

    // column context and B cursors are decoded once and then stepped from tile to tile (no divisions)
    typename P::Col col;
    typename P::Row myrow;
    // when four consecutive k are four consecutive channels (C % 4 == 0) a thread gathers 16 bytes at a time:
    // a quarter of the address arithmetic, which is what this kernel mostly executes besides FFMA
    constexpr int CPR = BK / 4;                   // 16-byte chunks per tile row
    constexpr int RPP = 256 / CPR;                // rows covered per pass of the vector loader
    static_assert(A_PER % 4 == 0 && BM % RPP == 0 && BM / RPP == A_PER / 4, "vector loader shape");
    const bool v4 = !A_ROWS && p.vec4_ok();
    // row-wise loader, 16 bytes = four consecutive rows (channels of one tap) of one k
    constexpr int RC = BM / 4;                    // row chunks per k
    constexpr int KG4 = 256 / RC;                 // k-groups of the vector row-wise loader
    static_assert(!A_ROWS || (BK % KG4 == 0 && BK / KG4 == A_PER / 4), "vector row-wise loader shape");
    const bool r4 = A_ROWS && p.a_rows_vec4_ok();
    const int a_col = A_ROWS ? 0 : (v4 ? 4 * (tid % CPR) : tid % BK), a_rg = A_ROWS ? 0 : (v4 ? tid / CPR : tid / BK);
    if constexpr (A_ROWS) {
        p.make_row(r4 ? m0 + 4 * (tid % RC) : m0 + tid % BM, M, myrow);
        if (tid < BK) { p.make_col(kbeg + tid, kend, col); cols[0][tid] = col; }
    } else {
        for (int i = tid; i < BM; i += 256) p.make_row(m0 + i, M, rows[i]);
        p.make_col(kbeg + a_col, kend, col);
    }
    __syncthreads();

    const int tx = tid % (BN / TN), ty = tid / (BN / TN);

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    // Filter gradient whose tap rows fill the tiles exactly: the all-ones bias row would cost a whole extra
    // row of tiles, so the first row of CTAs sums the columns of its B tiles (dz) instead -- same result.
    const int bias_row = P::kBiasRow ? p.colsum_row(BM) : -1;
    const bool do_colsum = bias_row >= 0 && blockIdx.x == 0 && ty == 0;
    float colsum[TN];
#pragma unroll
    for (int j = 0; j < TN; ++j) colsum[j] = 0.f;

    float a_reg[A_PER], b_reg[B_PER];
    typename P::BCursor bcur[B_PER];
    bool b_live[B_PER];
    auto b_kk = [&](int i) { int e = tid + 256 * i; return B_KLANES ? e % BK : e / BN; };
    auto b_nn = [&](int i) { int e = tid + 256 * i; return B_KLANES ? e / BK : e % BN; };
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
        b_live[i] = tid + 256 * i < BK * BN && n0 + b_nn(i) < N;
        bcur[i] = p.b_begin(kbeg + b_kk(i), b_live[i] ? n0 + b_nn(i) : 0);
    }
    int parity = 0;
    auto gather = [&](int k0) {
        if constexpr (A_ROWS) {
            if (r4) {
#pragma unroll
                for (int i = 0; i < A_PER / 4; ++i) {
                    const typename P::Col &c = cols[parity][tid / RC + KG4 * i];
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (c.valid && myrow.valid == 2) v.x = 1.f;                  // the bias row opens its chunk
                    else if (c.valid && myrow.valid) v = load4(p.a_ptr(myrow, c));
                    a_reg[4 * i] = v.x; a_reg[4 * i + 1] = v.y; a_reg[4 * i + 2] = v.z; a_reg[4 * i + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < A_PER; ++i) {
                    const typename P::Col &c = cols[parity][tid / BM + KG * i];
                    a_reg[i] = (c.valid && myrow.valid) ? p.a(myrow, c) : 0.f;
                }
            }
        } else {
            if (v4) {
#pragma unroll
                for (int i = 0; i < A_PER / 4; ++i) {
                    const typename P::Row &row = rows[a_rg + RPP * i];
                    float4 v = (col.valid && row.valid) ? p.a4(row, col) : make_float4(0.f, 0.f, 0.f, 0.f);
                    a_reg[4 * i] = v.x; a_reg[4 * i + 1] = v.y; a_reg[4 * i + 2] = v.z; a_reg[4 * i + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < A_PER; ++i) {
                    const typename P::Row &row = rows[a_rg + RG * i];
                    a_reg[i] = (col.valid && row.valid) ? p.a(row, col) : 0.f;
                }
            }
            p.next_col(k0 + BK + a_col, BK, kend, col);
        }
#pragma unroll
        for (int i = 0; i < B_PER; ++i)
            b_reg[i] = (b_live[i] && k0 + b_kk(i) < kend) ? __ldg(p.b_ptr(bcur[i])) : 0.f;
#pragma unroll
        for (int i = 0; i < B_PER; ++i) p.b_next(bcur[i], BK);
    };

    if (kbeg < kend) gather(kbeg);
    for (int k0 = kbeg; k0 < kend; k0 += BK) {
        if constexpr (A_ROWS) {
            if (r4) {
#pragma unroll
                for (int i = 0; i < A_PER / 4; ++i) {
                    float *dst = &As[tid / RC + KG4 * i][4 * (tid % RC)];
                    if constexpr ((BM + PAD) % 4 == 0) {
                        *reinterpret_cast<float4 *>(dst) = make_float4(a_reg[4 * i], a_reg[4 * i + 1], a_reg[4 * i + 2], a_reg[4 * i + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) dst[j] = a_reg[4 * i + j];
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < A_PER; ++i) As[tid / BM + KG * i][tid % BM] = a_reg[i];
            }
            parity ^= 1;                              // contexts of the next tile, read after the barrier
            if (tid < BK) { p.next_col(k0 + BK + tid, BK, kend, col); cols[parity][tid] = col; }
        } else if (v4) {
#pragma unroll
            for (int i = 0; i < A_PER / 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) As[a_col + j][a_rg + RPP * i] = a_reg[4 * i + j];
        } else {
#pragma unroll
            for (int i = 0; i < A_PER; ++i) As[a_col][a_rg + RG * i] = a_reg[i];
        }
#pragma unroll
        for (int i = 0; i < B_PER; ++i)
            if (tid + 256 * i < BK * BN) Bs[b_kk(i)][b_nn(i)] = b_reg[i];
        __syncthreads();
        if (k0 + BK < kend) gather(k0 + BK);      // in flight while this tile is multiplied
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[TM], b[TN];
            if constexpr (TM % 4 == 0) {
#pragma unroll
                for (int i = 0; i < TM; i += 4) {
                    float4 t = *reinterpret_cast<const float4 *>(&As[kk][ty * TM + i]);
                    a[i] = t.x; a[i + 1] = t.y; a[i + 2] = t.z; a[i + 3] = t.w;
                }
            } else if constexpr (TM == 2) {
                float2 t = *reinterpret_cast<const float2 *>(&As[kk][ty * 2]);
                a[0] = t.x; a[1] = t.y;
            } else {
                a[0] = As[kk][ty];
            }
            if constexpr (TN % 4 == 0) {
#pragma unroll
                for (int j = 0; j < TN; j += 4) {
                    float4 t = *reinterpret_cast<const float4 *>(&Bs[kk][tx * TN + j]);
                    b[j] = t.x; b[j + 1] = t.y; b[j + 2] = t.z; b[j + 3] = t.w;
                }
            } else if constexpr (TN == 2) {
                float2 t = *reinterpret_cast<const float2 *>(&Bs[kk][tx * 2]);
                b[0] = t.x; b[1] = t.y;
            } else {
#pragma unroll
                for (int j = 0; j < TN; ++j) b[j] = Bs[kk][tx * TN + j];
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            if (do_colsum) {
#pragma unroll
                for (int j = 0; j < TN; ++j) colsum[j] += b[j];
            }
        }
        __syncthreads();
    }
    if (do_colsum) {
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            int n = n0 + tx * TN + j;
            if (n >= N) continue;
            if (partial) partial[((long long)blockIdx.z * M + bias_row) * N + n] = colsum[j];
            else p.store(bias_row, n, colsum[j]);
        }
    }

    if (!partial && p.has_pre()) {
        // data gradient that also carries an activation derivative: fetch all multipliers first (independent
        // loads in flight together), then store -- a load between the stores would serialise on memory latency
        float mult[TM][TN];
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                int m = m0 + ty * TM + i, n = n0 + tx * TN + j;
                mult[i][j] = (m < M && n < N) ? p.pre(m, n) : 0.f;
            }
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                int m = m0 + ty * TM + i, n = n0 + tx * TN + j;
                if (m < M && n < N) p.store_pre(m, n, acc[i][j], mult[i][j]);
            }
        return;
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        int m = m0 + ty * TM + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            int n = n0 + tx * TN + j;
            if (n >= N) continue;
            if (partial) partial[((long long)blockIdx.z * M + m) * N + n] = acc[i][j];
            else p.store(m, n, acc[i][j]);
        }
    }
}

struct SplitPlan {
    int bm, bn, splits, k_per_split, bk, m_tiles;
    size_t ws_bytes;
};

// bias_row: the last of the M rows is the filter gradient's all-ones row (see do_colsum in the kernel)
static SplitPlan plan_for(long long M, long long N, long long Kg, bool bias_row = false) {
    SplitPlan s;
    s.bn = N <= 8 ? 8 : (N <= 16 ? 16 : (N <= 32 ? 32 : 64));
    int big = s.bn <= 16 ? 256 : 128;
    long long ntiles_n = ceil_div<long long>(N, s.bn);
    // full-size tiles when they fill two waves of 148 SMs; otherwise smaller tiles (less padding on
    // skinny outputs, more CTAs), and the reduction axis is split to occupy the remaining SMs
    if (ceil_div<long long>(M, big) * ntiles_n >= 2 * kNumSMs) { s.bm = big; s.bk = 16; }
    else if (M > 32) { s.bm = 64; s.bk = 32; }
    else { s.bm = 32; s.bk = 64; }
    const long long tiled_rows = (bias_row && M > 1 && (M - 1) % s.bm == 0) ? M - 1 : M;
    s.m_tiles = (int)ceil_div<long long>(tiled_rows, s.bm);
    long long tiles = s.m_tiles * ntiles_n;
    long long splits = 1;
    if (tiles < 3 * kNumSMs && Kg >= 4 * s.bk) {
        splits = 3 * kNumSMs / tiles;               // one wave of three CTAs per SM (launch bounds), never a few CTAs more
        long long most = Kg / (2 * s.bk);          // at least two k-tiles per split
        if (splits > most) splits = most;
        if (splits < 1) splits = 1;
    }
    long long per = ceil_div<long long>(ceil_div<long long>(Kg, splits), s.bk) * s.bk;
    s.k_per_split = (int)per;
    s.splits = (int)ceil_div<long long>(Kg, per);
    s.ws_bytes = s.splits > 1 ? (size_t)s.splits * M * N * sizeof(float) : 0;
    return s;
}

template <class P, int BM, int BN, int TM, int TN, int BKT>
static void launch_tile(const P &p, dim3 grid, int m, int n, int kg, int per, float *partial, cudaStream_t stream) {
    launch_pdl(igemm_simt_kernel<P, BM, BN, TM, TN, BKT>, dim3(grid), dim3(256), 0, stream, p, m, n, kg, per, partial);
}

template <class P>
static int run_problem(const P &p, long long M, long long N, long long Kg, void *ws, size_t ws_bytes,
                       cudaStream_t stream, const char *what) {
    if (M <= 0 || N <= 0) return FNET_OK;
    if (M > 0x7fffffffLL || Kg > 0x7fffffffLL || N > 0x7fffffffLL) {
        set_error("%s: GEMM extent exceeds 2^31", what);
        return FNET_ERR_UNSUPPORTED;
    }
    SplitPlan plan = plan_for(M, N, Kg, P::kBiasRow && p.has_bias_row());
    float *partial = nullptr;
    if (plan.splits > 1) {
        if (ws == nullptr || ws_bytes < plan.ws_bytes) {
            set_error("%s: workspace of %zu bytes needed, %zu given", what, plan.ws_bytes, ws_bytes);
            return FNET_ERR_WORKSPACE;
        }
        partial = reinterpret_cast<float *>(ws);
    }
    dim3 grid((unsigned)plan.m_tiles, (unsigned)ceil_div<long long>(N, plan.bn), plan.splits);
    if (grid.y > 65535u || grid.z > 65535u) {
        set_error("%s: too many tiles (%u x %u)", what, grid.y, grid.z);
        return FNET_ERR_UNSUPPORTED;
    }
    int m = (int)M, n = (int)N, kg = (int)Kg, per = plan.k_per_split;
    switch (plan.bn * 1000 + plan.bm) {
        case 8 * 1000 + 256: launch_tile<P, 256, 8, 4, 2, 16>(p, grid, m, n, kg, per, partial, stream); break;
        case 8 * 1000 + 64: launch_tile<P, 64, 8, 2, 1, 32>(p, grid, m, n, kg, per, partial, stream); break;
        case 8 * 1000 + 32: launch_tile<P, 32, 8, 1, 1, 64>(p, grid, m, n, kg, per, partial, stream); break;
        case 16 * 1000 + 256: launch_tile<P, 256, 16, 4, 4, 16>(p, grid, m, n, kg, per, partial, stream); break;
        case 16 * 1000 + 64: launch_tile<P, 64, 16, 2, 2, 32>(p, grid, m, n, kg, per, partial, stream); break;
        case 16 * 1000 + 32: launch_tile<P, 32, 16, 1, 2, 64>(p, grid, m, n, kg, per, partial, stream); break;
        case 32 * 1000 + 128: launch_tile<P, 128, 32, 4, 4, 16>(p, grid, m, n, kg, per, partial, stream); break;
        case 32 * 1000 + 64: launch_tile<P, 64, 32, 2, 4, 32>(p, grid, m, n, kg, per, partial, stream); break;
        case 32 * 1000 + 32: launch_tile<P, 32, 32, 2, 2, 64>(p, grid, m, n, kg, per, partial, stream); break;
        case 64 * 1000 + 128: launch_tile<P, 128, 64, 8, 4, 16>(p, grid, m, n, kg, per, partial, stream); break;
        case 64 * 1000 + 64: launch_tile<P, 64, 64, 4, 4, 32>(p, grid, m, n, kg, per, partial, stream); break;
        default: launch_tile<P, 32, 64, 2, 4, 64>(p, grid, m, n, kg, per, partial, stream); break;
    }
    int rc = check_launch(what);
    if (rc != FNET_OK) return rc;
    if (plan.splits > 1) rc = launch_reduce_splits(p, m, n, plan.splits, partial, stream, what);
    return rc;
}

static int geom_from_desc(const fnet_conv2d_desc *d, ConvGeom &g, const char *what) {
    if (d == nullptr) { set_error("%s: null descriptor", what); return FNET_ERR_ARG; }
    if (d->batch < 0 || d->in_h <= 0 || d->in_w <= 0 || d->in_c <= 0 || d->out_h <= 0 || d->out_w <= 0 ||
        d->out_c <= 0 || d->kh <= 0 || d->kw <= 0 || d->stride_h <= 0 || d->stride_w <= 0 || d->dil_h <= 0 ||
        d->dil_w <= 0 || d->pad_top < 0 || d->pad_left < 0) {
        set_error("%s: bad convolution geometry", what);
        return FNET_ERR_ARG;
    }
    // the last window must start inside the padded input
    if ((d->out_h - 1) * d->stride_h - d->pad_top >= d->in_h || (d->out_w - 1) * d->stride_w - d->pad_left >= d->in_w) {
        set_error("%s: output extent inconsistent with input, stride and padding", what);
        return FNET_ERR_ARG;
    }
    g = ConvGeom{d->batch, d->in_h, d->in_w, d->in_c, d->out_h, d->out_w, d->out_c, d->kh, d->kw,
                 d->stride_h, d->stride_w, d->pad_top, d->pad_left, d->dil_h, d->dil_w};
    return FNET_OK;
}

size_t simt_conv_workspace(const ConvGeom &g) {
    long long mo = (long long)g.B * g.Ho * g.Wo, mi = (long long)g.B * g.H * g.W;
    long long taps = (long long)g.kh * g.kw;
    size_t a = plan_for(mo, g.K, taps * g.C).ws_bytes;
    size_t b = plan_for(mi, g.C, taps * g.K).ws_bytes;
    size_t c = plan_for(taps * g.C + 1, g.K, mo, true).ws_bytes; // with the bias row ...
    size_t c0 = plan_for(taps * g.C, g.K, mo).ws_bytes;          // ... and without (dbias == nullptr)
    if (c0 > c) c = c0;
    return a > b ? (a > c ? a : c) : (b > c ? b : c);
}

int simt_conv_fwd(const ConvGeom &g, int act, const float *x, const float *w, const float *bias, float *y, void *ws,
                  size_t ws_bytes, cudaStream_t stream) {
    FwdProblem p{g, x, w, bias, y, act == FNET_ACT_SOFTMAX ? FNET_ACT_NONE : act};
    return run_problem(p, (long long)g.B * g.Ho * g.Wo, g.K, (long long)g.kh * g.kw * g.C, ws, ws_bytes, stream,
                       "conv2d_fwd");
}

int simt_conv_dgrad(const ConvGeom &g, const float *dz, const float *w, float *dx, DxEpilogue ep, void *ws,
                    size_t ws_bytes, cudaStream_t stream) {
    DgradProblem p{g, dz, w, dx, ep.accumulate, ep.act, ep.y};
    return run_problem(p, (long long)g.B * g.H * g.W, g.C, (long long)g.kh * g.kw * g.K, ws, ws_bytes, stream,
                       "conv2d_bwd_data");
}

int simt_conv_wgrad(const ConvGeom &g, const float *x, const float *dz, float *dw, float *dbias, void *ws,
                    size_t ws_bytes, cudaStream_t stream) {
    int taps_c = g.kh * g.kw * g.C;
    WgradProblem p{g, x, dz, dw, dbias, taps_c};
    return run_problem(p, (long long)taps_c + (dbias ? 1 : 0), g.K, (long long)g.B * g.Ho * g.Wo, ws, ws_bytes,
                       stream, "conv2d_bwd_filter");
}

int simt_reduce_fwd_partials(const ConvGeom &g, int act, const float *bias, float *y, int splits, const float *partial,
                             cudaStream_t stream) {
    FwdProblem p{g, nullptr, nullptr, bias, y, act};
    long long M = (long long)g.B * g.Ho * g.Wo;
    return launch_reduce_splits(p, (int)M, g.K, splits, partial, stream, "reduce_partials");
}

int simt_reduce_wgrad_partials(const ConvGeom &g, float *dw, float *dbias, int M, int splits, const float *partial,
                               cudaStream_t stream) {
    WgradProblem p{g, nullptr, nullptr, dw, dbias, g.kh * g.kw * g.C};
    return launch_reduce_splits(p, M, g.K, splits, partial, stream, "reduce_partials");
}

int geom_checked(const fnet_conv2d_desc *d, ConvGeom &g, const char *what) { return geom_from_desc(d, g, what); }

ConvGeom dense_geom(long long m, long long k, long long n) {
    // a dense layer is a 1x1 convolution over m "pixels" of one row each
    return ConvGeom{(int)m, 1, 1, (int)k, 1, 1, (int)n, 1, 1, 1, 1, 0, 0, 1, 1};
}

}  // namespace fnet

// =====================================================================================
// Allocation-free dense forward.
//
// Who needs it: a host that cannot hand this library a workspace because it cannot allocate
// device memory itself.  The case this was written for is libfnet loaded into the reference's
// own runtime: a TensorFlow 1.x session reserves nearly all of the GPU when it is created, so a
// plain cudaMalloc from a plug-in may simply fail, and TF's allocator is not reachable from C.
// The same holds for an evaluation worker that shares the GPU with any framework whose memory
// pool already owns the device, and for loops that must avoid the implicit device-wide
// synchronisation of cudaMalloc/cudaFree.
//
// The split-reduction partials, which fnet_dense_fwd keeps in the caller's workspace, live here
// in a fixed array inside the library image.  Every GPU that loads libfnet.so carries these
// FNET_STATIC_SCRATCH_BYTES (4 MiB) until the process exits.
// =====================================================================================
namespace fnet {

constexpr size_t kStaticScratchBytes = FNET_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) float g_static_scratch[kStaticScratchBytes / sizeof(float)];

constexpr int kMaxDevices = 64;
static float *g_static_scratch_ptr[kMaxDevices];   // device address of the array, per device ordinal

// In host code `g_static_scratch` is only a handle for the runtime, not a device address; the
// address a kernel can use comes from cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_for_current_device(float **out) {
    int device = -1;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) { set_error("fnet_dense_fwd_static: cudaGetDevice: %s", cudaGetErrorString(e)); return FNET_ERR_CUDA; }
    if (device < 0 || device >= kMaxDevices) { set_error("fnet_dense_fwd_static: device ordinal %d not supported", device); return FNET_ERR_UNSUPPORTED; }
    if (g_static_scratch_ptr[device] == nullptr) {
        void *address = nullptr;
        e = cudaGetSymbolAddress(&address, g_static_scratch);
        if (e != cudaSuccess || address == nullptr) {
            set_error("fnet_dense_fwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return FNET_ERR_CUDA;
        }
        g_static_scratch_ptr[device] = reinterpret_cast<float *>(address);
    }
    *out = g_static_scratch_ptr[device];
    return FNET_OK;
}

}  // namespace fnet

extern "C" int fnet_dense_fwd_static(int64_t m, int64_t k, int64_t n, int act, const float *x, const float *w,
                                     const float *bias, float *y, void *stream) {
    using namespace fnet;
    if (m < 0 || k <= 0 || n <= 0 || m > 0x7fffffffLL || k > 0x7fffffffLL || n > 0x7fffffffLL) {
        set_error("fnet_dense_fwd_static: bad GEMM shape m=%lld k=%lld n=%lld", (long long)m, (long long)k, (long long)n);
        return FNET_ERR_ARG;
    }
    if (act < FNET_ACT_NONE || act > FNET_ACT_SOFTMAX) { set_error("fnet_dense_fwd_static: unknown activation %d", act); return FNET_ERR_ARG; }
    if (m == 0) return FNET_OK;
    if (!x || !w || !y) { set_error("fnet_dense_fwd_static: null pointer"); return FNET_ERR_ARG; }
    size_t need = plan_for(m, n, k).ws_bytes;
    if (need > kStaticScratchBytes) {
        set_error("fnet_dense_fwd_static: this shape needs %zu bytes of scratch, the built-in array holds %zu; "
                  "use fnet_dense_fwd with a caller-provided workspace", need, kStaticScratchBytes);
        return FNET_ERR_WORKSPACE;
    }
    float *scratch = nullptr;
    int rc = static_scratch_for_current_device(&scratch);
    if (rc != FNET_OK) return rc;
    // Nothing in the scratch has to start at zero: every partial the reduction reads is written
    // earlier in this same call, so whatever the previous call left behind is never looked at.
    rc = simt_conv_fwd(dense_geom(m, k, n), act, x, w, bias, y, scratch, kStaticScratchBytes, as_stream(stream));
    if (rc == FNET_OK && act == FNET_ACT_SOFTMAX) rc = run_softmax_rows(m, n, y, y, as_stream(stream));
    return rc;
}
