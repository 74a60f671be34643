// fp32 kernels of the AWD-LSTM engine (precision mode AWD_PREC_FP32): the 1e-4-parity path.
//
// Everything here runs on the FP32 FMA pipes with accurate expf/tanhf so that logits and states
// match the reference's CPU fp32 arithmetic to ~1e-6 and greedy decoding is identical
// (BASELINE.json north_star).  The bf16 tensor-core path lives in gemm_tc.cu.
//
//   gather_f32_kernel   K1  EmbeddingDropout.forward eval (awd_lstm.py:114-122)
//   gemm_f32_kernel     K2/K3/K4  x*W_ih^T (+b) for all timesteps; h*W_hh^T + fused LSTM cell per
//                       step (nn.LSTM at awd_lstm.py:162); decoder Linear (awd_lstm.py:45)
//   skinny_f32_kernel   the same three contractions when only a few streams are live (bs <= 8):
//                       one warp per hidden unit streams its four gate rows once (HBM-bound),
//                       reduces with warp shuffles and applies the cell update in registers
//   sample_f32_kernel   multinomial(exp(z),10) + first-allowed rule (main.py:61-68)
#include <type_traits>

#include "common.cuh"

namespace awd {

// ------------------------------------------------------------------------------------------ K1
__global__ void gather_f32_vec_kernel(const float4* __restrict__ table, int ld4, int E4, int V,
                                      const int64_t* __restrict__ tok, int64_t n, float4* __restrict__ out,
                                      int64_t ldo4, int* err_flag) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t row = gid / E4;
    int col = (int)(gid - row * E4);
    if (row >= n) return;
    int64_t t = tok[row];
    if (t < 0 || t >= V) {
        if (col == 0) atomicExch(err_flag, 1);
        t = 0;
    }
    out[row * ldo4 + col] = __ldg(table + t * ld4 + col);
}

__global__ void gather_f32_scalar_kernel(const float* __restrict__ table, int ld, int E, int V,
                                         const int64_t* __restrict__ tok, int64_t n, float* __restrict__ out,
                                         int64_t ldo, int* err_flag) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t row = gid / E;
    int col = (int)(gid - row * E);
    if (row >= n) return;
    int64_t t = tok[row];
    if (t < 0 || t >= V) {
        if (col == 0) atomicExch(err_flag, 1);
        t = 0;
    }
    out[row * ldo + col] = __ldg(table + t * ld + col);
}

int launch_gather_f32(const float* table, int ld, int E, int V, const int64_t* tok, int64_t n,
                      float* out, int64_t ldo, int* err_flag, cudaStream_t st) {
    if (n == 0) return AWD_OK;
    bool vec = (E % 4 == 0) && (ldo % 4 == 0) && (((uintptr_t)out) % 16 == 0) && (ld % 4 == 0);
    if (vec) {
        int64_t total = n * (E / 4);
        gather_f32_vec_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(
            (const float4*)table, ld / 4, E / 4, V, tok, n, (float4*)out, ldo / 4, err_flag);
    } else {
        int64_t total = n * E;
        gather_f32_scalar_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(table, ld, E, V, tok, n, out, ldo,
                                                                              err_flag);
    }
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// ------------------------------------------------------------------------------- cell arithmetic
__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

// gates in nn.LSTM order i,f,g,o.  Returns h; updates c.
__device__ __forceinline__ float lstm_cell(float gi, float gf, float gg, float go, float& c) {
    float i = sigmoid_acc(gi), f = sigmoid_acc(gf), g = tanhf(gg), o = sigmoid_acc(go);
    c = f * c + i * g;
    return o * tanhf(c);
}

__device__ __forceinline__ void epilogue_one(const EpiStore& e, int m, int n, int N, const float (&v)[4]) {
    float* p = e.C + (int64_t)m * e.ldc + n;
    float b[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 4; ++q)
        if (e.bias != nullptr && n + q < N) b[q] = __ldg(e.bias + n + q);
    if (n + 3 < N && (((uintptr_t)p) & 15) == 0) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0] + b[0], v[1] + b[1], v[2] + b[2], v[3] + b[3]);
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (n + q < N) p[q] = v[q] + b[q];
    }
}

__device__ __forceinline__ void epilogue_one(const EpiCell& e, int m, int n, int N, const float (&v)[4]) {
    if (n >= N) return;
    int j = n >> 2;
    if (e.start != nullptr && e.t < __ldg(e.start + m)) {      // this stream's prompt has not begun yet
        e.h_out[(int64_t)m * e.ld_state + j] = 0.f;
        if (e.raw_out != nullptr) e.raw_out[(int64_t)m * e.ld_raw + j] = 0.f;
        return;
    }
    float4 add;
    if (e.addend != nullptr)
        add = *reinterpret_cast<const float4*>(e.addend + (int64_t)m * e.ld_add + n);
    else
        add = __ldg(reinterpret_cast<const float4*>(e.bias + n));
    float c = e.c[(int64_t)m * e.ld_state + j];
    float h = lstm_cell(v[0] + add.x, v[1] + add.y, v[2] + add.z, v[3] + add.w, c);
    e.c[(int64_t)m * e.ld_state + j] = c;
    e.h_out[(int64_t)m * e.ld_state + j] = h;
    if (e.raw_out != nullptr) e.raw_out[(int64_t)m * e.ld_raw + j] = h;
}

__device__ __forceinline__ const float* operand_row(const Operand& op, int m, int M) {
    if (m >= M) return nullptr;
    if (op.idx == nullptr) return op.A + (int64_t)m * op.lda;
    int64_t t = op.idx[m];
    if (t < 0 || t >= op.idx_limit) {
        atomicExch(op.err_flag, 1);
        t = 0;
    }
    return op.A + t * op.lda;
}

// ------------------------------------------------------------------------- sampling primitives
// Philox4x32-10 (Salmon et al. 2011): counter (column/4, stream, step_lo, step_hi), key = seed.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}

// key of column `col`: z/T for greedy, z/T - log(Exp(1)) otherwise.  The largest keys of an
// exponential race are exactly multinomial(exp(z/T)) draws without replacement (main.py:61).
__device__ __forceinline__ float race_key(float z, uint32_t bits, const SampleParams& sp) {
    float k = z * sp.inv_temperature;
    if (!sp.greedy) {
        float u = uniform_open01(bits);
        k -= logf(-logf(u));
    }
    return k;
}

struct KeyIdx {
    float key;
    int idx;
};
__device__ __forceinline__ bool better(float ka, int ia, float kb, int ib) {
    return (ka > kb) || (ka == kb && ia < ib);   // ties: lower id first
}
__device__ __forceinline__ KeyIdx warp_best(KeyIdx v) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        float k = __shfl_xor_sync(0xffffffffu, v.key, s);
        int i = __shfl_xor_sync(0xffffffffu, v.idx, s);
        if (better(k, i, v.key, v.idx)) { v.key = k; v.idx = i; }
    }
    return v;
}


// Race over the columns a thread holds for one row of a decoder tile: 4-aligned groups of 4 logits = one Philox call each, keyed
// exactly as sample_f32_kernel keys them (counter = column / 4, stream, step).
__device__ __forceinline__ void race_group(const EpiRace& e, const SampleParams& sp, int m, int n, int N, const float (&v)[4], KeyIdx& all, KeyIdx& ok) {
    if (n >= N) return;
    uint4 r = make_uint4(0, 0, 0, 0);
    if (!sp.greedy) r = philox4x32_10(make_uint4((uint32_t)(n >> 2), (uint32_t)m, sp.step_lo, sp.step_hi), sp.seed_lo, sp.seed_hi);
    const uint32_t bits[4] = {r.x, r.y, r.z, r.w};
    const int ban_lo = sp.n_banned ? sp.banned[0] : N, ban_hi = sp.n_banned ? sp.banned[sp.n_banned - 1] : -1;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int id = n + q;
        if (id >= N) continue;
        const float z = v[q] + (e.bias != nullptr ? __ldg(e.bias + id) : 0.f);
        const bool maybe_banned = id >= ban_lo && id <= ban_hi;
        // The exact key costs two accurate logarithms per logit.  A bound from the fast ones (error far below the 2e-3 margin)
        // decides first whether this id can beat the allowed best of the thread at all -- all.key >= ok.key, so one test covers
        // both; ids in the disallowed range always get their exact key, the merge needs it.
        if (!sp.greedy && !maybe_banned && fmaf(z, sp.inv_temperature, gumbel_from_bits(bits[q]) + 2e-3f) < ok.key) continue;
        const float k = race_key(z, bits[q], sp);
        if (better(k, id, all.key, all.idx)) { all.key = k; all.idx = id; }
        bool banned = false;
        if (maybe_banned)
            for (int t = 0; t < sp.n_banned; ++t)
                if (sp.banned[t] == id) { banned = true; e.banned_keys[(int64_t)m * 64 + t] = k; }
        if (!banned && better(k, id, ok.key, ok.idx)) { ok.key = k; ok.idx = id; }
    }
}

// --------------------------------------------------------------------------------- tiled GEMM
// C[M][N] = sum over operands of A[M][K] * W[N][K]^T, 256 threads, BMxBN tile, 16-deep K slabs,
// (BM/16)x(BN/16) accumulators per thread arranged as float4 groups 64 apart.
template <int BM, int BN, typename Epi>
__global__ void __launch_bounds__(256, 2) gemm_f32_kernel(Operand opa, Operand opb, int nops, int M, int N, Epi epi) {
    constexpr int BK = 16;
    constexpr int TM = BM / 16, TN = BN / 16;
    constexpr int GM = TM / 4, GN = TN / 4;
    constexpr int LA = BM * (BK / 4) / 256, LW = BN * (BK / 4) / 256;
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Ws[BK][BN + 4];

    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    for (int o = 0; o < nops; ++o) {
        const Operand& op = (o == 0) ? opa : opb;
        const float* arow[LA];
        const float* wrow[LW];
        int ar[LA], ak[LA], wr[LW], wk[LW];
#pragma unroll
        for (int i = 0; i < LA; ++i) {
            int id = tid + i * 256;
            ar[i] = id >> 2;
            ak[i] = (id & 3) * 4;
            arow[i] = operand_row(op, m0 + ar[i], M);
        }
#pragma unroll
        for (int i = 0; i < LW; ++i) {
            int id = tid + i * 256;
            wr[i] = id >> 2;
            wk[i] = (id & 3) * 4;
            wrow[i] = (n0 + wr[i] < N) ? op.W + (int64_t)(n0 + wr[i]) * op.ldw : nullptr;
        }
        float4 ra[LA], rw[LW];
        auto fetch = [&](int k0) {
#pragma unroll
            for (int i = 0; i < LA; ++i)
                ra[i] = (arow[i] != nullptr && k0 + ak[i] < op.K)
                            ? *reinterpret_cast<const float4*>(arow[i] + k0 + ak[i])
                            : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < LW; ++i)
                rw[i] = (wrow[i] != nullptr && k0 + wk[i] < op.K)
                            ? __ldg(reinterpret_cast<const float4*>(wrow[i] + k0 + wk[i]))
                            : make_float4(0.f, 0.f, 0.f, 0.f);
        };
        fetch(0);
        for (int k0 = 0; k0 < op.K; k0 += BK) {
            __syncthreads();
#pragma unroll
            for (int i = 0; i < LA; ++i) {
                As[ak[i] + 0][ar[i]] = ra[i].x;
                As[ak[i] + 1][ar[i]] = ra[i].y;
                As[ak[i] + 2][ar[i]] = ra[i].z;
                As[ak[i] + 3][ar[i]] = ra[i].w;
            }
#pragma unroll
            for (int i = 0; i < LW; ++i) {
                Ws[wk[i] + 0][wr[i]] = rw[i].x;
                Ws[wk[i] + 1][wr[i]] = rw[i].y;
                Ws[wk[i] + 2][wr[i]] = rw[i].z;
                Ws[wk[i] + 3][wr[i]] = rw[i].w;
            }
            __syncthreads();
            if (k0 + BK < op.K) fetch(k0 + BK);
#pragma unroll
            for (int k = 0; k < BK; ++k) {
                float a[TM], w[TN];
#pragma unroll
                for (int g = 0; g < GM; ++g) {
                    float4 t = *reinterpret_cast<const float4*>(&As[k][g * 64 + ty * 4]);
                    a[g * 4 + 0] = t.x; a[g * 4 + 1] = t.y; a[g * 4 + 2] = t.z; a[g * 4 + 3] = t.w;
                }
#pragma unroll
                for (int g = 0; g < GN; ++g) {
                    float4 t = *reinterpret_cast<const float4*>(&Ws[k][g * 64 + tx * 4]);
                    w[g * 4 + 0] = t.x; w[g * 4 + 1] = t.y; w[g * 4 + 2] = t.z; w[g * 4 + 3] = t.w;
                }
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
            }
        }
    }

    if constexpr (std::is_same<Epi, EpiRace>::value) {
        // per row: race over this thread's columns, then over the 16 threads (tx) that share the row -- they are the 16 lanes of a
        // half warp -- and one candidate pair per (row, tile) goes to `part`
        const SampleParams& sp = epi.sp;
#pragma unroll
        for (int gm = 0; gm < GM; ++gm)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int m = m0 + gm * 64 + ty * 4 + i;
                KeyIdx all{-INFINITY, 0x7fffffff}, ok{-INFINITY, 0x7fffffff};
                if (m < M) {
#pragma unroll
                    for (int gn = 0; gn < GN; ++gn) {
                        const float v[4] = {acc[gm * 4 + i][gn * 4 + 0], acc[gm * 4 + i][gn * 4 + 1], acc[gm * 4 + i][gn * 4 + 2], acc[gm * 4 + i][gn * 4 + 3]};
                        race_group(epi, sp, m, n0 + gn * 64 + tx * 4, N, v, all, ok);
                    }
                }
#pragma unroll
                for (int sft = 8; sft > 0; sft >>= 1) {
                    float k = __shfl_xor_sync(0xffffffffu, all.key, sft);
                    int id = __shfl_xor_sync(0xffffffffu, all.idx, sft);
                    if (better(k, id, all.key, all.idx)) { all.key = k; all.idx = id; }
                    k = __shfl_xor_sync(0xffffffffu, ok.key, sft);
                    id = __shfl_xor_sync(0xffffffffu, ok.idx, sft);
                    if (better(k, id, ok.key, ok.idx)) { ok.key = k; ok.idx = id; }
                }
                if (tx == 0 && m < M)
                    epi.part[(int64_t)m * epi.tiles_n + blockIdx.x] = make_float4(all.key, __int_as_float(all.idx), ok.key, __int_as_float(ok.idx));
            }
        return;
    } else {
#pragma unroll
    for (int gm = 0; gm < GM; ++gm)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int m = m0 + gm * 64 + ty * 4 + i;
            if (m >= M) continue;
#pragma unroll
            for (int gn = 0; gn < GN; ++gn) {
                int n = n0 + gn * 64 + tx * 4;
                float v[4] = {acc[gm * 4 + i][gn * 4 + 0], acc[gm * 4 + i][gn * 4 + 1], acc[gm * 4 + i][gn * 4 + 2],
                              acc[gm * 4 + i][gn * 4 + 3]};
                epilogue_one(epi, m, n, N, v);
            }
        }
    }
}

// ------------------------------------------------------------------------- skinny (few streams)
// One warp owns 4 consecutive weight rows (the four gates of one hidden unit, or 4 vocabulary
// rows) and all MB streams.  Lanes stride over K with 128-bit loads; every weight byte is read
// exactly once, so with few streams the kernel is bound by the weight stream from HBM/L2.
template <int MB, typename Epi>
__global__ void __launch_bounds__(256) skinny_f32_kernel(Operand opa, Operand opb, int nops, int M, int N, Epi epi) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n0 = warp * 4;
    if (n0 >= N) return;
    float acc[4][MB];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int m = 0; m < MB; ++m) acc[r][m] = 0.f;

    for (int o = 0; o < nops; ++o) {
        const Operand& op = (o == 0) ? opa : opb;
        const float4* w[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            int n = min(n0 + r, N - 1);
            w[r] = reinterpret_cast<const float4*>(op.W + (int64_t)n * op.ldw);
        }
        const float4* a[MB];
#pragma unroll
        for (int m = 0; m < MB; ++m) a[m] = reinterpret_cast<const float4*>(operand_row(op, min(m, M - 1), M));
        const int K4 = op.K >> 2;
        for (int k = lane; k < K4; k += 32) {
            float4 wv[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) wv[r] = __ldg(w[r] + k);
#pragma unroll
            for (int m = 0; m < MB; ++m) {
                float4 av = a[m][k];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    acc[r][m] = fmaf(av.x, wv[r].x, acc[r][m]);
                    acc[r][m] = fmaf(av.y, wv[r].y, acc[r][m]);
                    acc[r][m] = fmaf(av.z, wv[r].z, acc[r][m]);
                    acc[r][m] = fmaf(av.w, wv[r].w, acc[r][m]);
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int m = 0; m < MB; ++m) {
            float v = acc[r][m];
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
            acc[r][m] = v;
        }
    // lane m finishes stream m
    float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int m = 0; m < MB; ++m)
        if (lane == m) {
            v[0] = acc[0][m]; v[1] = acc[1][m]; v[2] = acc[2][m]; v[3] = acc[3][m];
        }
    if (lane < MB && lane < M) epilogue_one(epi, lane, n0, N, v);
}

template <typename Epi>
static int launch_gemm(const Operand& a, const Operand* b, int M, int N, const Epi& epi, cudaStream_t st) {
    if (M <= 0 || N <= 0) return AWD_OK;
    Operand ob = b ? *b : a;
    int nops = b ? 2 : 1;
    if (M <= 8) {
        int warps = (int)cdiv(N, 4);
        unsigned blocks = (unsigned)cdiv(warps, 8);
        if (M == 1) skinny_f32_kernel<1, Epi><<<blocks, 256, 0, st>>>(a, ob, nops, M, N, epi);
        else if (M == 2) skinny_f32_kernel<2, Epi><<<blocks, 256, 0, st>>>(a, ob, nops, M, N, epi);
        else if (M <= 4) skinny_f32_kernel<4, Epi><<<blocks, 256, 0, st>>>(a, ob, nops, M, N, epi);
        else skinny_f32_kernel<8, Epi><<<blocks, 256, 0, st>>>(a, ob, nops, M, N, epi);
    } else {
        // 128x128 tiles once they fill the 148 SMs about twice over, 64x64 below that
        int64_t big_tiles = cdiv(M, 128) * cdiv(N, 128);
        if (big_tiles >= 2 * 148) {
            dim3 grid((unsigned)cdiv(N, 128), (unsigned)cdiv(M, 128));
            gemm_f32_kernel<128, 128, Epi><<<grid, 256, 0, st>>>(a, ob, nops, M, N, epi);
        } else {
            dim3 grid((unsigned)cdiv(N, 64), (unsigned)cdiv(M, 64));
            gemm_f32_kernel<64, 64, Epi><<<grid, 256, 0, st>>>(a, ob, nops, M, N, epi);
        }
    }
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_gemm_store_f32(const Operand& a, int M, int N, const EpiStore& epi, cudaStream_t st) {
    return launch_gemm<EpiStore>(a, nullptr, M, N, epi, st);
}
int launch_gemm_cell_f32(const Operand& a, const Operand* b, int M, int N, const EpiCell& epi, cudaStream_t st) {
    return launch_gemm<EpiCell>(a, b, M, N, epi, st);
}

// ------------------------------------------------------------------------------------ sampling
// One CTA per stream.  Tracks the best key overall and the best key over allowed ids; the
// reference rule "first of the n_draws candidates that is allowed, else the first candidate"
// (main.py:63-68) equals: allowed-best unless n_draws or more disallowed ids outrank it.
__global__ void __launch_bounds__(256) sample_f32_kernel(const float* __restrict__ logits, int64_t ld, int V,
                                                          SampleParams sp, int64_t* __restrict__ out_a,
                                                          int64_t* __restrict__ out_b) {
    const int b = blockIdx.x;
    const float* z = logits + (int64_t)b * ld;
    const int ban_lo = sp.n_banned ? sp.banned[0] : V, ban_hi = sp.n_banned ? sp.banned[sp.n_banned - 1] : -1;
    KeyIdx all{-INFINITY, 0x7fffffff}, ok{-INFINITY, 0x7fffffff};
    const bool vec = ((ld & 3) == 0) && ((((uintptr_t)logits) & 15) == 0);
    for (int c4 = threadIdx.x; c4 * 4 < V; c4 += blockDim.x) {
        int c = c4 * 4;
        float zv[4];
        if (vec && c + 3 < V) {
            float4 t = *reinterpret_cast<const float4*>(z + c);
            zv[0] = t.x; zv[1] = t.y; zv[2] = t.z; zv[3] = t.w;
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) zv[q] = (c + q < V) ? z[c + q] : -INFINITY;
        }
        uint4 r = make_uint4(0, 0, 0, 0);
        if (!sp.greedy) r = philox4x32_10(make_uint4((uint32_t)c4, (uint32_t)b, sp.step_lo, sp.step_hi), sp.seed_lo, sp.seed_hi);
        uint32_t bits[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            int id = c + q;
            if (id >= V) continue;
            float k = race_key(zv[q], bits[q], sp);
            if (better(k, id, all.key, all.idx)) { all.key = k; all.idx = id; }
            bool banned = false;
            if (id >= ban_lo && id <= ban_hi)
                for (int t = 0; t < sp.n_banned; ++t) banned |= (sp.banned[t] == id);
            if (!banned && better(k, id, ok.key, ok.idx)) { ok.key = k; ok.idx = id; }
        }
    }
    __shared__ KeyIdx s_all[8], s_ok[8];
    all = warp_best(all);
    ok = warp_best(ok);
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_all[warp] = all; s_ok[warp] = ok; }
    __syncthreads();
    if (warp == 0) {
        all = (lane < 8) ? s_all[lane] : KeyIdx{-INFINITY, 0x7fffffff};
        ok = (lane < 8) ? s_ok[lane] : KeyIdx{-INFINITY, 0x7fffffff};
        all = warp_best(all);
        ok = warp_best(ok);
        if (lane == 0) {
            int pick = ok.idx;
            if (ok.idx == 0x7fffffff) {
                pick = all.idx;
            } else {
                int outrank = 0;
                for (int t = 0; t < sp.n_banned; ++t) {
                    int id = sp.banned[t];
                    if (id >= V) continue;
                    uint32_t bits = 0;
                    if (!sp.greedy) {
                        uint4 r = philox4x32_10(make_uint4((uint32_t)(id >> 2), (uint32_t)b, sp.step_lo, sp.step_hi),
                                                sp.seed_lo, sp.seed_hi);
                        int q = id & 3;
                        bits = q == 0 ? r.x : q == 1 ? r.y : q == 2 ? r.z : r.w;
                    }
                    float k = race_key(z[id], bits, sp);
                    outrank += better(k, id, ok.key, ok.idx) ? 1 : 0;
                }
                if (outrank >= sp.n_draws) pick = all.idx;
            }
            out_a[b] = pick;
            if (out_b != nullptr) out_b[b] = pick;
        }
    }
}

// One warp per stream: merge the per-tile candidate pairs of the fused decoder + race, apply the reference rule (main.py:63-68:
// the best allowed id unless n_draws or more disallowed ids outrank it, else the first candidate overall).
__global__ void __launch_bounds__(128) race_merge_f32_kernel(const float4* __restrict__ part, const float* __restrict__ banned_keys, int tiles_n,
                                                              int bs, int V, const SampleParams sp, int64_t* __restrict__ out) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= bs) return;
    KeyIdx all{-INFINITY, 0x7fffffff}, ok{-INFINITY, 0x7fffffff};
    for (int t = lane; t < tiles_n; t += 32) {
        const float4 q = __ldcg(part + (int64_t)row * tiles_n + t);
        if (better(q.x, __float_as_int(q.y), all.key, all.idx)) { all.key = q.x; all.idx = __float_as_int(q.y); }
        if (better(q.z, __float_as_int(q.w), ok.key, ok.idx)) { ok.key = q.z; ok.idx = __float_as_int(q.w); }
    }
    all = warp_best(all);
    ok = warp_best(ok);
    int outrank = 0;
    for (int t = lane; t < sp.n_banned; t += 32)
        if (sp.banned[t] < V) outrank += better(__ldcg(banned_keys + (int64_t)row * 64 + t), sp.banned[t], ok.key, ok.idx) ? 1 : 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) outrank += __shfl_xor_sync(0xffffffffu, outrank, s);
    if (lane == 0) out[row] = (ok.idx == 0x7fffffff || outrank >= sp.n_draws) ? all.idx : ok.idx;
}

int race_tiles_f32(int M, int N) {
    return (int)(cdiv(M, 128) * cdiv(N, 128) >= 2 * 148 ? cdiv(N, 128) : cdiv(N, 64));
}

int launch_gemm_race_f32(const Operand& a, int M, int N, const float* bias, const SampleParams& sp, float4* part, float* banned_keys,
                         int64_t* out, cudaStream_t st) {
    if (M <= 8) return set_error(AWD_EINVAL, "fused fp32 decoder + race: more than 8 streams (fewer take the skinny kernels)");
    EpiRace epi{bias, part, banned_keys, race_tiles_f32(M, N), sp};
    if (cdiv(M, 128) * cdiv(N, 128) >= 2 * 148) {
        dim3 grid((unsigned)cdiv(N, 128), (unsigned)cdiv(M, 128));
        gemm_f32_kernel<128, 128, EpiRace><<<grid, 256, 0, st>>>(a, a, 1, M, N, epi);
    } else {
        dim3 grid((unsigned)cdiv(N, 64), (unsigned)cdiv(M, 64));
        gemm_f32_kernel<64, 64, EpiRace><<<grid, 256, 0, st>>>(a, a, 1, M, N, epi);
    }
    AWD_LAUNCH_CHECK();
    race_merge_f32_kernel<<<(unsigned)cdiv(M, 4), 128, 0, st>>>(part, banned_keys, epi.tiles_n, M, N, sp, out);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_sample_f32(const float* logits, int64_t ld, int V, int bs, const SampleParams& sp, int64_t* out_a,
                      int64_t* out_b, cudaStream_t st) {
    if (bs <= 0) return AWD_OK;
    sample_f32_kernel<<<bs, 256, 0, st>>>(logits, ld, V, sp, out_a, out_b);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// ------------------------------------------------------------------------------ weight packing
// dst[r'][k] = src[r][k] for k < cols, 0 for cols <= k < ld_dst.  With H_interleave = H > 0 the
// rows are reordered from torch's gate-blocked order (g*H + j) to gate-interleaved (4*j + g).
__global__ void pack_rows_f32_kernel(const float* __restrict__ src, int rows, int cols, float* __restrict__ dst,
                                     int ld_dst, int H) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t r = gid / ld_dst;
    int k = (int)(gid - r * ld_dst);
    if (r >= rows) return;
    int64_t rd = r;
    if (H > 0) {
        int g = (int)(r / H), j = (int)(r - (int64_t)g * H);
        rd = 4 * (int64_t)j + g;
    }
    dst[rd * ld_dst + k] = (k < cols) ? src[r * cols + k] : 0.f;
}

int launch_pack_rows_f32(const float* src, int rows, int cols, float* dst, int ld_dst, int H_interleave,
                         cudaStream_t st) {
    int64_t total = (int64_t)rows * ld_dst;
    if (total == 0) return AWD_OK;
    pack_rows_f32_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, rows, cols, dst, ld_dst, H_interleave);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

__global__ void bias_sum_f32_kernel(const float* __restrict__ b_ih, const float* __restrict__ b_hh, int H,
                                    float* __restrict__ dst) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= 4 * H) return;
    int g = r / H, j = r - g * H;
    dst[4 * j + g] = b_ih[r] + b_hh[r];
}

int launch_bias_sum_f32(const float* b_ih, const float* b_hh, int H, float* dst, cudaStream_t st) {
    bias_sum_f32_kernel<<<(unsigned)cdiv(4 * H, 256), 256, 0, st>>>(b_ih, b_hh, H, dst);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

}  // namespace awd
