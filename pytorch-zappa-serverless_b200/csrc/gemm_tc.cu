// Host side of the tcgen05 GEMM family (gemm_tc.cuh), the fused sampling epilogue and the small bf16
// helper kernels (weight packing, embedding gather, sampler finalisation).
#include <mutex>

#include "gemm_tc.cuh"
#include "tc_api.h"

namespace awd {
namespace tc {

// ------------------------------------------------------------------------------- tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static EncodeTiledFn encode_fn() {
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    });
    return g_encode;
}

// bf16 row-major [rows][cols] with row pitch ld elements; box = box_rows x 64 columns, 128B swizzle.
int make_tmap_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return set_error(AWD_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * 2) % 16)
        return set_error(AWD_EINVAL, "TMA operand must be 16-byte aligned with a 16-byte row pitch");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(AWD_ECUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
    return AWD_OK;
}

static int g_num_sms = 0;
static int num_sms() {
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

// ------------------------------------------------------------------- fused sampling epilogue
// Philox4x32-10, identical to the fp32 sampler (kernels_f32.cu) so both precisions run the same race.
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
__device__ __forceinline__ float race_key(float z, uint32_t bits, float inv_t, int greedy) {
    float k = z * inv_t;
    if (!greedy) {
        float u = ((float)(bits >> 8) + 0.5f) * (1.0f / 16777216.0f);
        k -= __logf(-__logf(u));
    }
    return k;
}
__device__ __forceinline__ bool better(float ka, int ia, float kb, int ib) { return (ka > kb) || (ka == kb && ia < ib); }

// Per (row, N-tile) partial of the race: best key overall and best key over allowed ids.  Keys of
// banned ids are parked in banned_keys[row][slot] for the reference's fallback rule.
struct TcEpiSample {
    SampleParams sp;
    const float* bias;       // decoder bias [V] or nullptr
    float4* part;            // [M][tiles_n] (key_all, idx_all, key_ok, idx_ok)
    float* banned_keys;      // [M][64]
    int tiles_n;
    struct State {
        float ka, ko;
        int ia, io;
    };
    __device__ __forceinline__ void tile_begin(State& s) const {
        s.ka = s.ko = -INFINITY;
        s.ia = s.io = 0x7fffffff;
    }
    __device__ __forceinline__ void chunk(State& s, int row, bool row_ok, int n, int N, const float (&v)[32]) const {
        if (!row_ok) return;
        float ka = s.ka, ko = s.ko;
        int ia = s.ia, io = s.io;
        const int ban_lo = sp.n_banned ? sp.banned[0] : N, ban_hi = sp.n_banned ? sp.banned[sp.n_banned - 1] : -1;
#pragma unroll
        for (int g4 = 0; g4 < 8; ++g4) {
            const int c = n + 4 * g4;
            uint4 r = make_uint4(0, 0, 0, 0);
            if (!sp.greedy)
                r = philox4x32_10(make_uint4((uint32_t)(c >> 2), (uint32_t)row, sp.step_lo, sp.step_hi), sp.seed_lo, sp.seed_hi);
            const uint32_t bits[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int id = c + q;
                if (id >= N) continue;
                float z = v[4 * g4 + q] + (bias ? __ldg(bias + id) : 0.f);
                float k = race_key(z, bits[q], sp.inv_temperature, sp.greedy);
                if (better(k, id, ka, ia)) { ka = k; ia = id; }
                bool banned = false;
                if (id >= ban_lo && id <= ban_hi) {
                    for (int t = 0; t < sp.n_banned; ++t)
                        if (sp.banned[t] == id) { banned = true; banned_keys[(int64_t)row * 64 + t] = k; }
                }
                if (!banned && better(k, id, ko, io)) { ko = k; io = id; }
            }
        }
        s.ka = ka; s.ko = ko; s.ia = ia; s.io = io;
    }
    __device__ __forceinline__ void tile_end(State& s, int row, bool row_ok, int tn) const {
        if (row_ok) part[(int64_t)row * tiles_n + tn] = make_float4(s.ka, __int_as_float(s.ia), s.ko, __int_as_float(s.io));
    }
};

// One warp per stream: merge the per-tile partials, apply "first allowed of the n_draws candidates,
// else the first candidate" (reference main.py:63-68), emit the id, and gather its embedding row
// (bf16, 128-bit loads) as the next step's layer-0 input.
__global__ void __launch_bounds__(256) sample_finalize_kernel(const float4* __restrict__ part, const float* __restrict__ banned_keys,
                                                               int tiles_n, int bs, SampleParams sp, int64_t* __restrict__ out,
                                                               const uint4* __restrict__ table, int row_vec4 /*uint4 per row*/,
                                                               uint4* __restrict__ xnext) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= bs) return;
    float ka = -INFINITY, ko = -INFINITY;
    int ia = 0x7fffffff, io = 0x7fffffff;
    for (int t = lane; t < tiles_n; t += 32) {
        float4 p = part[(int64_t)row * tiles_n + t];
        int pia = __float_as_int(p.y), pio = __float_as_int(p.w);
        if (better(p.x, pia, ka, ia)) { ka = p.x; ia = pia; }
        if (better(p.z, pio, ko, io)) { ko = p.z; io = pio; }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        float k = __shfl_xor_sync(0xffffffffu, ka, s);
        int i = __shfl_xor_sync(0xffffffffu, ia, s);
        if (better(k, i, ka, ia)) { ka = k; ia = i; }
        k = __shfl_xor_sync(0xffffffffu, ko, s);
        i = __shfl_xor_sync(0xffffffffu, io, s);
        if (better(k, i, ko, io)) { ko = k; io = i; }
    }
    int pick = io;
    if (io == 0x7fffffff) {
        pick = ia;
    } else {
        int outrank = 0;
        for (int t = lane; t < sp.n_banned; t += 32)
            outrank += better(banned_keys[(int64_t)row * 64 + t], sp.banned[t], ko, io) ? 1 : 0;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) outrank += __shfl_xor_sync(0xffffffffu, outrank, s);
        if (outrank >= sp.n_draws) pick = ia;
    }
    if (lane == 0) out[row] = pick;
    if (xnext != nullptr)
        for (int i = lane; i < row_vec4; i += 32) xnext[(int64_t)row * row_vec4 + i] = __ldg(table + (int64_t)pick * row_vec4 + i);
}

// ----------------------------------------------------------------------------- small kernels
// K1 in bf16 mode: out[n][:] = table[tok[n]][:], rows are Epad64 bf16 = 16-byte multiples.
__global__ void gather_bf16_kernel(const uint4* __restrict__ table, int row_vec4, int V, const int64_t* __restrict__ tok,
                                   int64_t n, uint4* __restrict__ out, int* err_flag) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t row = gid / row_vec4;
    int col = (int)(gid - row * row_vec4);
    if (row >= n) return;
    int64_t t = tok[row];
    if (t < 0 || t >= V) {
        if (col == 0) atomicExch(err_flag, 1);
        t = 0;
    }
    out[row * row_vec4 + col] = __ldg(table + t * row_vec4 + col);
}

// dst[r'][k0 + k] = bf16(src[r][k]) for k < cols; columns [k0+cols, k0+kpad) are zeroed.  H > 0 reorders
// rows from torch's gate-blocked (g*H + j) to gate-interleaved (4*j + g).
__global__ void pack_rows_bf16_kernel(const float* __restrict__ src, int rows, int cols, __nv_bfloat16* __restrict__ dst,
                                      int ld_dst, int k0, int kpad, int H) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t r = gid / kpad;
    int k = (int)(gid - r * kpad);
    if (r >= rows) return;
    int64_t rd = r;
    if (H > 0) {
        int g = (int)(r / H), j = (int)(r - (int64_t)g * H);
        rd = 4 * (int64_t)j + g;
    }
    dst[rd * ld_dst + k0 + k] = __float2bfloat16_rn(k < cols ? src[r * cols + k] : 0.f);
}

__global__ void f32_to_bf16_rows_kernel(const float* __restrict__ src, int64_t ld_src, int rows, int cols,
                                        __nv_bfloat16* __restrict__ dst, int ld_dst) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t r = gid / ld_dst;
    int k = (int)(gid - r * ld_dst);
    if (r >= rows) return;
    dst[r * ld_dst + k] = __float2bfloat16_rn(k < cols ? src[r * ld_src + k] : 0.f);
}

int launch_gather_bf16(const void* table, int ld, int V, const int64_t* tok, int64_t n, void* out, int* err_flag,
                       cudaStream_t st) {
    if (n == 0) return AWD_OK;
    int v4 = ld / 8;
    int64_t total = n * v4;
    gather_bf16_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>((const uint4*)table, v4, V, tok, n, (uint4*)out, err_flag);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_pack_rows_bf16(const float* src, int rows, int cols, void* dst, int ld_dst, int k0, int kpad, int H,
                          cudaStream_t st) {
    int64_t total = (int64_t)rows * kpad;
    if (total == 0) return AWD_OK;
    pack_rows_bf16_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, rows, cols, (__nv_bfloat16*)dst, ld_dst, k0, kpad, H);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_f32_to_bf16_rows(const float* src, int64_t ld_src, int rows, int cols, void* dst, int ld_dst, cudaStream_t st) {
    int64_t total = (int64_t)rows * ld_dst;
    if (total == 0) return AWD_OK;
    f32_to_bf16_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, ld_src, rows, cols, (__nv_bfloat16*)dst, ld_dst);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// --------------------------------------------------------------------------------- launchers
template <int BN, class Epi>
static int launch_tc(const TcOperands& o, int M, int N, const Epi& epi, cudaStream_t st) {
    using Cfg = TileCfg<BN>;
    static bool attr_set = false;
    auto kern = tc_gemm_kernel<BN, Epi>;
    if (!attr_set) {
        AWD_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        attr_set = true;
    }
    GemmShape g{};
    g.M = M; g.N = N;
    g.kb0 = o.k0 / BK; g.kb1 = o.k1 / BK;
    g.a0_row = o.a0_row; g.a1_row = o.a1_row; g.b_k0 = o.b_k0;
    g.tiles_m = (int)cdiv(M, BM); g.tiles_n = (int)cdiv(N, BN);
    CUtensorMap tmA0, tmA1, tmB;
    int rc = make_tmap_bf16(&tmA0, o.a0, o.a0_rows, o.k0, o.lda0, BM);
    if (rc) return rc;
    if (o.k1 > 0) rc = make_tmap_bf16(&tmA1, o.a1, o.a1_rows, o.k1, o.lda1, BM);
    else tmA1 = tmA0;
    if (rc) return rc;
    rc = make_tmap_bf16(&tmB, o.w, o.w_rows, o.w_cols, o.ldw, BN);
    if (rc) return rc;
    int grid = std::min(g.tiles_m * g.tiles_n, num_sms());
    kern<<<grid, NUM_THREADS, Cfg::SMEM_BYTES, st>>>(tmA0, tmA1, tmB, g, epi);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

static int check_operands(const TcOperands& o) {
    if (o.k0 <= 0 || o.k0 % BK || o.k1 % BK || o.k1 < 0) return set_error(AWD_EINVAL, "tc gemm: K extents must be multiples of 64");
    return AWD_OK;
}

int launch_tc_store(const TcOperands& o, int M, int N, float* C, int64_t ldc, const float* bias, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    TcEpiStore epi{C, ldc, bias};
    // wide tiles when there are enough of them to fill the SMs
    if (cdiv(M, BM) * cdiv(N, 256) >= num_sms()) return launch_tc<256, TcEpiStore>(o, M, N, epi, st);
    return launch_tc<128, TcEpiStore>(o, M, N, epi, st);
}

int launch_tc_cell(const TcOperands& o, int M, int N, const TcCellArgs& a, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    TcEpiCell epi{a.addend, a.ld_add, a.bias, a.c, a.h, a.ld_state, (__nv_bfloat16*)a.hb, a.ld_hb, a.raw_out, a.ld_raw};
    if (cdiv(M, BM) * cdiv(N, 256) >= num_sms() - 8) return launch_tc<256, TcEpiCell>(o, M, N, epi, st);
    return launch_tc<128, TcEpiCell>(o, M, N, epi, st);
}

int launch_tc_sample(const TcOperands& o, int M, int N, const SampleParams& sp, const float* bias, float* part,
                     float* banned_keys, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    TcEpiSample epi{};
    epi.sp = sp; epi.bias = bias; epi.part = (float4*)part; epi.banned_keys = banned_keys;
    epi.tiles_n = tc_sample_tiles(N);
    return launch_tc<256, TcEpiSample>(o, M, N, epi, st);
}

int tc_sample_tiles(int N) { return (int)cdiv(N, 256); }

int launch_sample_finalize(const float* part, const float* banned_keys, int tiles_n, int bs, const SampleParams& sp,
                           int64_t* out, const void* table, int ld_table, void* xnext, cudaStream_t st) {
    if (bs <= 0) return AWD_OK;
    sample_finalize_kernel<<<(unsigned)cdiv(bs, 8), 256, 0, st>>>((const float4*)part, banned_keys, tiles_n, bs, sp, out,
                                                                   (const uint4*)table, ld_table / 8, (uint4*)xnext);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

}  // namespace tc
}  // namespace awd
