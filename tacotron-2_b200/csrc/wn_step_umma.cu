// wn_step_umma.cu - the two GEMMs of a residual layer of WaveNet.step (wavenet.py:650-721, modules.py:305-325, :471-521
// with is_incremental = False; BASELINE.json config 5) on the sm_100a tensor cores: tcgen05.mma, accumulators in TMEM.
//
//   gate     : D[128 positions x 2*NT]  = Xcat[128 x Kcat] * Wg[2*NT x Kcat]^T,  Xcat(n, :) = [x(t-2d) | x(t-d) | c(t) | x(t)]
//              epilogue bias, tanh * sigmoid (modules.py:510)                                      -> z [N][H]
//   out/skip : D[128 positions x (R+S)] = z[128 x H] * [W_out ; W_skip]^T, epilogue bias, residual add / scale
//              (modules.py:515-520) -> x' [N][R], and the skip recursion of wavenet.py:833-836     -> skips [N][S]
//
// Precision.  The criterion of this row is fp32 parity with the sample-by-sample engines (1e-3 relative on the
// distribution parameters after L layers), which one bf16 or tf32 pass does not meet.  Every fp32 operand is therefore
// split into two bf16 numbers, v = hi + lo (16 significant bits), and a product is accumulated as three tensor-core
// passes hi*hi + hi*lo + lo*hi into the same fp32 TMEM accumulator (the dropped lo*lo term is below 2^-16 relative).
// Weights are split once on the host; activations are split while they are staged into shared memory.
//
// One CTA = one tile of 128 positions (MMA M = 128, one TMEM lane per position) x 2*NT or R+S <= 256 columns (MMA N),
// 256 threads, 2 CTAs per SM (96 KB of shared memory and 256 TMEM columns each) so that one CTA stages while the other
// one's MMAs run.  K is walked in blocks of 64 (16 for the tail of the 80 conditioning channels); per block
//   * thread 0 asks for the block's packed weights ([hi | lo] x [bk/8][N][8] bf16, exactly the shared-memory image) with
//     bulk global->shared copies counted on an mbarrier,
//   * all threads gather their position's 32 activations (loaded one block ahead into registers), split them and store
//     them in the K-major core-matrix layout of wn_umma.cuh, fence.proxy.async, barrier,
//   * thread 0 waits for the weights and issues 3 x bk/16 tcgen05.mma, then tcgen05.commit on the "done" mbarrier that
//     everybody waits on before overwriting the operands.
// The epilogue reads the accumulator with tcgen05.ld (32 lanes x 16 columns per instruction): a thread owns one
// position and half of the columns.
#include <algorithm>
#include <array>
#include <cstdlib>
#include <cstring>

#include "wn_common.cuh"
#include "wn_umma.cuh"

namespace {

constexpr int UM = 128, UBK = 32, UMAXBLK = 32, NSTAGE = 2;
constexpr int USTAGERS = 256, UTHREADS = USTAGERS + 32;   // 8 staging / epilogue warps + the issuing warp
constexpr int LBO_A = UM * 16 + 16;                        // k-chunk stride of the activation operand (16 B of padding: the
                                                           // 8 chunks a quarter warp stores land in 8 different bank groups)
constexpr int A_HALF_BYTES = (UBK / 8) * LBO_A;            // one of A_hi / A_lo
constexpr int B_OFF = 2 * A_HALF_BYTES;
constexpr int STAGE_BYTES = B_OFF + 2 * 256 * UBK * 2;     // + B_hi, B_lo at N = 256
constexpr int U_SMEM = NSTAGE * STAGE_BYTES;
constexpr int STG_LD = 132;                                // row stride (floats) of the epilogue staging tile [128][128]
static_assert(UBK == 32, "the staging threads are mapped for blocks of 32 (tail 16)");
static_assert(UM * STG_LD * 4 <= U_SMEM, "the epilogue staging tile reuses the operand stages");

enum { UEPI_GATE = 0, UEPI_OUTSKIP = 1 };

struct UBlock {
    int seg;        // 0: src row of position n - back (zeros before the utterance starts), 1: cond row of position n
    int koff, bk, back;
    unsigned w_off; // byte offset of the block's packed weights inside one column tile's weight set
};

struct UArgs {
    const unsigned char *W;      // packed bf16 weights of this GEMM: [column tile][block][hi | lo][bk/8][ncols][8]
    long long tile_w_bytes;
    const float *bias;           // [column tile][ncols]
    const float *src, *cond;
    int src_ld, cond_ld;
    long long N;
    int T, ncols, nblk;
    UBlock blk[UMAXBLK];
    float *out, *skips;
    const float *xin;
    int H, R, S, NT, first_layer, legacy, residual_legacy;
    int dbg;
};

// tanh(t) * sigmoid(s) (modules.py:510) with the hardware exponential and reciprocal (ex2.approx / rcp.approx, relative
// error ~1e-6).  The sample-by-sample engines use the bit-reproducible functions of include/wn_math.h, ~3x the
// instructions; with those the gate epilogue, not the tensor pipe, bounded this kernel (ncu: 59 % of the issue slots).
// This path is compared with them by tolerance (1e-3, observed ~2e-5), so it can afford the fast forms.  The small-|t|
// polynomial is the one of wn_tanhf, which keeps the relative accuracy of tanh near zero.
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float gate_fast(float t, float s) {
    constexpr float LOG2E = 1.4426950408889634f;
    const float sg = rcp_approx(1.0f + ex2_approx(-LOG2E * s));
    const float at = fabsf(t);
    float th = 1.0f - 2.0f * rcp_approx(ex2_approx((2.0f * LOG2E) * at) + 1.0f);   // exp overflow -> rcp(inf) = 0 -> 1
    const float z = t * t;
    float p = -5.70498872745e-3f;
    p = __fmaf_rn(p, z, 2.06390887954e-2f);
    p = __fmaf_rn(p, z, -5.37397155531e-2f);
    p = __fmaf_rn(p, z, 1.33314422036e-1f);
    p = __fmaf_rn(p, z, -3.33332819422e-1f);
    th = copysignf(th, t);
    if (at < 0.625f) th = __fmaf_rn(p * z, t, t);
    return th * sg;
}

template <int EPI>
__global__ void __launch_bounds__(UTHREADS, 2) umma_layer_kernel(const __grid_constant__ UArgs a) {
    extern __shared__ __align__(128) unsigned char usmem[];
    __shared__ __align__(8) unsigned long long bars[3 * NSTAGE];   // full (weights landed) | ready (activations staged) | done (MMAs read the stage)
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar0 = umma::smem_addr(&bars[0]);
    auto full = [&](int s) { return bar0 + 8u * (uint32_t)s; };
    auto ready = [&](int s) { return bar0 + 8u * (uint32_t)(NSTAGE + s); };
    auto done = [&](int s) { return bar0 + 8u * (uint32_t)(2 * NSTAGE + s); };
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) { umma::mbar_init(full(s), 1); umma::mbar_init(ready(s), USTAGERS); umma::mbar_init(done(s), 1); }
        umma::mbar_fence_init();
    }
    if (warp == 8) umma::tmem_alloc(umma::smem_addr(&tmem_slot), 256);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_slot;

    const long long n0 = (long long)blockIdx.x * UM;
    const int ty = blockIdx.y, ncols = a.ncols, nblk = a.nblk;
    const uint32_t smem0 = umma::smem_addr(usmem);

    if (warp == 8) {
        // ---- issuing warp: weights by bulk copy, tcgen05.mma, commit ----------------------------------------------------
        if (lane == 0) {
            const unsigned char *Wt = a.W + (long long)ty * a.tile_w_bytes;
            const uint32_t idesc = umma::idesc_bf16(UM, ncols);
            const uint32_t lboB = (uint32_t)ncols * 16;
            for (int b = 0; b < nblk; ++b) {
                const int s = b % NSTAGE, bk = a.blk[b].bk;
                const uint32_t use = (uint32_t)(b / NSTAGE);
                if (b >= NSTAGE) umma::mbar_wait(done(s), (use - 1u) & 1u);   // the MMAs of block b - NSTAGE have read the stage
                const uint32_t st = smem0 + (uint32_t)s * STAGE_BYTES, Bs_a = st + B_OFF;
                const uint32_t bytes = ((uint32_t)bk * (uint32_t)ncols * 4u) >> (a.dbg & 7);
                umma::mbar_expect_tx(full(s), bytes);
                const unsigned char *src = Wt + a.blk[b].w_off;
                for (uint32_t o = 0; o < bytes; o += 16384u)
                    umma::bulk_g2s(Bs_a + o, src + o, bytes - o < 16384u ? bytes - o : 16384u, full(s));
                umma::mbar_wait(ready(s), use & 1u);
                umma::mbar_wait(full(s), use & 1u);
                umma::fence_after_sync();
                const uint32_t Ah_a = st, Al_a = st + A_HALF_BYTES, Bl_a = Bs_a + (uint32_t)bk * (uint32_t)ncols * 2u;
                for (int ks = 0; ks < ((a.dbg & 512) ? 0 : (bk >> 4)); ++ks) {
                    const uint64_t ah = umma::smem_desc(Ah_a + ks * 2 * LBO_A, LBO_A, 128), al = umma::smem_desc(Al_a + ks * 2 * LBO_A, LBO_A, 128);
                    const uint64_t bh = umma::smem_desc(Bs_a + ks * 2 * lboB, lboB, 128), bl = umma::smem_desc(Bl_a + ks * 2 * lboB, lboB, 128);
                    umma::mma_bf16(tmem, ah, bh, idesc, (b > 0 || ks > 0) ? 1u : 0u);
                    umma::mma_bf16(tmem, ah, bl, idesc, 1u);
                    umma::mma_bf16(tmem, al, bh, idesc, 1u);
                }
                umma::mma_commit(done(s));
            }
        }
        __syncwarp();
    } else {
        // ---- staging warps: gather, split into bf16 (hi, lo), store in the K-major core-matrix layout -----------------------
        // thread = (8-float chunk c of a 32-wide block, rows r0 and r0 + 64): the lanes of a warp read whole 128-byte row
        // segments.  Everything that depends only on the thread (row pointers, position inside the utterance) is computed
        // once; a 16-wide tail block keeps the mapping with the chunks c >= 2 idle.
        const int c = tid & 3, r0 = tid >> 2;
        const float *srow[2], *crow[2];
        int trow[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const long long n = n0 + r0 + 64 * j;
            const bool ok = n < a.N;
            trow[j] = ok ? (int)(n % a.T) : -1;               // -1: no such position (tail tile) -> zeros
            srow[j] = a.src + (ok ? n : 0) * a.src_ld + 8 * c;
            crow[j] = a.cond ? a.cond + (ok ? n : 0) * a.cond_ld + 8 * c : nullptr;
        }
        const int soff = (c * LBO_A + r0 * 16);
        float4 pf[4];
        auto load_block = [&](int b) {
            const UBlock &k = a.blk[b];
            const bool c_ok = 8 * c < k.bk;
            const int back_off = k.koff - k.back * a.src_ld;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float *p = k.seg == 0 ? srow[j] + back_off : crow[j] + k.koff;
                float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                if (c_ok && trow[j] >= k.back && !(a.dbg & 256)) { v0 = __ldg(reinterpret_cast<const float4 *>(p)); v1 = __ldg(reinterpret_cast<const float4 *>(p + 4)); }
                pf[2 * j] = v0; pf[2 * j + 1] = v1;
            }
        };
        auto store_block = [&](int b, int s) {
            if (8 * c >= a.blk[b].bk) return;
            unsigned char *Ah = usmem + (size_t)s * STAGE_BYTES + soff, *Al = Ah + A_HALF_BYTES;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint4 hi, lo;
                umma::split8(pf[2 * j], pf[2 * j + 1], hi, lo);
                *reinterpret_cast<uint4 *>(Ah + j * 64 * 16) = hi;
                *reinterpret_cast<uint4 *>(Al + j * 64 * 16) = lo;
            }
        };
        load_block(0);
        for (int b = 0; b < nblk; ++b) {
            const int s = b % NSTAGE;
            if (b >= NSTAGE) umma::mbar_wait(done(s), (uint32_t)(b / NSTAGE - 1) & 1u);
            store_block(b, s);
            umma::fence_proxy_async();
            umma::mbar_arrive(ready(s));
            if (b + 1 < nblk) load_block(b + 1);   // in flight while the tensor core works
        }
        // the last commit covers every MMA of the tile
        umma::mbar_wait(done((nblk - 1) % NSTAGE), (uint32_t)((nblk - 1) / NSTAGE) & 1u);
        umma::fence_after_sync();
    }

    // ---- epilogue: TMEM -> registers (thread = position 32q + lane, half of the columns) -> staging tile in shared memory
    //      -> coalesced rows (one warp per position row) -------------------------------------------------------------------
    float *stg = reinterpret_cast<float *>(usmem);
    const int q = warp & 3, half = (warp >> 2) & 1;
    const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
    const float *bias = a.bias + (size_t)ty * ncols;
    if (a.dbg & 1024) { umma::fence_before_sync(); __syncthreads(); if (warp == 8) umma::tmem_dealloc(tmem, 256); return; }
    if (EPI == UEPI_GATE) {
        const int NT = a.NT;
        if (warp < 8) {
            float *srow = stg + (size_t)(32 * q + lane) * STG_LD;
            for (int c = half * (NT / 2); c < (half + 1) * (NT / 2); c += 16) {
                float ta[16], sg[16];
                umma::tmem_ld16(tq + (uint32_t)c, ta);
                umma::tmem_ld16(tq + (uint32_t)(NT + c), sg);
                umma::tmem_ld_wait();
#pragma unroll
                for (int i4 = 0; i4 < 4; ++i4) {
                    const float4 bt = __ldg(reinterpret_cast<const float4 *>(bias + c + 4 * i4));
                    const float4 bs = __ldg(reinterpret_cast<const float4 *>(bias + NT + c + 4 * i4));
                    *reinterpret_cast<float4 *>(srow + c + 4 * i4) =
                        make_float4(gate_fast(ta[4 * i4] + bt.x, sg[4 * i4] + bs.x), gate_fast(ta[4 * i4 + 1] + bt.y, sg[4 * i4 + 1] + bs.y),
                                    gate_fast(ta[4 * i4 + 2] + bt.z, sg[4 * i4 + 2] + bs.z), gate_fast(ta[4 * i4 + 3] + bt.w, sg[4 * i4 + 3] + bs.w));
                }
            }
        }
        __syncthreads();
        if (warp < 8 && 4 * lane < NT)
            for (int r = warp; r < UM; r += 8) {
                const long long n = n0 + r;
                if (n < a.N)
                    *reinterpret_cast<float4 *>(a.out + n * a.H + ty * NT + 4 * lane) = *reinterpret_cast<const float4 *>(stg + (size_t)r * STG_LD + 4 * lane);
            }
    } else {
        for (int g = 0; g < ncols / 128; ++g) {
            if (g > 0) __syncthreads();
            if (warp < 8) {
                float *srow = stg + (size_t)(32 * q + lane) * STG_LD;
                for (int c = half * 64; c < half * 64 + 64; c += 16) {
                    float v[16];
                    umma::tmem_ld16(tq + (uint32_t)(g * 128 + c), v);
                    umma::tmem_ld_wait();
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4)
                        *reinterpret_cast<float4 *>(srow + c + 4 * i4) = make_float4(v[4 * i4], v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                }
            }
            __syncthreads();
            if (warp < 8) {
                const int m = ty * ncols + g * 128 + 4 * lane;            // first of this lane's 4 output rows of the GEMM
                const float4 bb = __ldg(reinterpret_cast<const float4 *>(bias + g * 128 + 4 * lane));
                // what is added to a row: the layer input (residual, modules.py:515-520) or the running skip sum
                // (wavenet.py:833-836); 8 rows of loads are in flight before the first one is used
                const bool is_out = m < a.R, has_add = is_out || !a.first_layer;
                const float *addp = is_out ? a.xin + m : a.skips + (m - a.R);
                float *dstp = is_out ? a.out + m : a.skips + (m - a.R);
                const int ld = is_out ? a.R : a.S;
                const bool scale = is_out ? a.residual_legacy != 0 : (a.legacy != 0 && !a.first_layer);
                for (int rb = warp; rb < UM; rb += 64) {
                    float4 add[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const long long n = n0 + rb + 8 * k;
                        add[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (has_add && n < a.N) add[k] = *reinterpret_cast<const float4 *>(addp + n * ld);
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int r = rb + 8 * k;
                        const long long n = n0 + r;
                        if (n >= a.N) continue;
                        const float4 v = *reinterpret_cast<const float4 *>(stg + (size_t)r * STG_LD + 4 * lane);
                        float4 y = make_float4(__fadd_rn(v.x, bb.x), __fadd_rn(v.y, bb.y), __fadd_rn(v.z, bb.z), __fadd_rn(v.w, bb.w));
                        if (has_add) {
                            if (is_out) { y.x = __fadd_rn(y.x, add[k].x); y.y = __fadd_rn(y.y, add[k].y); y.z = __fadd_rn(y.z, add[k].z); y.w = __fadd_rn(y.w, add[k].w); }
                            else { y.x = __fadd_rn(add[k].x, y.x); y.y = __fadd_rn(add[k].y, y.y); y.z = __fadd_rn(add[k].z, y.z); y.w = __fadd_rn(add[k].w, y.w); }
                        }
                        if (scale) {
                            y.x = __fmul_rn(y.x, WN_SQRT_HALF); y.y = __fmul_rn(y.y, WN_SQRT_HALF);
                            y.z = __fmul_rn(y.z, WN_SQRT_HALF); y.w = __fmul_rn(y.w, WN_SQRT_HALF);
                        }
                        *reinterpret_cast<float4 *>(dstp + n * ld) = y;
                    }
                }
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 8) umma::tmem_dealloc(tmem, 256);
}

unsigned short bf16_rn_host(float f) {
    unsigned u;
    memcpy(&u, &f, 4);
    if ((u & 0x7f800000u) == 0x7f800000u) return (unsigned short)(u >> 16);   // inf / nan: truncate
    return (unsigned short)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
}
float bf16_to_float_host(unsigned short h) {
    const unsigned u = (unsigned)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

struct KBlockHost { int seg, tap, koff, bk, kcat; unsigned w_off; };

// cuts the segments of K into blocks of 32 (16 for a tail); w_off for `ncols` columns per tile
std::vector<KBlockHost> cut_blocks(const std::vector<std::array<int, 3>> &segs /* seg, tap, len */, int ncols, unsigned *total) {
    std::vector<KBlockHost> out;
    unsigned off = 0;
    int kcat = 0;
    for (const auto &s : segs) {
        for (int k = 0; k < s[2];) {
            int bk = UBK;                                    // 32, then 16 for a tail (every length is a multiple of 16)
            while (bk > s[2] - k) bk >>= 1;
            out.push_back({s[0], s[1], k, bk, kcat + k, off});
            off += (unsigned)bk * (unsigned)ncols * 4u;
            k += bk;
        }
        kcat += s[2];
    }
    *total = off;
    return out;
}

}  // namespace

struct WnUmmaPlan {
    unsigned char *d_w = nullptr;
    float *d_bias = nullptr;
    int NT = 0, gate_ncols = 0, gate_tiles = 0, os_ncols = 0, os_tiles = 0;
    unsigned gate_tile_bytes = 0, os_tile_bytes = 0;
    std::vector<KBlockHost> gate_blk, os_blk;
    std::vector<size_t> gate_w, os_w, gate_b, os_b;   // per layer: byte offsets into d_w, float offsets into d_bias
};

void wn_umma_free(wn_handle *h) {
    WnUmmaPlan *u = h->umma;
    if (!u) return;
    if (u->d_w) cudaFree(u->d_w);
    if (u->d_bias) cudaFree(u->d_bias);
    delete u;
    h->umma = nullptr;
}

bool wn_umma_supported(const wn_handle *h) {
    const WnModelDev &m = h->m;
    if (getenv("WN_STEP_FFMA") && atoi(getenv("WN_STEP_FFMA"))) return false;   // fp32 FFMA2 GEMMs (wn_step_gemm.cu)
    if (!wn_gemm_supported(h)) return false;
    if (m.H % 64 || (m.R + m.S) % 128 || m.R % 16 || m.cin % 16) return false;
    if (3 * ((m.R + UBK - 1) / UBK) + (m.cin + UBK - 1) / UBK > UMAXBLK || (m.H + UBK - 1) / UBK > UMAXBLK) return false;
    return true;
}

static int umma_plan(wn_handle *h) {
    if (h->umma) return WN_OK;
    const WnModelDev &m = h->m;
    auto Wt = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    WnUmmaPlan *u = new WnUmmaPlan();
    const int R = m.R, S = m.S, H = m.H, G = 2 * H, cin = m.cin;
    u->NT = H % 128 == 0 ? 128 : 64;
    u->gate_ncols = 2 * u->NT; u->gate_tiles = H / u->NT;
    u->os_ncols = (R + S) % 256 == 0 ? 256 : 128; u->os_tiles = (R + S) / u->os_ncols;
    u->gate_blk = cut_blocks({{0, 0, R}, {0, 1, R}, {1, 0, cin}, {0, 2, R}}, u->gate_ncols, &u->gate_tile_bytes);
    u->os_blk = cut_blocks({{0, 2, H}}, u->os_ncols, &u->os_tile_bytes);
    std::vector<unsigned short> w;
    std::vector<float> bias;
    auto pack = [&](const std::vector<KBlockHost> &blks, int ncols, int tiles, unsigned tile_bytes, auto &&weight /* (k, row) */) {
        const size_t base = w.size();
        w.resize(base + (size_t)tiles * tile_bytes / 2, 0);
        for (int ty = 0; ty < tiles; ++ty)
            for (const auto &b : blks) {
                unsigned short *hi = w.data() + base + ((size_t)ty * tile_bytes + b.w_off) / 2, *lo = hi + (size_t)b.bk * ncols;
                for (int k = 0; k < b.bk; ++k)
                    for (int j = 0; j < ncols; ++j) {
                        const float v = weight(b.kcat + k, ty * ncols + j);
                        const unsigned short vh = bf16_rn_host(v), vl = bf16_rn_host(v - bf16_to_float_host(vh));
                        const size_t o = ((size_t)(k / 8) * ncols + j) * 8 + k % 8;
                        hi[o] = vh; lo[o] = vl;
                    }
            }
        return base * 2;
    };
    for (int l = 0; l < m.L; ++l) {
        const std::string pre = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        const float *cw = Wt(pre + "causal_conv/kernel"), *cb = Wt(pre + "causal_conv/bias");
        const float *iw = Wt(pre + "cin_conv/kernel"), *ib = Wt(pre + "cin_conv/bias");
        const float *ow = Wt(pre + "out_conv/kernel"), *ob = Wt(pre + "out_conv/bias");
        const float *sw = Wt(pre + "skip_conv/kernel"), *sb = Wt(pre + "skip_conv/bias");
        const int NT = u->NT;
        // packed gate row `row` of the layer: tile row / (2 NT), inside a tile NT tanh rows then the NT sigmoid rows of the
        // same channels (modules.py:494: first half of the conv output -> tanh)
        auto gate_col = [&](int row) { const int ty = row / (2 * NT), j = row % (2 * NT); return (j < NT ? 0 : H) + ty * NT + (j % NT); };
        u->gate_w.push_back(pack(u->gate_blk, u->gate_ncols, u->gate_tiles, u->gate_tile_bytes, [&](int k, int row) {
            const int col = gate_col(row);
            if (k < 2 * R) return cw[(size_t)k * G + col];
            if (k < 2 * R + cin) return iw[(size_t)(k - 2 * R) * G + col];
            return cw[(size_t)(k - cin) * G + col];
        }));
        u->gate_b.push_back(bias.size());
        for (int row = 0; row < G; ++row) { const int col = gate_col(row); bias.push_back((cb ? cb[col] : 0.0f) + (ib ? ib[col] : 0.0f)); }
        u->os_w.push_back(pack(u->os_blk, u->os_ncols, u->os_tiles, u->os_tile_bytes, [&](int k, int row) {
            return row < R ? ow[(size_t)k * R + row] : sw[(size_t)k * S + (row - R)];
        }));
        u->os_b.push_back(bias.size());
        for (int r = 0; r < R; ++r) bias.push_back(ob ? ob[r] : 0.0f);
        for (int s = 0; s < S; ++s) bias.push_back(sb ? sb[s] : 0.0f);
    }
    if (cudaMalloc((void **)&u->d_w, w.size() * 2) != cudaSuccess ||
        cudaMemcpy(u->d_w, w.data(), w.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMalloc((void **)&u->d_bias, bias.size() * sizeof(float)) != cudaSuccess ||
        cudaMemcpy(u->d_bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaFuncSetAttribute(umma_layer_kernel<UEPI_GATE>, cudaFuncAttributeMaxDynamicSharedMemorySize, U_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(umma_layer_kernel<UEPI_OUTSKIP>, cudaFuncAttributeMaxDynamicSharedMemorySize, U_SMEM) != cudaSuccess) {
        h->err = std::string("step tensor-core weights: ") + cudaGetErrorString(cudaGetLastError());
        if (u->d_w) cudaFree(u->d_w);
        if (u->d_bias) cudaFree(u->d_bias);
        delete u;
        return WN_ERR_NOMEM;
    }
    h->umma = u;
    return WN_OK;
}

// One residual layer: z = gate(x, c), then x' and the skip sum.  x, z, x_out, skips are position-major fp32.
int wn_umma_layer(wn_handle *h, int l, const float *x, const float *c_up, float *z, float *x_out, float *skips, long long N, int T,
                  cudaStream_t st) {
    int rc = umma_plan(h);
    if (rc) return rc;
    const WnUmmaPlan *u = h->umma;
    const WnModelDev &m = h->m;
    UArgs a;
    memset(&a, 0, sizeof(a));
    a.dbg = getenv("WN_UMMA_DBG") ? atoi(getenv("WN_UMMA_DBG")) : 0;
    a.N = N; a.T = T; a.H = m.H; a.R = m.R; a.S = m.S; a.NT = u->NT;
    a.legacy = m.legacy; a.residual_legacy = m.residual_legacy; a.first_layer = l == 0;
    const unsigned grid_x = (unsigned)((N + UM - 1) / UM);
    // gate
    a.W = u->d_w + u->gate_w[l]; a.tile_w_bytes = u->gate_tile_bytes; a.bias = u->d_bias + u->gate_b[l];
    a.src = x; a.src_ld = m.R; a.cond = c_up; a.cond_ld = m.cin; a.ncols = u->gate_ncols;
    a.nblk = (int)u->gate_blk.size();
    for (int i = 0; i < a.nblk; ++i) {
        const KBlockHost &b = u->gate_blk[i];
        a.blk[i] = {b.seg, b.koff, b.bk, b.seg == 0 ? (2 - b.tap) * m.dil[l] : 0, b.w_off};
    }
    a.out = z;
    umma_layer_kernel<UEPI_GATE><<<dim3(grid_x, (unsigned)u->gate_tiles), UTHREADS, U_SMEM, st>>>(a);
    WN_CUDA_TRY(h, cudaGetLastError());
    // out / skip
    a.W = u->d_w + u->os_w[l]; a.tile_w_bytes = u->os_tile_bytes; a.bias = u->d_bias + u->os_b[l];
    a.src = z; a.src_ld = m.H; a.cond = nullptr; a.cond_ld = 0; a.ncols = u->os_ncols;
    a.nblk = (int)u->os_blk.size();
    for (int i = 0; i < a.nblk; ++i) {
        const KBlockHost &b = u->os_blk[i];
        a.blk[i] = {0, b.koff, b.bk, 0, b.w_off};
    }
    a.out = x_out; a.skips = skips; a.xin = x;
    umma_layer_kernel<UEPI_OUTSKIP><<<dim3(grid_x, (unsigned)u->os_tiles), UTHREADS, U_SMEM, st>>>(a);
    WN_CUDA_TRY(h, cudaGetLastError());
    h->launches += 2;
    return WN_OK;
}
