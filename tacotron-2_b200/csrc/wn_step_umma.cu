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
// Persistent kernel, one CTA per SM (215 KB of shared memory, all 512 TMEM columns), 16 warps with fixed roles; a work
// item is a tile of 128 positions (MMA M = 128, one TMEM lane per position) x 2*NT or R+S <= 256 columns (MMA N):
//   * copy thread: the packed weights of a K block ([hi | lo] x [bk/8][N][8] bf16, exactly the shared-memory image, L2
//     resident) go into one of 3 operand stages by bulk global->shared copies counted on the stage's "full" mbarrier, as
//     soon as the MMAs that last read the stage have completed ("done");
//   * 6 staging warps in two groups that take alternate K blocks (32 wide, 16 for the tail of the 80 conditioning
//     channels): gather the positions' activations (requested one block ahead into registers), split, store in the
//     K-major core-matrix layout of wn_umma.cuh, fence.proxy.async, arrive on "ready";
//   * MMA thread: waits for "ready" and "full", issues 3 x bk/16 tcgen05.mma into accumulator (tile & 1), tcgen05.commit
//     on "done"; after the last block of a tile a second commit on "tfull";
//   * 8 epilogue warps: wait for "tfull", tcgen05.ld (32 lanes x 16 columns per instruction; a thread owns one position and
//     half of the columns), gate / bias in registers, release the accumulator ("tempty"), pass the tile through a
//     shared-memory staging tile so that one warp writes (and, for the residual and the skip sum, reads) whole 512-byte
//     rows - while the other warps already work on the next tile in the other accumulator.
// Measured (profiles/README_r1.md): C5 253 ms (fp32 FFMA2 GEMMs) -> 46 ms with this kernel (the out/skip launch moves
// 5 x N x 512 bytes for one K = 128 GEMM and runs at ~72 % of the HBM roof) -> 38 ms with umma_fused_layer_kernel below,
// which keeps z on the SM (default block).
#include <algorithm>
#include <array>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "wn_common.cuh"
#include "wn_umma.cuh"

namespace {

constexpr int UM = 128, UBK = 32, UMAXBLK = 32, NSTAGE = 3, PD = 1;
constexpr int EPI_WARPS = 8, UEPI = EPI_WARPS * 32, USTAGERS = 192, ISSUE_WARP = (UEPI + USTAGERS) / 32;
constexpr int SG_THREADS = USTAGERS / 2, SG_ROWS = SG_THREADS / 4, SROWS = (UM + SG_ROWS - 1) / SG_ROWS;   // per staging group: threads,
                                                           // rows per pass (4 threads = the 4 chunks of a row), passes
constexpr int COPY_WARP = ISSUE_WARP + 1;
constexpr int UTHREADS = UEPI + USTAGERS + 64;             // 8 epilogue warps | 6 staging warps | the MMA warp | the copy warp:
                                                           // 16 warps, so that a thread may use 128 registers (18 warps: 96)
constexpr int LBO_A = UM * 16 + 16;                        // k-chunk stride of the activation operand (16 B of padding: the
                                                           // 8 chunks a quarter warp stores land in 8 different bank groups)
constexpr int A_HALF_BYTES = (UBK / 8) * LBO_A;            // one of A_hi / A_lo
constexpr int B_OFF = 2 * A_HALF_BYTES;
constexpr int STAGE_BYTES = B_OFF + 2 * 256 * UBK * 2;     // + B_hi, B_lo at N = 256
constexpr int STG_LD = 132;                                // row stride (floats) of the epilogue staging tile [128][128]
constexpr int U_SMEM = NSTAGE * STAGE_BYTES + UM * STG_LD * 4;
static_assert(UBK == 32, "the staging threads are mapped for blocks of 32 (tail 16)");
static_assert(U_SMEM <= 227 * 1024, "one CTA per SM");

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
    int H, R, S, NT, nty, first_layer, legacy, residual_legacy;
    int relu_in, plain, out_ld;  // head 1x1 (wavenet.py:838-842): relu on the operand, epilogue = bias + relu -> out [N][out_ld]
    unsigned long long *stats;   // development aid (WN_UMMA_STATS): clock64 sums per role, [EPI][16]
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

// Development aid: -DWN_UMMA_STATS_BUILD adds clock64 sums of the time every role of umma_layer_kernel waits on its
// mbarriers (printed by wn_umma_free when WN_UMMA_STATS is set); production builds carry neither the counters nor their
// registers.
#ifdef WN_UMMA_STATS_BUILD
#define TWAIT(slot, bar, par) do { const long long t0_ = clock64(); umma::mbar_wait(bar, par); st_acc[slot] += clock64() - t0_; } while (0)
#define STATS_ONLY(...) __VA_ARGS__
#else
#define TWAIT(slot, bar, par) umma::mbar_wait(bar, par)
#define STATS_ONLY(...)
#endif

template <int EPI>
__global__ void __launch_bounds__(UTHREADS, 1) umma_layer_kernel(const __grid_constant__ UArgs a) {
    extern __shared__ __align__(128) unsigned char usmem[];
    // full (weights landed) | ready (activations staged) | done (MMAs read the stage) per operand stage;
    // tfull (accumulator finished) | tempty (epilogue has read it) per TMEM accumulator
    __shared__ __align__(8) unsigned long long bars[3 * NSTAGE + 4];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar0 = umma::smem_addr(&bars[0]);
    auto full = [&](int s) { return bar0 + 8u * (uint32_t)s; };
    auto ready = [&](int s) { return bar0 + 8u * (uint32_t)(NSTAGE + s); };
    auto done = [&](int s) { return bar0 + 8u * (uint32_t)(2 * NSTAGE + s); };
    auto tfull = [&](int i) { return bar0 + 8u * (uint32_t)(3 * NSTAGE + i); };
    auto tempty = [&](int i) { return bar0 + 8u * (uint32_t)(3 * NSTAGE + 2 + i); };
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) { umma::mbar_init(full(s), 1); umma::mbar_init(ready(s), USTAGERS / 2); umma::mbar_init(done(s), 1); }
        for (int i = 0; i < 2; ++i) { umma::mbar_init(tfull(i), 1); umma::mbar_init(tempty(i), UEPI); }
        umma::mbar_fence_init();
    }
    if (warp == ISSUE_WARP) umma::tmem_alloc(umma::smem_addr(&tmem_slot), 512);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_slot;
    const int ncols = a.ncols, nblk = a.nblk, nty = a.nty;
    const long long total = ((a.N + UM - 1) / UM) * nty;     // work items: (position tile, column tile), column tile minor
    const uint32_t smem0 = umma::smem_addr(usmem);
    STATS_ONLY(long long st_acc[4] = {0, 0, 0, 0}; const long long st_begin = clock64();)

    if (warp == COPY_WARP) {
        // ---- copy thread: the packed weights of block g go to stage g % NSTAGE as soon as the MMAs that last read that stage
        //      have completed (up to NSTAGE blocks ahead of the MMA thread) ----------------------------------------------------
        if (lane == 0) {
            uint32_t s = 0, ph = 0, g = 0;                       // stage, parity of its current use, blocks requested
            int ty = (int)(blockIdx.x % nty);
            const int ty_step = (int)(gridDim.x % nty);
            for (long long w = blockIdx.x; w < total; w += gridDim.x) {
                const unsigned char *Wt = a.W + (long long)ty * a.tile_w_bytes;
                for (int b = 0; b < nblk; ++b, ++g) {
                    if (g >= NSTAGE) TWAIT(3, done(s), ph ^ 1u);   // use - 1 of this stage has been read
                    const uint32_t bytes = (uint32_t)a.blk[b].bk * (uint32_t)ncols * 4u;
                    const uint32_t Bs_a = smem0 + s * STAGE_BYTES + B_OFF;
                    const unsigned char *src = Wt + a.blk[b].w_off;
                    umma::mbar_expect_tx(full(s), bytes);
                    for (uint32_t o = 0; o < bytes; o += 16384u)
                        umma::bulk_g2s(Bs_a + o, src + o, bytes - o < 16384u ? bytes - o : 16384u, full(s));
                    if (++s == NSTAGE) { s = 0; ph ^= 1u; }
                }
                ty += ty_step;
                if (ty >= nty) ty -= nty;
            }
            STATS_ONLY(if (a.stats) atomicAdd(a.stats + EPI * 16 + 3, (unsigned long long)st_acc[3]);)
        }
        __syncwarp();
    } else if (warp == ISSUE_WARP) {
        // ---- MMA thread: tcgen05.mma into accumulator it & 1, commits ------------------------------------------------------
        if (lane == 0) {
            const uint32_t idesc = umma::idesc_bf16(UM, ncols);
            const uint32_t lboB = (uint32_t)ncols * 16;
            const uint64_t dA = umma::smem_desc(0, LBO_A, 128), dB = umma::smem_desc(0, lboB, 128);   // + (address >> 4)
            const uint32_t kA = (2 * LBO_A) >> 4, kB = (2 * lboB) >> 4;                             // one K step of 16
            uint32_t s = 0, ph = 0;
            int it = 0;
            for (long long w = blockIdx.x; w < total; w += gridDim.x, ++it) {
                const int ab = it & 1;
                if (it >= 2) { TWAIT(0, tempty(ab), (uint32_t)((it >> 1) - 1) & 1u); umma::fence_after_sync(); }
                const uint32_t d_tmem = tmem + (uint32_t)ab * 256u;
                for (int b = 0; b < nblk; ++b) {
                    const int bk = a.blk[b].bk;
                    TWAIT(1, ready(s), ph);
                    TWAIT(2, full(s), ph);
                    umma::fence_after_sync();
                    const uint32_t st16 = (smem0 + s * STAGE_BYTES) >> 4;
                    const uint64_t ah = dA + st16, al = ah + (A_HALF_BYTES >> 4), bh = dB + st16 + (B_OFF >> 4),
                                   bl = bh + (((uint32_t)bk * (uint32_t)ncols * 2u) >> 4);
                    umma::mma_bf16(d_tmem, ah, bh, idesc, b > 0 ? 1u : 0u);
                    umma::mma_bf16(d_tmem, ah, bl, idesc, 1u);
                    umma::mma_bf16(d_tmem, al, bh, idesc, 1u);
                    if (bk > 16) {
                        umma::mma_bf16(d_tmem, ah + kA, bh + kB, idesc, 1u);
                        umma::mma_bf16(d_tmem, ah + kA, bl + kB, idesc, 1u);
                        umma::mma_bf16(d_tmem, al + kA, bh + kB, idesc, 1u);
                    }
                    umma::mma_commit(done(s));
                    if (++s == NSTAGE) { s = 0; ph ^= 1u; }
                }
                umma::mma_commit(tfull(ab));                      // every MMA of the tile has completed
            }
            STATS_ONLY(if (a.stats) { for (int i = 0; i < 3; ++i) atomicAdd(a.stats + EPI * 16 + i, (unsigned long long)st_acc[i]); atomicAdd(a.stats + EPI * 16 + 4, (unsigned long long)(clock64() - st_begin)); atomicAdd(a.stats + EPI * 16 + 5, 1ull); })
        }
        __syncwarp();
    } else if (warp >= EPI_WARPS) {
        // ---- staging warps: gather, split into bf16 (hi, lo), store in the K-major core-matrix layout -----------------------
        // Two groups of 4 warps take alternate blocks (a block is one dependent chain load -> split -> store -> fence -> arrive
        // per warp, so two blocks in flight nearly double the staging rate).  thread = (8-float chunk c of a 32-wide block,
        // rows r0 + 32 j): the lanes of a warp read whole 128-byte row segments.  What depends only on the tile (row pointers,
        // position inside the utterance) is computed once per tile; a 16-wide tail block keeps the mapping with the chunks
        // c >= 2 idle.
        const int st_tid = tid - EPI_WARPS * 32, grp = st_tid / SG_THREADS, c = st_tid & 3, r0 = (st_tid % SG_THREADS) >> 2;
        const int soff = c * LBO_A + r0 * 16;
        uint32_t gb0 = 0;                                        // global index of the tile's first block
        for (long long w = blockIdx.x; w < total; w += gridDim.x, gb0 += (uint32_t)nblk) {
            const long long n0 = (w / nty) * UM;
            const float *srow[SROWS], *crow[SROWS];
            int trow[SROWS];
#pragma unroll
            for (int j = 0; j < SROWS; ++j) {
                const long long n = n0 + r0 + SG_ROWS * j;
                const bool ok = n < a.N && r0 + SG_ROWS * j < UM;
                trow[j] = ok ? (int)(n % a.T) : -1;               // -1: no such position (tail tile) -> zeros
                srow[j] = a.src + (ok ? n : 0) * a.src_ld + 8 * c;
                crow[j] = a.cond ? a.cond + (ok ? n : 0) * a.cond_ld + 8 * c : nullptr;
            }
            // the activations of a block are requested PD of the group's blocks before they are split and stored
            float4 pf[PD][2 * SROWS];
            auto load_block = [&](int b, float4 (&pfx)[2 * SROWS]) {
                const UBlock &k = a.blk[b];
                const bool c_ok = 8 * c < k.bk;
                const int back_off = k.koff - k.back * a.src_ld;
#pragma unroll
                for (int j = 0; j < SROWS; ++j) {
                    const float *p = k.seg == 0 ? srow[j] + back_off : crow[j] + k.koff;
                    float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                    if (c_ok && trow[j] >= k.back) { v0 = __ldg(reinterpret_cast<const float4 *>(p)); v1 = __ldg(reinterpret_cast<const float4 *>(p + 4)); }
                    if (a.relu_in) {
                        v0.x = fmaxf(v0.x, 0.f); v0.y = fmaxf(v0.y, 0.f); v0.z = fmaxf(v0.z, 0.f); v0.w = fmaxf(v0.w, 0.f);
                        v1.x = fmaxf(v1.x, 0.f); v1.y = fmaxf(v1.y, 0.f); v1.z = fmaxf(v1.z, 0.f); v1.w = fmaxf(v1.w, 0.f);
                    }
                    pfx[2 * j] = v0; pfx[2 * j + 1] = v1;
                }
            };
            auto step = [&](int b, float4 (&pfx)[2 * SROWS]) {
                const uint32_t g = gb0 + (uint32_t)b, s = g % NSTAGE, use = g / NSTAGE;
                if (g >= NSTAGE) TWAIT(0, done(s), (use - 1u) & 1u);
                if (8 * c < a.blk[b].bk) {
                    unsigned char *Ah = usmem + (size_t)s * STAGE_BYTES + soff, *Al = Ah + A_HALF_BYTES;
#pragma unroll
                    for (int j = 0; j < SROWS; ++j) {
                        if (r0 + SG_ROWS * j >= UM) continue;
                        uint4 hi, lo;
                        umma::split8(pfx[2 * j], pfx[2 * j + 1], hi, lo);
                        *reinterpret_cast<uint4 *>(Ah + j * SG_ROWS * 16) = hi;
                        *reinterpret_cast<uint4 *>(Al + j * SG_ROWS * 16) = lo;
                    }
                }
                umma::fence_proxy_async();
                umma::mbar_arrive(ready(s));
                if (b + 2 * PD < nblk) load_block(b + 2 * PD, pfx);
            };
            const int bfirst = (int)((gb0 ^ (uint32_t)grp) & 1u);   // the group's first block of this tile
#pragma unroll
            for (int i = 0; i < PD; ++i)
                if (bfirst + 2 * i < nblk) load_block(bfirst + 2 * i, pf[i]);
            for (int b = bfirst; b < nblk; b += 2 * PD) {
#pragma unroll
                for (int i = 0; i < PD; ++i)
                    if (b + 2 * i < nblk) step(b + 2 * i, pf[i]);
            }
        }
    } else {
        // ---- epilogue warps: TMEM -> registers (thread = position 32q + lane, half of the columns) -> staging tile in shared
        //      memory -> coalesced rows (one warp per position row), while the other warps work on the next tile ----------------
        float *stg = reinterpret_cast<float *>(usmem + NSTAGE * STAGE_BYTES);
        const int q = warp & 3, half = warp >> 2;
        float *srow = stg + (size_t)(32 * q + lane) * STG_LD;
        auto epi_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(UEPI) : "memory"); };
        int it = 0;
        for (long long w = blockIdx.x; w < total; w += gridDim.x, ++it) {
            const long long n0 = (w / nty) * UM;
            const int ty = (int)(w % nty), ab = it & 1;
            const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)ab * 256u;
            const float *bias = a.bias + (size_t)ty * ncols;
            TWAIT(0, tfull(ab), (uint32_t)(it >> 1) & 1u);
            umma::fence_after_sync();
            if (EPI == UEPI_GATE) {
                const int NT = a.NT;
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
                umma::fence_before_sync();
                umma::mbar_arrive(tempty(ab));                    // the accumulator may be overwritten
                epi_sync();
                if (4 * lane < NT)
                    for (int r = warp; r < UM; r += EPI_WARPS) {
                        const long long n = n0 + r;
                        if (n < a.N)
                            *reinterpret_cast<float4 *>(a.out + n * a.H + ty * NT + 4 * lane) = *reinterpret_cast<const float4 *>(stg + (size_t)r * STG_LD + 4 * lane);
                    }
                epi_sync();
            } else {
                const int ngroups = ncols / 128;
                for (int g = 0; g < ngroups; ++g) {
                    for (int c = half * 64; c < half * 64 + 64; c += 16) {
                        float v[16];
                        umma::tmem_ld16(tq + (uint32_t)(g * 128 + c), v);
                        umma::tmem_ld_wait();
#pragma unroll
                        for (int i4 = 0; i4 < 4; ++i4)
                            *reinterpret_cast<float4 *>(srow + c + 4 * i4) = make_float4(v[4 * i4], v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                    }
                    if (g == ngroups - 1) { umma::fence_before_sync(); umma::mbar_arrive(tempty(ab)); }
                    epi_sync();
                    const int m = ty * ncols + g * 128 + 4 * lane;            // first of this lane's 4 output rows of the GEMM
                    const float4 bb = __ldg(reinterpret_cast<const float4 *>(bias + g * 128 + 4 * lane));
                    // what is added to a row: the layer input (residual, modules.py:515-520) or the running skip sum
                    // (wavenet.py:833-836); 8 rows of loads are in flight before the first one is used
                    const bool plain = a.plain != 0, is_out = plain || m < a.R, has_add = !plain && (is_out || !a.first_layer);
                    const float *addp = is_out ? a.xin + m : a.skips + (m - a.R);
                    float *dstp = is_out ? a.out + m : a.skips + (m - a.R);
                    const int ld = plain ? a.out_ld : (is_out ? a.R : a.S);
                    const bool scale = !plain && (is_out ? a.residual_legacy != 0 : (a.legacy != 0 && !a.first_layer));
                    for (int rb = warp; rb < UM; rb += 8 * EPI_WARPS) {
                        float4 add[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const long long n = n0 + rb + EPI_WARPS * k;
                            add[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (has_add && n < a.N) add[k] = *reinterpret_cast<const float4 *>(addp + n * ld);
                        }
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const int r = rb + EPI_WARPS * k;
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
                            if (plain) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
                            *reinterpret_cast<float4 *>(dstp + n * ld) = y;
                        }
                    }
                    epi_sync();
                }
            }
        }
    }
#ifdef WN_UMMA_STATS_BUILD
    if (a.stats && (tid == 0 || tid == UEPI)) {   // (the stagers' wait is that of group 0)
        const int o = EPI * 16 + (tid == 0 ? 6 : 8);
        atomicAdd(a.stats + o, (unsigned long long)st_acc[0]); atomicAdd(a.stats + o + 1, (unsigned long long)(clock64() - st_begin));
    }
#endif
    umma::fence_before_sync();
    __syncthreads();
    if (warp == ISSUE_WARP) umma::tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------------------------------------------------
// Fused residual layer: gate GEMM -> tanh * sigmoid -> out/skip GEMM in ONE persistent kernel; z never leaves the SM.
// For H = 128 and R + S = 256 (one column tile each: the default block).  TMEM: gate accumulator in columns [0, 256),
// out/skip accumulator in [256, 512).  Per tile of 128 positions:
//   G   the 15 K blocks of the gate as in umma_layer_kernel (copy thread, staging warps, MMA thread)       -> accG
//   E1  epilogue warps: accG -> gate -> z split into bf16 (hi, lo) and stored as the A operand of the second GEMM
//       (K-major core-matrix layout, 16 chunks of 8 channels); every 32 channels finished are announced on their own
//       mbarrier, so the MMA thread starts the out/skip K block b while E1 still works on later channels
//   O   4 K blocks: A = z from shared memory, B = [W_out ; W_skip] through the same operand stages           -> accO
//   E2  epilogue warps: accO -> staging tile (the z buffer, dead by then) -> residual / skip rows, coalesced
// G of the next tile starts as soon as E1 has drained accG and runs under E2.  HBM traffic per layer: x (+ its two
// older taps, mostly L2 hits), c, the skip sum in and out, x' out - the [N][H] round trip of z is gone (-36 %).
struct FArgs {
    const unsigned char *Wg, *Wo;     // packed weights of the layer: gate blocks, out/skip blocks
    const float *bias_g, *bias_o;
    const float *x, *cond;            // layer input [N][R], conditioning [N][cin]
    float *x_out, *skips;
    long long N;
    int T, R, S, cin, nblk_g, nblk_o, first_layer, legacy, residual_legacy;
    UBlock blk[UMAXBLK];              // gate blocks, then out/skip blocks
    unsigned long long *stats;
};

constexpr int F_NBAR = 3 * NSTAGE + 2 + 4 + 1;   // full | ready | done per stage, tgfull, tgempty, zready[4], tofull
constexpr int ERB = 8;                           // rows per warp whose residual / skip loads are in flight together
constexpr int Z_HALF_BYTES = 16 * LBO_A;         // z_hi or z_lo: 16 chunks of 8 channels x 128 positions
static_assert(2 * Z_HALF_BYTES <= UM * STG_LD * 4, "the z operand and the epilogue staging tile share one region");

__global__ void __launch_bounds__(UTHREADS, 1) umma_fused_layer_kernel(const __grid_constant__ FArgs a) {
    extern __shared__ __align__(128) unsigned char usmem[];
    __shared__ __align__(8) unsigned long long bars[F_NBAR];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar0 = umma::smem_addr(&bars[0]);
    auto full = [&](int s) { return bar0 + 8u * (uint32_t)s; };
    auto ready = [&](int s) { return bar0 + 8u * (uint32_t)(NSTAGE + s); };
    auto done = [&](int s) { return bar0 + 8u * (uint32_t)(2 * NSTAGE + s); };
    const uint32_t tgfull = bar0 + 8u * (3 * NSTAGE), tgempty = tgfull + 8u, zready0 = tgfull + 16u, tofull = tgfull + 48u;
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) { umma::mbar_init(full(s), 1); umma::mbar_init(ready(s), USTAGERS / 2); umma::mbar_init(done(s), 1); }
        umma::mbar_init(tgfull, 1); umma::mbar_init(tgempty, UEPI); umma::mbar_init(tofull, 1);
        for (int i = 0; i < 4; ++i) umma::mbar_init(zready0 + 8u * i, UEPI / 2);
        umma::mbar_fence_init();
    }
    if (warp == ISSUE_WARP) umma::tmem_alloc(umma::smem_addr(&tmem_slot), 512);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_slot;
    const int nbg = a.nblk_g, nbo = a.nblk_o, nb = nbg + nbo;
    const long long total = (a.N + UM - 1) / UM;
    const uint32_t smem0 = umma::smem_addr(usmem);
    unsigned char *zbuf = usmem + NSTAGE * STAGE_BYTES;          // z_hi | z_lo, later the staging tile of E2
    const uint32_t z_a = smem0 + NSTAGE * STAGE_BYTES;

    if (warp == COPY_WARP) {
        if (lane == 0) {
            uint32_t s = 0, ph = 0, g = 0;
            for (long long w = blockIdx.x; w < total; w += gridDim.x) {
                for (int b = 0; b < nb; ++b, ++g) {
                    if (g >= NSTAGE) umma::mbar_wait(done(s), ph ^ 1u);
                    const uint32_t bytes = (uint32_t)a.blk[b].bk * 256u * 4u;
                    const uint32_t Bs_a = smem0 + s * STAGE_BYTES + B_OFF;
                    const unsigned char *src = (b < nbg ? a.Wg : a.Wo) + a.blk[b].w_off;
                    umma::mbar_expect_tx(full(s), bytes);
                    for (uint32_t o = 0; o < bytes; o += 16384u)
                        umma::bulk_g2s(Bs_a + o, src + o, bytes - o < 16384u ? bytes - o : 16384u, full(s));
                    if (++s == NSTAGE) { s = 0; ph ^= 1u; }
                }
            }
        }
        __syncwarp();
    } else if (warp == ISSUE_WARP) {
        if (lane == 0) {
            const uint32_t idesc = umma::idesc_bf16(UM, 256);
            const uint32_t lboB = 256u * 16u;
            const uint64_t dA = umma::smem_desc(0, LBO_A, 128), dB = umma::smem_desc(0, lboB, 128);
            const uint32_t kA = (2 * LBO_A) >> 4, kB = (2 * lboB) >> 4;
            const uint64_t zh = dA + (z_a >> 4), zl = zh + (Z_HALF_BYTES >> 4);
            uint32_t s = 0, ph = 0;
            int it = 0;
            for (long long w = blockIdx.x; w < total; w += gridDim.x, ++it) {
                if (it >= 1) { umma::mbar_wait(tgempty, (uint32_t)(it - 1) & 1u); umma::fence_after_sync(); }
                for (int b = 0; b < nb; ++b) {
                    const int bk = a.blk[b].bk;
                    const bool gate = b < nbg;
                    if (!gate) { umma::mbar_wait(zready0 + 8u * (uint32_t)(b - nbg), (uint32_t)it & 1u); }
                    umma::mbar_wait(ready(s), ph);
                    umma::mbar_wait(full(s), ph);
                    umma::fence_after_sync();
                    const uint32_t st16 = (smem0 + s * STAGE_BYTES) >> 4;
                    const uint64_t bh = dB + st16 + (B_OFF >> 4), bl = bh + (((uint32_t)bk * 256u * 2u) >> 4);
                    uint64_t ah, al;
                    if (gate) { ah = dA + st16; al = ah + (A_HALF_BYTES >> 4); }
                    else { const uint32_t o = (uint32_t)(b - nbg) * 4u * (LBO_A >> 4); ah = zh + o; al = zl + o; }
                    const uint32_t d_tmem = tmem + (gate ? 0u : 256u);
                    const uint32_t acc0 = (gate ? b > 0 : b > nbg) ? 1u : 0u;
                    umma::mma_bf16(d_tmem, ah, bh, idesc, acc0);
                    umma::mma_bf16(d_tmem, ah, bl, idesc, 1u);
                    umma::mma_bf16(d_tmem, al, bh, idesc, 1u);
                    if (bk > 16) {
                        umma::mma_bf16(d_tmem, ah + kA, bh + kB, idesc, 1u);
                        umma::mma_bf16(d_tmem, ah + kA, bl + kB, idesc, 1u);
                        umma::mma_bf16(d_tmem, al + kA, bh + kB, idesc, 1u);
                    }
                    umma::mma_commit(done(s));
                    if (b == nbg - 1) umma::mma_commit(tgfull);
                    if (++s == NSTAGE) { s = 0; ph ^= 1u; }
                }
                umma::mma_commit(tofull);
            }
        }
        __syncwarp();
    } else if (warp >= EPI_WARPS) {
        // ---- staging warps (two groups on alternate blocks); the out/skip blocks have nothing to stage, their turn only
        //      keeps the stage bookkeeping uniform -------------------------------------------------------------------------
        const int st_tid = tid - EPI_WARPS * 32, grp = st_tid / SG_THREADS, c = st_tid & 3, r0 = (st_tid % SG_THREADS) >> 2;
        const int soff = c * LBO_A + r0 * 16;
        uint32_t gb0 = 0;
        for (long long w = blockIdx.x; w < total; w += gridDim.x, gb0 += (uint32_t)nb) {
            const long long n0 = w * UM;
            const float *srow[SROWS], *crow[SROWS];
            int trow[SROWS];
#pragma unroll
            for (int j = 0; j < SROWS; ++j) {
                const long long n = n0 + r0 + SG_ROWS * j;
                const bool ok = n < a.N && r0 + SG_ROWS * j < UM;
                trow[j] = ok ? (int)(n % a.T) : -1;
                srow[j] = a.x + (ok ? n : 0) * a.R + 8 * c;
                crow[j] = a.cond + (ok ? n : 0) * a.cin + 8 * c;
            }
            float4 pf[2 * SROWS];
            auto load_block = [&](int b) {
                if (b >= nbg) return;
                const UBlock &k = a.blk[b];
                const bool c_ok = 8 * c < k.bk;
                const int back_off = k.koff - k.back * a.R;
#pragma unroll
                for (int j = 0; j < SROWS; ++j) {
                    const float *p = k.seg == 0 ? srow[j] + back_off : crow[j] + k.koff;
                    float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                    if (c_ok && trow[j] >= k.back) { v0 = __ldg(reinterpret_cast<const float4 *>(p)); v1 = __ldg(reinterpret_cast<const float4 *>(p + 4)); }
                    pf[2 * j] = v0; pf[2 * j + 1] = v1;
                }
            };
            const int bfirst = (int)((gb0 ^ (uint32_t)grp) & 1u);
            load_block(bfirst);
            for (int b = bfirst; b < nb; b += 2) {
                const uint32_t g = gb0 + (uint32_t)b, s = g % NSTAGE, use = g / NSTAGE;
                if (g >= NSTAGE) umma::mbar_wait(done(s), (use - 1u) & 1u);
                if (b < nbg && 8 * c < a.blk[b].bk) {
                    unsigned char *Ah = usmem + (size_t)s * STAGE_BYTES + soff, *Al = Ah + A_HALF_BYTES;
#pragma unroll
                    for (int j = 0; j < SROWS; ++j) {
                        if (r0 + SG_ROWS * j >= UM) continue;
                        uint4 hi, lo;
                        umma::split8(pf[2 * j], pf[2 * j + 1], hi, lo);
                        *reinterpret_cast<uint4 *>(Ah + j * SG_ROWS * 16) = hi;
                        *reinterpret_cast<uint4 *>(Al + j * SG_ROWS * 16) = lo;
                    }
                }
                umma::fence_proxy_async();
                umma::mbar_arrive(ready(s));
                if (b + 2 < nb) load_block(b + 2);
            }
        }
    } else {
        // ---- epilogue warps -----------------------------------------------------------------------------------------------
        float *stg = reinterpret_cast<float *>(zbuf);
        const int q = warp & 3, half = warp >> 2, row = 32 * q + lane;
        float *srow = stg + (size_t)row * STG_LD;
        const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
        auto epi_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(UEPI) : "memory"); };
        int it = 0;
        for (long long w = blockIdx.x; w < total; w += gridDim.x, ++it) {
            const long long n0 = w * UM;
            // E1: gate -> z (bf16 hi / lo) as the A operand of the second GEMM
            umma::mbar_wait(tgfull, (uint32_t)it & 1u);
            umma::fence_after_sync();
            for (int c = half * 64; c < half * 64 + 64; c += 16) {
                float ta[16], sg[16];
                umma::tmem_ld16(tq + (uint32_t)c, ta);
                umma::tmem_ld16(tq + (uint32_t)(128 + c), sg);
                umma::tmem_ld_wait();
                float zz[16];
#pragma unroll
                for (int i4 = 0; i4 < 4; ++i4) {
                    const float4 bt = __ldg(reinterpret_cast<const float4 *>(a.bias_g + c + 4 * i4));
                    const float4 bs = __ldg(reinterpret_cast<const float4 *>(a.bias_g + 128 + c + 4 * i4));
                    zz[4 * i4] = gate_fast(ta[4 * i4] + bt.x, sg[4 * i4] + bs.x);
                    zz[4 * i4 + 1] = gate_fast(ta[4 * i4 + 1] + bt.y, sg[4 * i4 + 1] + bs.y);
                    zz[4 * i4 + 2] = gate_fast(ta[4 * i4 + 2] + bt.z, sg[4 * i4 + 2] + bs.z);
                    zz[4 * i4 + 3] = gate_fast(ta[4 * i4 + 3] + bt.w, sg[4 * i4 + 3] + bs.w);
                }
#pragma unroll
                for (int h8 = 0; h8 < 2; ++h8) {
                    uint4 hi, lo;
                    umma::split8(make_float4(zz[8 * h8], zz[8 * h8 + 1], zz[8 * h8 + 2], zz[8 * h8 + 3]),
                                 make_float4(zz[8 * h8 + 4], zz[8 * h8 + 5], zz[8 * h8 + 6], zz[8 * h8 + 7]), hi, lo);
                    const int off = ((c >> 3) + h8) * LBO_A + row * 16;
                    *reinterpret_cast<uint4 *>(zbuf + off) = hi;
                    *reinterpret_cast<uint4 *>(zbuf + Z_HALF_BYTES + off) = lo;
                }
                if (c & 16) {   // 32 channels of all this half's positions are in place: out/skip K block c / 32 may start
                    umma::fence_proxy_async();
                    umma::mbar_arrive(zready0 + 8u * (uint32_t)(c >> 5));
                }
            }
            umma::fence_before_sync();
            umma::mbar_arrive(tgempty);                       // accG may be overwritten by the next tile's gate
            // E2: out / skip
            umma::mbar_wait(tofull, (uint32_t)it & 1u);
            umma::fence_after_sync();
            for (int g = 0; g < 2; ++g) {
                for (int c = half * 64; c < half * 64 + 64; c += 16) {
                    float v[16];
                    umma::tmem_ld16(tq + (uint32_t)(256 + g * 128 + c), v);
                    umma::tmem_ld_wait();
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4)
                        *reinterpret_cast<float4 *>(srow + c + 4 * i4) = make_float4(v[4 * i4], v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                }
                epi_sync();
                const int m = g * 128 + 4 * lane;                         // first of this lane's 4 rows of [out ; skip]
                const float4 bb = __ldg(reinterpret_cast<const float4 *>(a.bias_o + m));
                // what is added to a row: the layer input (residual, modules.py:515-520) or the running skip sum
                // (wavenet.py:833-836); 8 rows of loads are in flight before the first one is used (16 in flight, with or without setmaxnreg moving registers from the staging to the epilogue warps, measured 25-30 % slower)
                const bool is_out = m < a.R, has_add = is_out || !a.first_layer;
                const float *addp = is_out ? a.x + m : a.skips + (m - a.R);
                float *dstp = is_out ? a.x_out + m : a.skips + (m - a.R);
                const int ld = is_out ? a.R : a.S;
                const bool scale = is_out ? a.residual_legacy != 0 : (a.legacy != 0 && !a.first_layer);
                for (int rb = warp; rb < UM; rb += ERB * EPI_WARPS) {
                    float4 add[ERB];
#pragma unroll
                    for (int k = 0; k < ERB; ++k) {
                        const long long n = n0 + rb + EPI_WARPS * k;
                        add[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (has_add && n < a.N) add[k] = *reinterpret_cast<const float4 *>(addp + n * ld);
                    }
#pragma unroll
                    for (int k = 0; k < ERB; ++k) {
                        const int r = rb + EPI_WARPS * k;
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
                epi_sync();
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == ISSUE_WARP) umma::tmem_dealloc(tmem, 512);
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
    unsigned long long *d_stats = nullptr;
    std::vector<size_t> gate_w, os_w, gate_b, os_b;   // per layer: byte offsets into d_w, float offsets into d_bias
    // head: final_convolution_1 on the tensor cores when S is a multiple of 128; final_convolution_2 by head2_kernel
    bool head = false;
    int h1_ncols = 0, h1_tiles = 0;
    unsigned h1_tile_bytes = 0;
    std::vector<KBlockHost> h1_blk;
    size_t h1_w = 0, h1_b = 0, h2_w = 0, h2_b = 0;    // h2_w: [S][O] floats inside d_bias
};

void wn_umma_free(wn_handle *h) {
    WnUmmaPlan *u = h->umma;
    if (!u) return;
    if (u->d_stats) {   // development aid: average clocks per CTA and launch
        unsigned long long st[32];
        cudaDeviceSynchronize();
        if (cudaMemcpy(st, u->d_stats, sizeof(st), cudaMemcpyDeviceToHost) == cudaSuccess)
            for (int e = 0; e < 2; ++e) {
                const double n = st[e * 16 + 5] ? (double)st[e * 16 + 5] : 1.0;
                fprintf(stderr, "umma stats %s: issuer waits tempty %.0f ready %.0f full %.0f done %.0f of %.0f clk | epilogue wait tfull %.0f of %.0f | stager wait done %.0f of %.0f (per CTA launch, %g launches)\n",
                        e ? "out/skip" : "gate", st[e * 16] / n, st[e * 16 + 1] / n, st[e * 16 + 2] / n, st[e * 16 + 3] / n, st[e * 16 + 4] / n,
                        st[e * 16 + 6] / n, st[e * 16 + 7] / n, st[e * 16 + 8] / n, st[e * 16 + 9] / n, n);
            }
        cudaFree(u->d_stats);
    }
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
    if (S % 128 == 0 && m.O <= 32) {
        const float *w1 = Wt("final_convolution_1/kernel"), *b1 = Wt("final_convolution_1/bias");
        const float *w2 = Wt("final_convolution_2/kernel"), *b2 = Wt("final_convolution_2/bias");
        u->head = true;
        u->h1_ncols = S % 256 == 0 ? 256 : 128; u->h1_tiles = S / u->h1_ncols;
        u->h1_blk = cut_blocks({{0, 2, S}}, u->h1_ncols, &u->h1_tile_bytes);
        u->h1_w = pack(u->h1_blk, u->h1_ncols, u->h1_tiles, u->h1_tile_bytes, [&](int k, int row) { return w1[(size_t)k * S + row]; });
        u->h1_b = bias.size();
        for (int n = 0; n < S; ++n) bias.push_back(b1 ? b1[n] : 0.0f);
        u->h2_w = bias.size();
        for (int k = 0; k < S; ++k)
            for (int o = 0; o < m.O; ++o) bias.push_back(w2[(size_t)k * m.O + o]);
        u->h2_b = bias.size();
        for (int o = 0; o < m.O; ++o) bias.push_back(b2 ? b2[o] : 0.0f);
    }
    if (cudaMalloc((void **)&u->d_w, w.size() * 2) != cudaSuccess ||
        cudaMemcpy(u->d_w, w.data(), w.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMalloc((void **)&u->d_bias, bias.size() * sizeof(float)) != cudaSuccess ||
        cudaMemcpy(u->d_bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaFuncSetAttribute(umma_layer_kernel<UEPI_GATE>, cudaFuncAttributeMaxDynamicSharedMemorySize, U_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(umma_layer_kernel<UEPI_OUTSKIP>, cudaFuncAttributeMaxDynamicSharedMemorySize, U_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(umma_fused_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, U_SMEM) != cudaSuccess) {
        h->err = std::string("step tensor-core weights: ") + cudaGetErrorString(cudaGetLastError());
        if (u->d_w) cudaFree(u->d_w);
        if (u->d_bias) cudaFree(u->d_bias);
        delete u;
        return WN_ERR_NOMEM;
    }
    if (getenv("WN_UMMA_STATS") && cudaMalloc((void **)&u->d_stats, 32 * sizeof(unsigned long long)) == cudaSuccess)
        cudaMemset(u->d_stats, 0, 32 * sizeof(unsigned long long));
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
    a.stats = u->d_stats;
    a.N = N; a.T = T; a.H = m.H; a.R = m.R; a.S = m.S; a.NT = u->NT;
    a.legacy = m.legacy; a.residual_legacy = m.residual_legacy; a.first_layer = l == 0;
    const long long ptiles = (N + UM - 1) / UM;
    // one fused kernel when each GEMM is a single column tile of 256 (the default block); WN_STEP_UNFUSED=1: two kernels
    const bool unfused = getenv("WN_STEP_UNFUSED") && atoi(getenv("WN_STEP_UNFUSED"));
    if (!unfused && m.H == 128 && m.R + m.S == 256 && u->gate_tiles == 1 && u->os_tiles == 1 &&
        (int)(u->gate_blk.size() + u->os_blk.size()) <= UMAXBLK) {
        FArgs f;
        memset(&f, 0, sizeof(f));
        f.Wg = u->d_w + u->gate_w[l]; f.Wo = u->d_w + u->os_w[l];
        f.bias_g = u->d_bias + u->gate_b[l]; f.bias_o = u->d_bias + u->os_b[l];
        f.x = x; f.cond = c_up; f.x_out = x_out; f.skips = skips;
        f.N = N; f.T = T; f.R = m.R; f.S = m.S; f.cin = m.cin;
        f.nblk_g = (int)u->gate_blk.size(); f.nblk_o = (int)u->os_blk.size();
        f.first_layer = l == 0; f.legacy = m.legacy; f.residual_legacy = m.residual_legacy;
        for (int i = 0; i < f.nblk_g; ++i) {
            const KBlockHost &b = u->gate_blk[i];
            f.blk[i] = {b.seg, b.koff, b.bk, b.seg == 0 ? (2 - b.tap) * m.dil[l] : 0, b.w_off};
        }
        for (int i = 0; i < f.nblk_o; ++i) {
            const KBlockHost &b = u->os_blk[i];
            f.blk[f.nblk_g + i] = {0, b.koff, b.bk, 0, b.w_off};
        }
        umma_fused_layer_kernel<<<(unsigned)std::min<long long>(ptiles, h->sm_count), UTHREADS, U_SMEM, st>>>(f);
        WN_CUDA_TRY(h, cudaGetLastError());
        h->launches += 1;
        return WN_OK;
    }
    auto grid_for = [&](int nty) { return (unsigned)std::min<long long>(ptiles * nty, h->sm_count); };   // persistent: one CTA per SM
    // gate
    a.W = u->d_w + u->gate_w[l]; a.tile_w_bytes = u->gate_tile_bytes; a.bias = u->d_bias + u->gate_b[l];
    a.src = x; a.src_ld = m.R; a.cond = c_up; a.cond_ld = m.cin; a.ncols = u->gate_ncols;
    a.nblk = (int)u->gate_blk.size();
    for (int i = 0; i < a.nblk; ++i) {
        const KBlockHost &b = u->gate_blk[i];
        a.blk[i] = {b.seg, b.koff, b.bk, b.seg == 0 ? (2 - b.tap) * m.dil[l] : 0, b.w_off};
    }
    a.out = z;
    a.nty = u->gate_tiles;
    umma_layer_kernel<UEPI_GATE><<<grid_for(a.nty), UTHREADS, U_SMEM, st>>>(a);
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
    a.nty = u->os_tiles;
    umma_layer_kernel<UEPI_OUTSKIP><<<grid_for(a.nty), UTHREADS, U_SMEM, st>>>(a);
    WN_CUDA_TRY(h, cudaGetLastError());
    h->launches += 2;
    return WN_OK;
}

namespace {
// final_convolution_2 (wavenet.py:842-844): raw[n][o] = b2[o] + sum_k h1[n][k] * W2[k][o], one warp per position (the row is
// read once, coalesced); K ascending inside a lane, lanes combined by a butterfly
__global__ void __launch_bounds__(256) head2_kernel(const float *__restrict__ h1, const float *__restrict__ w2, const float *__restrict__ b2,
                                                    long long N, int S, int O, float *__restrict__ raw) {
    const int lane = threadIdx.x & 31;
    const long long n = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (n >= N) return;
    for (int o0 = 0; o0 < O; o0 += 2) {
        float acc0 = 0.f, acc1 = 0.f;
        for (int k = 4 * lane; k < S; k += 128) {
            const float4 v = *reinterpret_cast<const float4 *>(h1 + n * S + k);
            const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                acc0 = __fmaf_rn(vv[e], __ldg(w2 + (size_t)(k + e) * O + o0), acc0);
                if (o0 + 1 < O) acc1 = __fmaf_rn(vv[e], __ldg(w2 + (size_t)(k + e) * O + o0 + 1), acc1);
            }
        }
#pragma unroll
        for (int msk = 16; msk >= 1; msk >>= 1) {
            acc0 = __fadd_rn(acc0, __shfl_xor_sync(0xffffffffu, acc0, msk));
            acc1 = __fadd_rn(acc1, __shfl_xor_sync(0xffffffffu, acc1, msk));
        }
        if (lane == 0) {
            raw[n * O + o0] = __fadd_rn(acc0, b2[o0]);
            if (o0 + 1 < O) raw[n * O + o0 + 1] = __fadd_rn(acc1, b2[o0 + 1]);
        }
    }
}
}  // namespace

bool wn_umma_head_supported(wn_handle *h) { return umma_plan(h) == WN_OK && h->umma->head; }

// relu -> final_convolution_1 -> relu (tensor cores) -> final_convolution_2: skips [N][S] -> raw [N][O]; h1 is scratch [N][S]
int wn_umma_head(wn_handle *h, const float *skips, float *h1, float *raw, long long N, int T, cudaStream_t st) {
    int rc = umma_plan(h);
    if (rc) return rc;
    const WnUmmaPlan *u = h->umma;
    const WnModelDev &m = h->m;
    UArgs a;
    memset(&a, 0, sizeof(a));
    a.stats = nullptr;
    a.N = N; a.T = T; a.H = m.H; a.R = m.R; a.S = m.S; a.NT = u->NT;
    a.W = u->d_w + u->h1_w; a.tile_w_bytes = u->h1_tile_bytes; a.bias = u->d_bias + u->h1_b;
    a.src = skips; a.src_ld = m.S; a.ncols = u->h1_ncols; a.nty = u->h1_tiles;
    a.nblk = (int)u->h1_blk.size();
    for (int i = 0; i < a.nblk; ++i) a.blk[i] = {0, u->h1_blk[i].koff, u->h1_blk[i].bk, 0, u->h1_blk[i].w_off};
    a.relu_in = 1; a.plain = 1; a.out = h1; a.out_ld = m.S;
    const long long ptiles = (N + UM - 1) / UM;
    umma_layer_kernel<UEPI_OUTSKIP><<<(unsigned)std::min<long long>(ptiles * a.nty, h->sm_count), UTHREADS, U_SMEM, st>>>(a);
    WN_CUDA_TRY(h, cudaGetLastError());
    head2_kernel<<<(unsigned)((N + 7) / 8), 256, 0, st>>>(h1, u->d_bias + u->h2_w, u->d_bias + u->h2_b, N, m.S, m.O, raw);
    WN_CUDA_TRY(h, cudaGetLastError());
    h->launches += 2;
    return WN_OK;
}
