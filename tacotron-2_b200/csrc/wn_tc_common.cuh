// wn_tc_common.cuh - device helpers shared by the tensor-core loop engines (wn_loop_tc.cu: one CTA per layer, 128-channel block;
// wn_loop_tc2.cu: four CTAs per layer, 256-channel block): {value, tag} mailbox words, the watchdog of every wait, distributed
// shared memory, fp16 (hi, lo) activation images, the gate non-linearity, the samplers.  Included inside each file's anonymous
// namespace.
#pragma once

constexpr unsigned long long WATCHDOG_NS_DEFAULT = 20ull * 1000 * 1000 * 1000;   // WN_WATCHDOG_MS overrides it (tests)

// ---- {value, tag} words -----------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long pack_ll(float v, unsigned tag) {
    return ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
}
__device__ __forceinline__ void st_ll2(unsigned long long *p, float v0, float v1, unsigned tag) {
    const unsigned long long a = pack_ll(v0, tag), b = pack_ll(v1, tag);
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void st_ll1(unsigned long long *p, float v, unsigned tag) {
    const unsigned long long a = pack_ll(v, tag);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(a) : "memory");
}
__device__ __forceinline__ void ld_ll2(const unsigned long long *p, unsigned long long &a, unsigned long long &b) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// shared watchdog of every wait in this kernel: a waiter that sees the abort flag, or that has waited WATCHDOG_NS of wall
// clock, gives up (and tells everybody else); the launch then ends with wrong output and WN_ERR_WATCHDOG instead of a hang
struct Watch {
    int *flag;
    unsigned spins = 0;
    unsigned long long t0 = 0;
    __device__ __forceinline__ bool expired() {
        if ((++spins & 0x3ffu) != 0) return false;
        if (*(volatile int *)flag) return true;
        const unsigned long long now = global_timer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > *reinterpret_cast<const volatile unsigned long long *>(flag + 2)) { atomicExch(flag, 1); return true; }
        return false;
    }
};
// the first `nu` of NU units of two words, `stride` words apart (the others read as 0); true if the watchdog fired
template <int NU>
__device__ __forceinline__ bool poll_units(const unsigned long long *p, long long stride, unsigned tag, float *v, int nu, int *abort_flag) {
    Watch wd{abort_flag};
    for (;;) {
        unsigned long long w[2 * NU];
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (u < nu) ld_ll2(p + u * stride, w[2 * u], w[2 * u + 1]);
        bool ok = true;
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (u < nu) ok = ok && ((unsigned)(w[2 * u] >> 32) == tag) && ((unsigned)(w[2 * u + 1] >> 32) == tag);
        if (ok) {
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                v[2 * u] = u < nu ? __uint_as_float((unsigned)(w[2 * u] & 0xffffffffu)) : 0.0f;
                v[2 * u + 1] = u < nu ? __uint_as_float((unsigned)(w[2 * u + 1] & 0xffffffffu)) : 0.0f;
            }
            return false;
        }
        if (wd.expired()) return true;
    }
}
// poll_units that also says whether the words were there at the first look (the reader is behind the writer)
template <int NU>
__device__ __forceinline__ bool poll_units_first(const unsigned long long *p, long long stride, unsigned tag, float *v, int nu, int *abort_flag, bool &first) {
    Watch wd{abort_flag};
    first = true;
    for (;;) {
        unsigned long long w[2 * NU];
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (u < nu) ld_ll2(p + u * stride, w[2 * u], w[2 * u + 1]);
        bool ok = true;
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (u < nu) ok = ok && ((unsigned)(w[2 * u] >> 32) == tag) && ((unsigned)(w[2 * u + 1] >> 32) == tag);
        if (ok) {
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                v[2 * u] = u < nu ? __uint_as_float((unsigned)(w[2 * u] & 0xffffffffu)) : 0.0f;
                v[2 * u + 1] = u < nu ? __uint_as_float((unsigned)(w[2 * u + 1] & 0xffffffffu)) : 0.0f;
            }
            return false;
        }
        first = false;
        if (wd.expired()) return true;
    }
}
// ONE look at LB mailboxes `box_stride` words apart (the first `nl` of them), all loads in flight together: returns how many
// leading mailboxes were complete (their values in v).  A reader that is behind pays one L2 round trip per LB layers this way.
template <int NU, int LB>
__device__ __forceinline__ int look_boxes(const unsigned long long *p, long long box_stride, long long stride, unsigned tag, float (&v)[LB][2 * NU],
                                          int nu, int nl) {
    unsigned long long w[LB][2 * NU];
#pragma unroll
    for (int b = 0; b < LB; ++b)
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (b < nl && u < nu) ld_ll2(p + b * box_stride + u * stride, w[b][2 * u], w[b][2 * u + 1]);
    int m = 0;
    bool run = true;
#pragma unroll
    for (int b = 0; b < LB; ++b) {
        bool ok = b < nl;
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (b < nl && u < nu) ok = ok && ((unsigned)(w[b][2 * u] >> 32) == tag) && ((unsigned)(w[b][2 * u + 1] >> 32) == tag);
        run = run && ok;
        m += run ? 1 : 0;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            v[b][2 * u] = (b < nl && u < nu) ? __uint_as_float((unsigned)(w[b][2 * u] & 0xffffffffu)) : 0.0f;
            v[b][2 * u + 1] = (b < nl && u < nu) ? __uint_as_float((unsigned)(w[b][2 * u + 1] & 0xffffffffu)) : 0.0f;
        }
    }
    return m;
}
__device__ __forceinline__ bool mbar_wait_wd(uint32_t bar, uint32_t parity, int *abort_flag) {
    Watch wd{abort_flag};
    while (!umma::mbar_try(bar, parity))
        if (wd.expired()) return true;
    return false;
}

// ---- thread-block cluster / distributed shared memory: consecutive stages of a replica that share a cluster hand x over
// with st.async straight into the next CTA's shared memory (counted in bytes on that CTA's mbarrier) ----------------------
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_nctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_async2(uint32_t raddr, float v0, float v1, uint32_t rmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                 ::"r"(raddr), "f"(v0), "f"(v1), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t raddr, uint32_t v, uint32_t rmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr), "r"(v), "r"(rmbar) : "memory");
}
__device__ __forceinline__ bool mbar_try_cluster(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ bool mbar_wait_cluster_wd(uint32_t bar, uint32_t parity, int *abort_flag) {
    Watch wd{abort_flag};
    while (!mbar_try_cluster(bar, parity))
        if (wd.expired()) return true;
    return false;
}

// ---- activation images: [K/8][NB][8] fp16, hi plane then lo plane; one thread per channel writes 2-byte elements ---------
// v -> (hi, lo) with hi = v rounded to 11 significant bits (what fp16 holds), lo = fp16(v - hi): hi + lo carries 22 bits of v, so
// the three MMA passes hi*hi + hi*lo + lo*hi reproduce the fp32 product to ~2^-22.  hi is rounded with integer operations
// (round to nearest even on bit 13): exactly representable in fp16 for |v| in [6.1e-5, 65504], which the conversions below
// then pack without a second rounding; below that range the absolute error is < 6e-8.
__device__ __forceinline__ float round11(float v) {
    const uint32_t u = __float_as_uint(v);
    return __uint_as_float((u + 0x00000fffu + ((u >> 13) & 1u)) & 0xffffe000u);
}
// WIDE images (groups of <= 8 utterances with N = 16 MMAs): the unused half of the columns carries the lo parts, i.e. plane 0 holds
// hi at column n and lo at column n + 8, plane 1 holds hi at column n + 8 (columns 0..7 stay zero).  One MMA of A_hi with plane 0
// then yields A_hi x_hi (columns 0..7) AND A_hi x_lo (columns 8..15), a second one of A_lo with plane 1 adds A_lo x_hi to columns
// 8..15: two MMAs per k step instead of three; the epilogue adds column n + 8 to column n.  The cross terms (2^-11 of the main
// term) accumulate among themselves: the tensor core truncates when it aligns an addend to the accumulator, which costs them
// bits when they are added one by one to the large main sum (measured on the 256-channel block: 1.5-2x smaller end-to-end error
// than with a common accumulator; an exact tanh / sigmoid instead of gate_fast changes nothing).
template <bool WIDE>
__device__ __forceinline__ void img_put(unsigned char *img, int plane, int lbo, int j, int n, float v) {
    const float hf = round11(v);
    const int off = (j >> 3) * lbo + n * 16 + (j & 7) * 2;
    const unsigned short h = __half_as_ushort(__float2half_rn(hf)), l = __half_as_ushort(__float2half_rn(__fsub_rn(v, hf)));
    *reinterpret_cast<unsigned short *>(img + off) = h;
    if (WIDE) {
        *reinterpret_cast<unsigned short *>(img + off + 8 * 16) = l;
        *reinterpret_cast<unsigned short *>(img + plane + off + 8 * 16) = h;
    } else {
        *reinterpret_cast<unsigned short *>(img + plane + off) = l;
    }
}
// two neighbouring utterances n, n + 1 of one channel: one packed conversion per plane (float -> half conversions run on the
// 16-lane special-function pipe, which bounds the epilogues of this kernel)
template <bool WIDE>
__device__ __forceinline__ void img_put2(unsigned char *p /* element (j, n) of plane 0 */, int plane, float v0, float v1) {
    const float h0 = round11(v0), h1 = round11(v1);
    uint32_t h, l;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(h1), "f"(h0));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(l) : "f"(__fsub_rn(v1, h1)), "f"(__fsub_rn(v0, h0)));
    *reinterpret_cast<unsigned short *>(p) = (unsigned short)h;
    *reinterpret_cast<unsigned short *>(p + 16) = (unsigned short)(h >> 16);
    if (WIDE) {
        *reinterpret_cast<unsigned short *>(p + 8 * 16) = (unsigned short)l;
        *reinterpret_cast<unsigned short *>(p + 9 * 16) = (unsigned short)(l >> 16);
        *reinterpret_cast<unsigned short *>(p + plane + 8 * 16) = (unsigned short)h;
        *reinterpret_cast<unsigned short *>(p + plane + 9 * 16) = (unsigned short)(h >> 16);
    } else {
        *reinterpret_cast<unsigned short *>(p + plane) = (unsigned short)l;
        *reinterpret_cast<unsigned short *>(p + plane + 16) = (unsigned short)(l >> 16);
    }
}
// NH accumulator columns of this thread's lane from column `taddr` on; WIDE: plus the partial sums 8 columns further
template <int NH, bool WIDE>
__device__ __forceinline__ void acc_load(uint32_t taddr, float (&v)[NH]) {
    if constexpr (NH == 8) {
        umma::tmem_ld8(taddr, v);
        if constexpr (WIDE) {   // at most 4 utterances per thread: 4 columns (a wider load could leave the allocation)
            float w[4];
            umma::tmem_ld4(taddr + 8, w);
            umma::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 4; ++i) v[i] = __fadd_rn(v[i], w[i]);
            return;
        }
    } else {
        umma::tmem_ld16(taddr, v);
    }
    umma::tmem_ld_wait();
}

// tanh(t) * sigmoid(s) (modules.py:510) with ex2.approx / rcp.approx; same form as wn_step_umma.cu
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float gate_fast(float t, float s) {
    // tanh(t) * sigmoid(s) = sign(t) (1 - E) / ((1 + E)(1 + F)), E = exp(-2|t|) in (0, 1], F = exp(-s): two ex2 and ONE rcp per
    // element (the MUFU pipe, 16 lanes per clock and SM, bounds this epilogue); F = inf gives rcp(inf) = 0, the right limit
    constexpr float LOG2E = 1.4426950408889634f;
    const float at = fabsf(t);
    const float E = ex2_approx(__fmul_rn(-2.0f * LOG2E, at)), F = ex2_approx(__fmul_rn(-LOG2E, s));
    const float den = __fmul_rn(__fadd_rn(1.0f, E), __fadd_rn(1.0f, F));
    const float r = rcp_approx(den);
    // 1 - E cancels for small |t|: tanh(|t|) (1 + E) from the polynomial of wn_tanhf instead (select, no branch)
    const float z = __fmul_rn(at, at);
    float p = -5.70498872745e-3f;
    p = __fmaf_rn(p, z, 2.06390887954e-2f);
    p = __fmaf_rn(p, z, -5.37397155531e-2f);
    p = __fmaf_rn(p, z, 1.33314422036e-1f);
    p = __fmaf_rn(p, z, -3.33332819422e-1f);
    const float small = __fmul_rn(__fmaf_rn(__fmul_rn(p, z), at, at), __fadd_rn(1.0f, E));
    const float num = at < 0.625f ? small : __fsub_rn(1.0f, E);
    return copysignf(__fmul_rn(num, r), t);
}

__device__ __forceinline__ float tc_sample_gaussian(float log_scale_min_gauss, const float *out, float eps) {   // gaussian.py:39-52
    const float mean = out[0];
    const float ls = out[1] > log_scale_min_gauss ? out[1] : log_scale_min_gauss;
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), eps));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}
__device__ float tc_sample_mol(float log_scale_min, int O, const float *out, const float *u) {   // mixture.py:76-107
    const int K = O / 3;
    int best = 0;
    float bestv = 0.0f;
    for (int k = 0; k < K; ++k) {
        const float v = __fsub_rn(out[k], wn_logf(-wn_logf(u[k])));
        if (k == 0 || v > bestv) { bestv = v; best = k; }
    }
    const float mean = out[K + best], lsr = out[2 * K + best];
    const float ls = lsr > log_scale_min ? lsr : log_scale_min;
    const float u2 = u[K];
    const float lg = __fsub_rn(wn_logf(u2), wn_logf(__fsub_rn(1.0f, u2)));
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), lg));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

// two independent IEEE round-to-nearest fp32 fmas in one instruction (FFMA2)
__device__ __forceinline__ float2 ffma2v(float2 a, float2 b, float2 c) {
    float2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(*reinterpret_cast<unsigned long long *>(&d))
        : "l"(*reinterpret_cast<const unsigned long long *>(&a)), "l"(*reinterpret_cast<const unsigned long long *>(&b)),
          "l"(*reinterpret_cast<const unsigned long long *>(&c)));
    return d;
}
