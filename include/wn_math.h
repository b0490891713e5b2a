/* wn_math.h - float32 transcendental functions with ONE definition for nvcc (device) and gcc (host).
 *
 * Why: the WaveNet loop feeds every sample back into the network, so a 1-ulp difference in tanh/exp between
 * the CUDA kernels and the CPU checker is amplified into completely different audio after a few hundred
 * samples.  libm (glibc) and the CUDA math library round differently, so neither can be the common definition.
 * Everything below is built only from IEEE-754 correctly rounded operations (add, mul, fma, div) written with
 * explicit rounding intrinsics on the device (never contracted or re-associated by nvcc) and plain C on the
 * host (compile the host side with -ffp-contract=off -mfma), therefore device and host results are bit-identical.
 *
 * Accuracy (vs. float64 libm, measured in tests/test_c_oracle.py): expf/tanhf/sigmoid/logf within 2 ulp over the
 * ranges the path uses.  The reference computes the same quantities with tf.nn.tanh / tf.nn.sigmoid / tf.exp /
 * tf.log (modules.py:510, gaussian.py:47, mixture.py:92,105), i.e. Eigen or CUDA libm at a few ulp as well.
 */
#ifndef WN_MATH_H_
#define WN_MATH_H_

#include <stdint.h>
#include <string.h>

#if defined(__CUDA_ARCH__)
#define WN_HD __device__ __forceinline__
#define WN_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define WN_MUL(a, b) __fmul_rn((a), (b))
#define WN_ADD(a, b) __fadd_rn((a), (b))
#define WN_SUB(a, b) __fsub_rn((a), (b))
#define WN_DIV(a, b) __fdiv_rn((a), (b))
#define WN_F2U(x) __float_as_uint(x)
#define WN_U2F(x) __uint_as_float(x)
#else
#include <math.h>
#if defined(__CUDACC__)
#define WN_HD __host__ inline
#else
#define WN_HD static inline
#endif
#define WN_FMA(a, b, c) fmaf((a), (b), (c))
#define WN_MUL(a, b) ((a) * (b))
#define WN_ADD(a, b) ((a) + (b))
#define WN_SUB(a, b) ((a) - (b))
#define WN_DIV(a, b) ((a) / (b))
static inline uint32_t wn_f2u_(float x) { uint32_t u; memcpy(&u, &x, 4); return u; }
static inline float wn_u2f_(uint32_t u) { float x; memcpy(&x, &u, 4); return x; }
#define WN_F2U(x) wn_f2u_(x)
#define WN_U2F(x) wn_u2f_(x)
#endif

#define WN_SQRT_HALF 0.70710677f /* float32(np.sqrt(0.5)), wavenet.py:834, modules.py:518 */

WN_HD float wn_pow2i(int n) { /* 2^n for -126 <= n <= 127 */
    return WN_U2F((uint32_t)(n + 127) << 23);
}

/* e^x.  Cody-Waite reduction x = n ln2 + r, |r| <= ln2/2, degree-5 polynomial for (e^r - 1 - r)/r^2. */
WN_HD float wn_expf(float x) {
    if (!(x == x)) return x;
    if (x > 88.72283f) return WN_U2F(0x7f800000u);
    if (x < -103.9f) return 0.0f;
    const float magic = 12582912.0f; /* 1.5 * 2^23: adding it rounds to nearest integer */
    float t = WN_FMA(x, 1.44269504088896341f, magic);
    float n = WN_SUB(t, magic);
    int ni = (int)n;
    float r = WN_FMA(n, -0.693359375f, x);
    r = WN_FMA(n, 2.12194440e-4f, r);
    float p = 1.9875691500e-4f;
    p = WN_FMA(p, r, 1.3981999507e-3f);
    p = WN_FMA(p, r, 8.3334519073e-3f);
    p = WN_FMA(p, r, 4.1665795894e-2f);
    p = WN_FMA(p, r, 1.6666665459e-1f);
    p = WN_FMA(p, r, 5.0000001201e-1f);
    float r2 = WN_MUL(r, r);
    float y = WN_FMA(p, r2, r);
    y = WN_ADD(y, 1.0f);
    if (ni < -125) {
        y = WN_MUL(y, wn_pow2i(ni + 100));
        return WN_MUL(y, wn_pow2i(-100));
    }
    if (ni > 126) {
        y = WN_MUL(y, wn_pow2i(ni - 2));
        return WN_MUL(y, 4.0f);
    }
    return WN_MUL(y, wn_pow2i(ni));
}

/* natural log for x > 0 (x <= 0 returns -inf; the path only takes logs of uniforms in [1e-5, 1-1e-5] and of -log u). */
WN_HD float wn_logf(float x) {
    if (!(x == x)) return x;
    if (!(x > 0.0f)) return WN_U2F(0xff800000u);
    int e = 0;
    if (x < 1.17549435e-38f) { x = WN_MUL(x, 8388608.0f); e = -23; }
    uint32_t u = WN_F2U(x);
    if (u >= 0x7f800000u) return x;
    e += (int)(u >> 23) - 126;
    float m = WN_U2F((u & 0x007fffffu) | 0x3f000000u); /* [0.5, 1) */
    if (m < 0.707106781186547524f) {
        e -= 1;
        m = WN_SUB(WN_ADD(m, m), 1.0f);
    } else {
        m = WN_SUB(m, 1.0f);
    }
    float z = WN_MUL(m, m);
    float p = 7.0376836292e-2f;
    p = WN_FMA(p, m, -1.1514610310e-1f);
    p = WN_FMA(p, m, 1.1676998740e-1f);
    p = WN_FMA(p, m, -1.2420140846e-1f);
    p = WN_FMA(p, m, 1.4249322787e-1f);
    p = WN_FMA(p, m, -1.6668057665e-1f);
    p = WN_FMA(p, m, 2.0000714765e-1f);
    p = WN_FMA(p, m, -2.4999993993e-1f);
    p = WN_FMA(p, m, 3.3333331174e-1f);
    float y = WN_MUL(WN_MUL(p, m), z);
    float fe = (float)e;
    y = WN_FMA(fe, -2.12194440e-4f, y);
    y = WN_FMA(z, -0.5f, y);
    float r = WN_ADD(m, y);
    return WN_FMA(fe, 0.693359375f, r);
}

/* tanh: odd polynomial below 0.625, 1 - 2/(e^{2|x|}+1) above, saturating at 9.02. */
WN_HD float wn_tanhf(float x) {
    float ax = x < 0.0f ? -x : x;
    if (!(x == x)) return x;
    if (ax < 0.625f) {
        float z = WN_MUL(x, x);
        float p = -5.70498872745e-3f;
        p = WN_FMA(p, z, 2.06390887954e-2f);
        p = WN_FMA(p, z, -5.37397155531e-2f);
        p = WN_FMA(p, z, 1.33314422036e-1f);
        p = WN_FMA(p, z, -3.33332819422e-1f);
        return WN_FMA(WN_MUL(p, z), x, x);
    }
    float r;
    if (ax > 9.02f) {
        r = 1.0f;
    } else {
        float e = wn_expf(WN_ADD(ax, ax));
        r = WN_SUB(1.0f, WN_DIV(2.0f, WN_ADD(e, 1.0f)));
    }
    return x < 0.0f ? -r : r;
}

/* logistic sigmoid 1 / (1 + e^{-x}) (tf.nn.sigmoid, modules.py:510). */
WN_HD float wn_sigmoidf(float x) {
    return WN_DIV(1.0f, WN_ADD(1.0f, wn_expf(-x)));
}

/* mu-law expansion sign(y)/255 * (256^{|y|} - 1), mu fixed at 255 (util.py:52-68). */
WN_HD float wn_inv_mulawf(float y) {
    float ay = y < 0.0f ? -y : y;
    float p = wn_expf(WN_MUL(ay, 5.545177444479562f)); /* ln 256 */
    float v = WN_MUL(WN_SUB(p, 1.0f), 0.00392156862745098f);
    return y < 0.0f ? -v : (y > 0.0f ? v : 0.0f);
}

#endif /* WN_MATH_H_ */
