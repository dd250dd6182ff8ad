// scg_math.cuh -- precision-generic math wrappers and the Philox4x32-10 draw contract.
// The draw contract (key = (seed, env id), counter = (index, episode, stream, block),
// u01 / Box-Muller / randint transforms) is specified in oracle/philox.py; the stream ids
// below must stay in sync with it.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace scg {

#define SCG_DEV __device__ __forceinline__

SCG_DEV float  m_sqrt(float x)  { return sqrtf(x); }
SCG_DEV double m_sqrt(double x) { return sqrt(x); }
SCG_DEV float  m_rsqrt(float x)  { return rsqrtf(x); }
SCG_DEV double m_rsqrt(double x) { return 1.0 / sqrt(x); }
SCG_DEV void m_sincos(float x, float* s, float* c)   { sincosf(x, s, c); }
SCG_DEV void m_sincos(double x, double* s, double* c) { sincos(x, s, c); }
SCG_DEV float  m_sin(float x)  { return sinf(x); }
SCG_DEV double m_sin(double x) { return sin(x); }
SCG_DEV float  m_atan2(float y, float x)   { return atan2f(y, x); }
SCG_DEV double m_atan2(double y, double x) { return atan2(y, x); }
SCG_DEV float  m_asin(float x)  { return asinf(x); }
SCG_DEV double m_asin(double x) { return asin(x); }
SCG_DEV float  m_exp(float x)  { return expf(x); }
SCG_DEV double m_exp(double x) { return exp(x); }
SCG_DEV float  m_log(float x)  { return logf(x); }
SCG_DEV double m_log(double x) { return log(x); }
SCG_DEV float  m_pow(float x, float y)   { return powf(x, y); }
SCG_DEV double m_pow(double x, double y) { return pow(x, y); }
SCG_DEV float  m_abs(float x)  { return fabsf(x); }
SCG_DEV double m_abs(double x) { return fabs(x); }
SCG_DEV float  m_max(float a, float b)   { return fmaxf(a, b); }
SCG_DEV double m_max(double a, double b) { return fmax(a, b); }
template <typename real> SCG_DEV real m_max3(real a, real b, real c) { return m_max(a, m_max(b, c)); }   // FMNMX3 in fp32
SCG_DEV float  m_min(float a, float b)   { return fminf(a, b); }
SCG_DEV double m_min(double a, double b) { return fmin(a, b); }
SCG_DEV float  m_fma(float a, float b, float c)    { return fmaf(a, b, c); }
SCG_DEV double m_fma(double a, double b, double c) { return fma(a, b, c); }
SCG_DEV float  m_rint(float x)  { return rintf(x); }
SCG_DEV double m_rint(double x) { return rint(x); }
// rint() by the 1.5 * 2^23 magic constant: two fma-pipe adds instead of FRND, which issues on the
// quarter-rate XU pipe (valid for |x| < 2^22; the kernels only feed it angles / (2 pi)).
SCG_DEV float fast_rint(float x) { return (x + 12582912.f) - 12582912.f; }
SCG_DEV double fast_rint(double x) { return rint(x); }
// sin/cos for |x| <= pi + small (the attitude angles after wrap_pi): quadrant reduction with a two-term
// pi/2 and the cephes single-precision minimax polynomials on [-pi/4, pi/4] (about 1 ulp).  Compact:
// no large-argument (Payne-Hanek) path, no F2I / FRND, so the substep loop stays small in the
// instruction cache and on the fma pipe.
SCG_DEV void sincos_pi(float x, float* s, float* c) {
    const float kf = fmaf(x, 0.63661977236758134f, 12582912.f);
    const int q = __float_as_int(kf);          // mantissa = 0x400000 + k: the low bits are k mod 4
    const float k = kf - 12582912.f;
    const float r = fmaf(-k, -4.371139000186243e-08f, fmaf(-k, 1.5707963705062866f, x));
    const float z = r * r;
    const float sp = fmaf(r * z, fmaf(z, fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f), -1.6666654611e-1f), r);
    const float cp = fmaf(z * z, fmaf(z, fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f), 4.166664568298827e-2f),
                          fmaf(z, -0.5f, 1.0f));
    const bool swap = (q & 1) != 0;
    const float ss = swap ? cp : sp, cc = swap ? sp : cp;
    const uint32_t uq = static_cast<uint32_t>(q);
    *s = __uint_as_float(__float_as_uint(ss) ^ ((uq & 2u) << 30));
    *c = __uint_as_float(__float_as_uint(cc) ^ (((uq + 1u) & 2u) << 30));
}
SCG_DEV void sincos_pi(double x, double* s, double* c) { sincos(x, s, c); }
// sqrt for the fp32 fast paths: MUFU.SQRT (about 1 ulp) instead of the IEEE sequence with its
// denormal / special-case branch
SCG_DEV float sqrt_fast(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
SCG_DEV double sqrt_fast(double x) { return sqrt(x); }
// atan2 for the fp32 fast paths: a = min/max in [0, 1], odd polynomial a * P(a^2) (degree 15, fitted
// to 1.2e-7 max abs error in fp32, tools/fit_atan.py), octant fix-ups by selects.  About 25
// instructions, one MUFU.RCP, no branches (libdevice's atan2f is ~3x that with a division).
SCG_DEV float atan2_fast(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = mx > 0.f ? __fdividef(mn, mx) : 0.f;
    const float z = a * a;
    float p = -4.3554287404e-03f;
    p = fmaf(p, z, 2.3040225729e-02f);
    p = fmaf(p, z, -5.7773731649e-02f);
    p = fmaf(p, z, 9.7942464054e-02f);
    p = fmaf(p, z, -1.3976587355e-01f);
    p = fmaf(p, z, 1.9962705672e-01f);
    p = fmaf(p, z, -3.3331659436e-01f);
    float r = fmaf(p * z, a, a);
    r = ay > ax ? 1.57079632679489661923f - r : r;
    r = x < 0.f ? 3.14159265358979323846f - r : r;
    return copysignf(r, y);
}
SCG_DEV double atan2_fast(double y, double x) { return atan2(y, x); }
// np.clip(x, lo, hi) == minimum(maximum(x, lo), hi)
template <typename real> SCG_DEV real m_clip(real x, real lo, real hi) { return m_min(m_max(x, lo), hi); }

// Packed fp32 pairs (sm_100a FFMA2 / FADD2 / FMUL2: two IEEE fp32 operations per lane in ONE issue slot; the fused
// rollout is issue-bound, not pipe-bound).  Each half is the same round-to-nearest operation as the scalar instruction, so
// pairing changes no result.  A pair lives in an aligned 64-bit register pair; ptxas folds pk2 of two equal values into
// the broadcast operand form (R.F32 / UR.F32 / immediate) and reads halves in place.
// one 16-byte row of kernel-parameter constants as ONE load (LDCU.128 into uniform registers)
template <typename real> struct Row4 { real x, y, z, w; };
SCG_DEV Row4<float> ldc4(const float* p) { const float4 v = *reinterpret_cast<const float4*>(p); return {v.x, v.y, v.z, v.w}; }
SCG_DEV Row4<double> ldc4(const double* p) { return {p[0], p[1], p[2], p[3]}; }
// the same row from global memory through the read-only path
SCG_DEV Row4<float> ldg4(const float* p) { const float4 v = __ldg(reinterpret_cast<const float4*>(p)); return {v.x, v.y, v.z, v.w}; }
SCG_DEV Row4<double> ldg4(const double* p) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(p)), b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return {a.x, a.y, b.x, b.y};
}
typedef unsigned long long f32x2;
SCG_DEV f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
SCG_DEV f32x2 bc2(float v) { return pk2(v, v); }
SCG_DEV float lo2(f32x2 v) { return __uint_as_float(unsigned(v)); }
SCG_DEV float hi2(f32x2 v) { return __uint_as_float(unsigned(v >> 32)); }
SCG_DEV f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
SCG_DEV f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
SCG_DEV f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

template <typename real> struct Consts;
template <> struct Consts<float> {
    static constexpr float pi = 3.14159265358979323846f;
    static constexpr float two_pi = 6.28318530717958647692f;
    static constexpr float half_pi = 1.57079632679489661923f;
};
template <> struct Consts<double> {
    static constexpr double pi = 3.14159265358979323846;
    static constexpr double two_pi = 6.28318530717958647692;
    static constexpr double half_pi = 1.57079632679489661923;
};

// ---------------------------------------------------------------- Philox4x32-10
constexpr uint32_t STREAM_RESET_INIT = 0, STREAM_RESET_INERTIA = 1, STREAM_RESET_OFFSETS = 2,
                   STREAM_RESET_GATES = 3, STREAM_RESET_OBSTACLES = 4, STREAM_DISTURB_BASE = 16;

struct U4 { uint32_t x, y, z, w; };

SCG_DEV U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    return U4{c0, c1, c2, c3};
}

template <typename real> SCG_DEV real u01(uint32_t w) {
    return (static_cast<real>(w >> 8) + real(0.5)) * real(5.9604644775390625e-08);  // 2^-24
}
template <typename real> SCG_DEV void normal_pair(uint32_t w0, uint32_t w1, real* n0, real* n1) {
    const real r = m_sqrt(real(-2) * m_log(u01<real>(w0)));
    real s, c;
    m_sincos(Consts<real>::two_pi * u01<real>(w1), &s, &c);
    *n0 = r * c;
    *n1 = r * s;
}
// Box-Muller for the fp32 fast paths: MUFU.LG2 / MUFU.SQRT / MUFU.SIN / MUFU.COS instead of libdevice's logf / sqrtf / sincosf
// (~100 instructions with their slow-path branches).  The angle is taken in [-pi, pi): cos(2 pi u) = -cos(2 pi (u - 1/2)).
// Same draws as normal_pair up to ~1e-6 relative (|ln u| >= 6e-8 keeps the MUFU.LG2 error relative to the radius small).
template <typename real, bool FAST> SCG_DEV void normal_pair_f(uint32_t w0, uint32_t w1, real* n0, real* n1) {
    if constexpr (FAST && sizeof(real) == 4) {
        const float r = sqrt_fast(-2.f * __logf(u01<float>(w0)));
        float s, c;
        __sincosf(Consts<float>::two_pi * (u01<float>(w1) - 0.5f), &s, &c);
        *n0 = -(r * c);
        *n1 = -(r * s);
    } else normal_pair<real>(w0, w1, n0, n1);
}
template <typename real, int DIM, bool FAST> SCG_DEV void draw_normals_f(const uint32_t seed, const uint32_t env, uint32_t index, uint32_t episode,
                                                                         uint32_t stream, real (&z)[DIM]) {
#pragma unroll
    for (int b = 0; b < (DIM + 3) / 4; ++b) {
        const U4 w = philox4x32_10(index, episode, stream, b, seed, env);
        real n[4];
        normal_pair_f<real, FAST>(w.x, w.y, &n[0], &n[1]);
        if (4 * b + 2 < DIM) normal_pair_f<real, FAST>(w.z, w.w, &n[2], &n[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * b + j < DIM) z[4 * b + j] = n[j];
    }
}
// sin / cos of any angle for the fp32 fast paths: wrap + ~1-ulp polynomial pair (no libdevice slow path)
SCG_DEV void sincos_any(float x, float* s, float* c) {
    const float k = fast_rint(x * 0.15915494309189535f);
    sincos_pi(fmaf(-k, -1.7484555e-07f, fmaf(-k, 6.2831855f, x)), s, c);
}
SCG_DEV void sincos_any(double x, double* s, double* c) { sincos(x, s, c); }
// decay^k for a small non-negative integer k (impulse disturbances, disturbances.py:133-139): exp2(k log2 decay) in fp32
SCG_DEV float pow_int(float decay, float l2decay, int k) { (void)decay; return exp2f(float(k) * l2decay); }
SCG_DEV double pow_int(double decay, double l2decay, int k) { (void)l2decay; return pow(decay, double(k)); }

SCG_DEV int32_t randint(uint32_t w, int32_t n) { return static_cast<int32_t>(__umulhi(w, static_cast<uint32_t>(n))); }

// value i (of a vector draw) lives in word i % 4 of block i / 4
SCG_DEV uint32_t word_of(const U4& b, int i) { return (i & 3) == 0 ? b.x : (i & 3) == 1 ? b.y : (i & 3) == 2 ? b.z : b.w; }

struct RngKey { uint32_t seed, env; };

template <typename real, int DIM> SCG_DEV void draw_uniforms(const RngKey& k, uint32_t index, uint32_t episode,
                                                              uint32_t stream, real (&u)[DIM]) {
#pragma unroll
    for (int b = 0; b < (DIM + 3) / 4; ++b) {
        const U4 w = philox4x32_10(index, episode, stream, b, k.seed, k.env);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * b + j < DIM) u[4 * b + j] = u01<real>(word_of(w, j));
    }
}
template <typename real, int DIM> SCG_DEV void draw_normals(const RngKey& k, uint32_t index, uint32_t episode,
                                                             uint32_t stream, real (&z)[DIM]) {
#pragma unroll
    for (int b = 0; b < (DIM + 3) / 4; ++b) {
        const U4 w = philox4x32_10(index, episode, stream, b, k.seed, k.env);
        real n[4];
        normal_pair<real>(w.x, w.y, &n[0], &n[1]);
        normal_pair<real>(w.z, w.w, &n[2], &n[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (4 * b + j < DIM) z[4 * b + j] = n[j];
    }
}

}  // namespace scg
