// math.cuh -- arithmetic policy of the two builds.
//
//   Math<double>: PARITY build.  Every operation is an individually rounded IEEE double op in the
//                 reference's order; the translation unit is compiled with -fmad=false so nvcc never
//                 contracts a*b+c.  The three places where the reference itself runs fused code (BLAS
//                 ddot / dgemv behind np.linalg.norm and R @ v, see DESIGN.md "numpy semantics") use
//                 explicit __fma_rn.
//   Math<float>:  THROUGHPUT build.  Fused multiply-adds where they are written (fmaf, mad), none added by the compiler
//                 (-fmad=false as well); short-range polynomial transcendentals.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace acb {

template <typename T> struct Math;

template <> struct Math<double> {
    using real = double;
    static constexpr bool kParity = true;
    static __device__ __forceinline__ double sq(double x) { return x * x; }  // numpy scalar x**2 (pow(x,2.0), 1 ulp)
    // a * b + c as the reference evaluates it: two roundings (this build is compiled with -fmad=false)
    static __device__ __forceinline__ double mad(double a, double b, double c) { return a * b + c; }
    static __device__ __forceinline__ double mul(double a, double b) { return a * b; }      // a product that feeds a sum
    static __device__ __forceinline__ double norm3(double u, double v, double w) {
        return ::sqrt(__fma_rn(w, w, __fma_rn(v, v, u * u)));                // np.linalg.norm -> ddot, fused tail
    }
    static __device__ __forceinline__ double norm2(double a, double b) { return ::sqrt(__fma_rn(b, b, a * a)); }
    static __device__ __forceinline__ double row_dot3(double r0, double r1, double r2, double x0, double x1, double x2) {
        return __fma_rn(r2, x2, __fma_rn(r0, x0, r1 * x1));                  // dgemv row as OpenBLAS evaluates it
    }
    static __device__ __forceinline__ void sincos(double x, double* s, double* c) { ::sincos(x, s, c); }
    static __device__ __forceinline__ double sin(double x) { return ::sin(x); }
    static __device__ __forceinline__ double cos(double x) { return ::cos(x); }
    static __device__ __forceinline__ double tan(double x) { return ::tan(x); }
    static __device__ __forceinline__ double atan2(double y, double x) { return ::atan2(y, x); }
    static __device__ __forceinline__ double asin(double x) { return ::asin(x); }
    static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
    static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double log(double x) { return ::log(x); }
    static __device__ __forceinline__ double div(double a, double b) { return a / b; }
    static __device__ __forceinline__ double fdiv(double a, double b) { return a / b; }     // guidance-level quotients
    static __device__ __forceinline__ double div_by(double a, double c) { return a / c; }   // division by a literal
    static __device__ __forceinline__ double exp_neg(double x) { return ::exp(x); }         // exp of a non-positive argument
    static __device__ __forceinline__ double cos_bounded(double x) { return ::cos(x); }
    static __device__ __forceinline__ double abs(double x) { return ::fabs(x); }
    static __device__ __forceinline__ double fmod(double a, double b) { return ::fmod(a, b); }
    static __device__ __forceinline__ bool finite(double x) { return isfinite(x); }
    // tan/cos of the pitch clipped to +-max_pitch (simplified_6dof.py:463-465)
    static __device__ __forceinline__ void tan_cos_safe(double theta_safe, double, double, bool, double* tn, double* cs) {
        *cs = ::cos(theta_safe);
        *tn = ::tan(theta_safe);
    }
    // sin/cos of the clipped angle of attack (simplified_6dof.py:369-373)
    static __device__ __forceinline__ void alpha_sincos(double alpha, double, double, bool, double* s, double* c) {
        ::sincos(alpha, s, c);
    }
    // roll/yaw := atan2(sin, cos) (simplified_6dof.py:266,270)
    static __device__ __forceinline__ double wrap_pi(double a) {
        double s, c;
        ::sincos(a, &s, &c);
        return ::atan2(s, c);
    }
};

// ---- float: short-range transcendental kernels --------------------------------------------------
// Flight angles stay within a few multiples of pi (roll/yaw are re-wrapped every substep, pitch is
// clamped), so a two-constant Cody-Waite reduction by pi/2 is exact enough; results are within
// ~1.5 ulp of the correctly rounded value on |x| < 1e4 (tests/test_math_f32.py measures it).
struct F32 {
    static __device__ __forceinline__ void sincos(float x, float* sn, float* cs) {
        const float j = rintf(x * 0.636619772367581343f);          // x / (pi/2)
        const int q = (int)j;
        float r = fmaf(j, -1.57079601287841796875f, x);            // pi/2 high (24 bits, 8 trailing zeros)
        r = fmaf(j, -3.1391647326017846e-07f, r);                  // pi/2 mid
        r = fmaf(j, -5.3903025299577648e-15f, r);                  // pi/2 low
        const float r2 = r * r;
        // minimax on [-pi/4, pi/4] (coefficients of the classic Cephes sinf/cosf kernels)
        float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
        ps = fmaf(ps, r2, -1.6666654611e-1f);
        const float s = fmaf(ps * r2, r, r);
        float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
        pc = fmaf(pc, r2, 4.166664568298827e-2f);
        const float c = fmaf(pc * r2, r2, fmaf(r2, -0.5f, 1.0f));
        const float a = (q & 1) ? c : s;
        const float b = (q & 1) ? s : c;
        *sn = (q & 2) ? -a : a;
        *cs = ((q + 1) & 2) ? -b : b;
    }
    // atan on [0,1] then octant fix-up; max err ~1.2 ulp
    static __device__ __forceinline__ float atan_unit(float t) {
        const float z = t * t;
        float p = fmaf(z, 0.00282363896258175373077393f, -0.0159569028764963150024414f);
        p = fmaf(p, z, 0.0425049886107444763183594f);
        p = fmaf(p, z, -0.0748900920152664184570312f);
        p = fmaf(p, z, 0.106347933411598205566406f);
        p = fmaf(p, z, -0.142027363181114196777344f);
        p = fmaf(p, z, 0.199926957488059997558594f);
        p = fmaf(p, z, -0.333331018686294555664062f);
        return fmaf(p * z, t, t);
    }
    static __device__ __forceinline__ float atan2(float y, float x) {
        const float ax = fabsf(x), ay = fabsf(y);
        const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
        float t = (mx == 0.0f) ? 0.0f : __fdividef(mn, mx);
        t = (mx == mn && mx > 1e30f) ? 1.0f : t;                      // inf/inf
        float a = atan_unit(t);
        a = (ay > ax) ? 1.57079637050628662109375f - a : a;
        a = (x < 0.0f) ? 3.14159274101257324f - a : a;
        return copysignf(a, y);
    }
    // asin via atan2(x, sqrt(1-x^2)) is wasteful; direct: asin(x) = x + x^3 P(x^2) on |x|<=0.5,
    // pi/2 - 2 asin(sqrt((1-|x|)/2)) beyond.
    static __device__ __forceinline__ float asin(float x) {
        const float ax = fabsf(x);
        const bool big = ax > 0.5f;
        const float z = big ? fmaf(ax, -0.5f, 0.5f) : ax * ax;
        const float t = big ? sqrtf(z) : ax;
        float p = fmaf(z, 4.2163199048e-2f, 2.4181311049e-2f);
        p = fmaf(p, z, 4.5470025998e-2f);
        p = fmaf(p, z, 7.4953002686e-2f);
        p = fmaf(p, z, 1.6666752422e-1f);
        float r = fmaf(p * z, t, t);
        r = big ? fmaf(r, -2.0f, 1.57079637050628662109375f) : r;
        return copysignf(r, x);
    }
};

template <> struct Math<float> {
    using real = float;
    static constexpr bool kParity = false;
    static __device__ __forceinline__ float sq(float x) { return x * x; }
    // a * b + c, ALWAYS fused: the throughput build leaves no multiply-add to the compiler's contraction heuristics, so that
    // two kernels (one aircraft per thread / a pair per thread) that inline the same function round it the same way
    static __device__ __forceinline__ float mad(float a, float b, float c) { return fmaf(a, b, c); }
    // a product that feeds a sum (or is stored AND summed): rounded on its own, never contracted into the addition
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float norm3(float u, float v, float w) { return sqrtf(fmaf(w, w, fmaf(v, v, u * u))); }
    static __device__ __forceinline__ float norm2(float a, float b) { return sqrtf(fmaf(b, b, a * a)); }
    static __device__ __forceinline__ float row_dot3(float r0, float r1, float r2, float x0, float x1, float x2) {
        return fmaf(r2, x2, fmaf(r0, x0, r1 * x1));
    }
    static __device__ __forceinline__ void sincos(float x, float* s, float* c) { F32::sincos(x, s, c); }
    static __device__ __forceinline__ float sin(float x) { float s, c; F32::sincos(x, &s, &c); return s; }
    static __device__ __forceinline__ float cos(float x) { float s, c; F32::sincos(x, &s, &c); return c; }
    static __device__ __forceinline__ float tan(float x) { float s, c; F32::sincos(x, &s, &c); return s / c; }
    static __device__ __forceinline__ float atan2(float y, float x) { return F32::atan2(y, x); }
    static __device__ __forceinline__ float asin(float x) { return F32::asin(x); }
    static __device__ __forceinline__ float exp(float x) { return ::expf(x); }
    static __device__ __forceinline__ float sqrt(float x) { return ::sqrtf(x); }
    static __device__ __forceinline__ float log(float x) { return ::logf(x); }
    static __device__ __forceinline__ float div(float a, float b) { return a / b; }
    // quotients of the guidance levels (turn radii, blend factors, unit vectors): MUFU.RCP x multiply, 2 ulp
    static __device__ __forceinline__ float fdiv(float a, float b) {
        float r;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
        return __fmul_rn(a, r);                       // (explicitly rounded: 1 - fdiv(..) must not become one FFMA in one kernel and two instructions in another)
    }
    // division by a literal: a multiplication by its reciprocal, folded at compile time (1 ulp from the quotient)
    static __device__ __forceinline__ float div_by(float a, float c) { return a * (1.0f / c); }
    // exp of a non-positive, bounded argument (the reward's stability terms): ex2.approx on x log2(e), 2 ulp
    static __device__ __forceinline__ float exp_neg(float x) { return __expf(x); }
    // cosine of a clipped bank / heading-sized angle (|x| <= 2 pi): the SFU's cos.approx, abs err ~5e-7
    static __device__ __forceinline__ float cos_bounded(float x) { return __cosf(x); }
    static __device__ __forceinline__ float abs(float x) { return ::fabsf(x); }
    static __device__ __forceinline__ float fmod(float a, float b) { return ::fmodf(a, b); }
    static __device__ __forceinline__ bool finite(float x) { return isfinite(x); }
    // When the pitch is inside its clamp (always, except RK4 stage states of an aircraft already at the
    // clamp) tan/cos of the clipped pitch are sin_theta/cos_theta and cos_theta: no second sincos.
    static __device__ __forceinline__ void tan_cos_safe(float theta_safe, float sin_theta, float cos_theta, bool clipped,
                                                        float* tn, float* cs) {
        float s = sin_theta, c = cos_theta;
        if (clipped) F32::sincos(theta_safe, &s, &c);
        *cs = c;
        *tn = s / c;
    }
    // sin/cos(atan2(w,u)) = (w,u)/hypot(w,u) when alpha was not clipped and u_safe > 0
    static __device__ __forceinline__ void alpha_sincos(float alpha, float w, float u_safe, bool direct, float* s, float* c) {
        if (direct) {
            const float inv = rsqrtf(fmaf(w, w, u_safe * u_safe));
            *s = w * inv;
            *c = u_safe * inv;
        } else {
            F32::sincos(alpha, s, c);
        }
    }
    // atan2(sin a, cos a) for |a| < 3*pi is a - 2*pi*round(a/(2*pi))
    static __device__ __forceinline__ float wrap_pi(float a) {
        const float k = rintf(a * 0.159154943091895346f);
        float r = fmaf(k, -6.28318548202514648f, a);
        r = fmaf(k, 1.74845553146951715e-07f, r);
        return r;
    }
};

// ---- scalar semantics shared by both builds (numpy / Python builtins the reference relies on) ----
// np.clip on a scalar: NaN propagates.
template <typename T> __device__ __forceinline__ T np_clip(T x, T lo, T hi) {
    T y = (x < lo) ? lo : x;
    y = (y > hi) ? hi : y;
    return y;  // NaN compares false twice -> returned unchanged
}
// Python max(a, b): a unless b > a
template <typename T> __device__ __forceinline__ T py_max(T a, T b) { return (b > a) ? b : a; }
template <typename T> __device__ __forceinline__ T np_sign(T x) {
    return (x > T(0)) ? T(1) : ((x < T(0)) ? T(-1) : ((x != x) ? x : T(0)));
}

}  // namespace acb
