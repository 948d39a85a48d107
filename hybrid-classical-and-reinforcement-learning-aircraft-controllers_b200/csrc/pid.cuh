// pid.cuh -- the fp32 PID core, bit-for-bit the arithmetic of PIDController::compute
// (cpp/src/pid_controller.cpp:24-60).  The reference is built -O3 without -march, i.e. with no FMA
// (CMakeLists.txt:15-16), so every operation here is an explicitly rounded single op: on the device
// the __f*_rn intrinsics (never contracted, whatever -fmad says), on the host plain float ops in a
// translation unit compiled with -ffp-contract=off.
#pragma once
#include "../../include/acb200.h"
#include <cuda_runtime.h>
#include <stdint.h>

namespace acb {

#ifdef __CUDA_ARCH__
#define ACB_FADD(a, b) __fadd_rn((a), (b))
#define ACB_FSUB(a, b) __fsub_rn((a), (b))
#define ACB_FMUL(a, b) __fmul_rn((a), (b))
#define ACB_FDIV(a, b) __fdiv_rn((a), (b))
#else
#define ACB_FADD(a, b) ((a) + (b))
#define ACB_FSUB(a, b) ((a) - (b))
#define ACB_FMUL(a, b) ((a) * (b))
#define ACB_FDIV(a, b) ((a) / (b))
#endif

// PIDController::clamp = std::max(min_val, std::min(value, max_val)) (pid_controller.cpp:80-82);
// with value = NaN: std::min(NaN, max) keeps NaN, std::max(min, NaN) keeps min.
__host__ __device__ __forceinline__ float pid_clamp(float value, float min_val, float max_val) {
    const float m = (max_val < value) ? max_val : value;
    return (min_val < m) ? m : min_val;
}

// The three words of controller memory that influence later outputs.
struct PidMem {
    float integral, error_prev, derivative_filtered;
};

// The arithmetic of compute(); DIV supplies the one division, derivative = (error - error_prev) / dt  (:37-41)
template <typename DIV>
__host__ __device__ __forceinline__ float pid_compute_impl(const acb200_pid_config& c, PidMem& m, float setpoint, float measurement,
                                                           float dt, const DIV& div) {
    const float error = ACB_FSUB(setpoint, measurement);                                   // :26
    const float p_term = ACB_FMUL(c.kp, error);                                            // :29
    m.integral = pid_clamp(ACB_FADD(m.integral, ACB_FMUL(error, dt)), c.integral_min, c.integral_max);  // :32-33
    const float i_term = ACB_FMUL(c.ki, m.integral);                                       // :34
    const float derivative = (dt > 1e-6f) ? div(ACB_FSUB(error, m.error_prev)) : 0.0f;     // :37-41
    m.derivative_filtered = ACB_FADD(ACB_FMUL(c.derivative_filter_alpha, derivative),
                                     ACB_FMUL(ACB_FSUB(1.0f, c.derivative_filter_alpha), m.derivative_filtered));  // :45-46
    const float d_term = ACB_FMUL(c.kd, m.derivative_filtered);                            // :48
    const float output = pid_clamp(ACB_FADD(ACB_FADD(p_term, i_term), d_term), c.output_min, c.output_max);  // :51-54
    m.error_prev = error;                                                                  // :57
    return output;
}

struct PidDivExact {
    float dt;
    __host__ __device__ __forceinline__ float operator()(float x) const { return ACB_FDIV(x, dt); }
};
__host__ __device__ __forceinline__ float pid_compute(const acb200_pid_config& c, PidMem& m, float setpoint, float measurement,
                                                      float dt) {
    return pid_compute_impl(c, m, setpoint, measurement, dt, PidDivExact{dt});
}

// ---- the same division, correctly rounded, without the divider ------------------------------------------------------------
// Every PID of a kernel divides by the same control period.  With r = RN(1 / dt) held in the kernel's parameters,
//     q0 = RN(x r);  rem = x - q0 dt (exact in one FMA);  q = RN(q0 + rem r)
// is the correctly rounded x / dt (Markstein's theorem: r correctly rounded, q0 within an ulp) as long as nothing under- or
// overflows on the way: |x| in [2^-100, 2^100] for dt in [2^-20, 16].  tests/test_gpu_pid_division.py sweeps every exponent,
// zeros, subnormals, infinities and NaN through a kernel and compares bit for bit with IEEE division; a CPU sweep over all
// 2^23 mantissas of a binade for 1 500 divisors found no exception.  x = +-0 is handled by giving q the sign of x; anything
// else outside the window (never seen in flight: an error difference below 1e-30 or above 1e30) takes the real division, out
// of line.  A period outside the proven range turns the window off (every non-zero x divides for real).
// Why: __fdiv_rn costs 16 issue slots and one MUFU per PID, and sends the WHOLE WARP through a 30-instruction slow path
// whenever one lane's numerator is zero -- which "hold the current yaw" makes the normal case for the yaw loop.
struct PidDt {
    float dt;          // the control period, as the caller gave it
    float dtf, rdt;    // divisor and reciprocal of the fused sequence (1, 1 with the window off)
    uint32_t tiny_m1;  // (|x| bits - 1) <  tiny_m1  <=>  0 < |x| < 2^-100
    uint32_t huge;     //  |x| bits      >  huge     <=>  |x| > 2^100, infinity, NaN
    __host__ __device__ operator float() const { return dt; }
};
static inline PidDt make_pid_dt(float dt) {
    PidDt t;
    t.dt = dt;
    const bool proven = dt >= 9.5367431640625e-07f && dt <= 16.0f;        // also false for NaN
    t.dtf = proven ? dt : 1.0f;
    t.rdt = proven ? (float)(1.0 / (double)dt) : 1.0f;
    t.tiny_m1 = proven ? 0x0d800000u - 1u : 0xffffffffu;
    t.huge = 0x71800000u;
    return t;
}
#ifdef __CUDACC__
static __device__ __noinline__ float pid_div_out_of_line(float x, float dt) { return __fdiv_rn(x, dt); }
struct PidDivFused {
    const PidDt& t;
    __device__ __forceinline__ float operator()(float x) const {
        const float q0 = __fmul_rn(x, t.rdt);
        const float rem = __fmaf_rn(-q0, t.dtf, x);
        const uint32_t xb = __float_as_uint(x), a = xb & 0x7fffffffu;
        float q = __uint_as_float((__float_as_uint(__fmaf_rn(rem, t.rdt, q0)) & 0x7fffffffu) | (xb & 0x80000000u));
        if ((a - 1u) < t.tiny_m1 || a > t.huge) q = pid_div_out_of_line(x, t.dt);
        return q;
    }
};
__device__ __forceinline__ float pid_compute(const acb200_pid_config& c, PidMem& m, float setpoint, float measurement,
                                             const PidDt& t) {
    return pid_compute_impl(c, m, setpoint, measurement, t.dt, PidDivFused{t});
}
#endif

}  // namespace acb
