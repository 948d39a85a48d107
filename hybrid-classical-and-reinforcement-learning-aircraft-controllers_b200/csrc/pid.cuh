// pid.cuh -- the fp32 PID core, bit-for-bit the arithmetic of PIDController::compute
// (cpp/src/pid_controller.cpp:24-60).  The reference is built -O3 without -march, i.e. with no FMA
// (CMakeLists.txt:15-16), so every operation here is an explicitly rounded single op: on the device
// the __f*_rn intrinsics (never contracted, whatever -fmad says), on the host plain float ops in a
// translation unit compiled with -ffp-contract=off.
#pragma once
#include "../../include/acb200.h"
#include <cuda_runtime.h>

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

__host__ __device__ __forceinline__ float pid_compute(const acb200_pid_config& c, PidMem& m, float setpoint, float measurement,
                                                      float dt) {
    const float error = ACB_FSUB(setpoint, measurement);                                   // :26
    const float p_term = ACB_FMUL(c.kp, error);                                            // :29
    m.integral = pid_clamp(ACB_FADD(m.integral, ACB_FMUL(error, dt)), c.integral_min, c.integral_max);  // :32-33
    const float i_term = ACB_FMUL(c.ki, m.integral);                                       // :34
    const float derivative = (dt > 1e-6f) ? ACB_FDIV(ACB_FSUB(error, m.error_prev), dt) : 0.0f;      // :37-41
    m.derivative_filtered = ACB_FADD(ACB_FMUL(c.derivative_filter_alpha, derivative),
                                     ACB_FMUL(ACB_FSUB(1.0f, c.derivative_filter_alpha), m.derivative_filtered));  // :45-46
    const float d_term = ACB_FMUL(c.kd, m.derivative_filtered);                            // :48
    const float output = pid_clamp(ACB_FADD(ACB_FADD(p_term, i_term), d_term), c.output_min, c.output_max);  // :51-54
    m.error_prev = error;                                                                  // :57
    return output;
}

}  // namespace acb
