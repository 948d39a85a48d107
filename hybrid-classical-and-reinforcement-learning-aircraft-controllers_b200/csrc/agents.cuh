// agents.cuh -- the cascaded control agents for one aircraft (controllers/rate_agent.py:65-124,
// controllers/attitude_agent.py:86-154, controllers/utils/pid_utils.py:49-65).
// Commands and measurements are clipped in the physics type, narrowed to float at the point where the
// reference builds acb.Vector3 (pybind11 narrowing), run through the fp32 PID, and widened again.
#pragma once
#include "math.cuh"
#include "params.cuh"
#include "pid.cuh"

namespace acb {

// wrap_angle: (a + pi) % (2 pi) - pi with Python's floored modulo (range [-pi, pi))
template <typename T> __device__ __forceinline__ T wrap_angle(T a) {
    using M = Math<T>;
    const T two_pi = T(2) * T(3.141592653589793);
    T m = M::fmod(a + T(3.141592653589793), two_pi);
    if (m != T(0)) {
        if (m < T(0)) m += two_pi;
    } else {
        m = T(0);
    }
    return m - T(3.141592653589793);
}

// RateAgent.compute_action.  surf = elevator, aileron, rudder, throttle (ControlSurfaces.to_array order)
template <typename T>
__device__ __forceinline__ void rate_agent(const CtrlParams<T>& C, PidMem (&mem)[3], T p_cmd_in, T q_cmd_in, T r_cmd_in,
                                           T throttle, T p, T q, T r, float dt, T (&surf)[4]) {
    const T p_cmd = np_clip(p_cmd_in, -C.max_roll_rate, C.max_roll_rate);      // rate_agent.py:92-94
    const T q_cmd = np_clip(q_cmd_in, -C.max_pitch_rate, C.max_pitch_rate);
    const T r_cmd = np_clip(r_cmd_in, -C.max_yaw_rate, C.max_yaw_rate);
    const float o_roll = pid_compute(C.rate[0], mem[0], (float)p_cmd, (float)p, dt);   // :96-104
    const float o_pitch = pid_compute(C.rate[1], mem[1], (float)q_cmd, (float)q, dt);
    const float o_yaw = pid_compute(C.rate[2], mem[2], (float)r_cmd, (float)r, dt);
    surf[1] = np_clip((T)o_roll, T(-1), T(1));                                   // :111-122
    surf[0] = np_clip(-(T)o_pitch, T(-1), T(1));
    surf[2] = np_clip(-(T)o_yaw, T(-1), T(1));
    surf[3] = np_clip(throttle, T(0), T(1));
}

// AttitudeAgent.compute_action.  yaw_cmd_in: NaN = yaw_angle None (-> 0.0, :119); +inf = hold the
// current yaw, which is the command HSAAgent issues every step (hsa_agent.py:214-216).
template <typename T>
__device__ __forceinline__ void attitude_agent(const CtrlParams<T>& C, PidMem (&ang)[3], PidMem (&rate)[3], T roll_cmd_in,
                                               T pitch_cmd_in, T yaw_cmd_in, T throttle, const T (&s)[12], float dt,
                                               T (&surf)[4]) {
    const T roll_cmd = np_clip(roll_cmd_in, -C.max_roll, C.max_roll);            // attitude_agent.py:116-117
    const T pitch_cmd = np_clip(pitch_cmd_in, -C.max_pitch, C.max_pitch);
    const bool hold = (yaw_cmd_in == yaw_cmd_in) && !Math<T>::finite(yaw_cmd_in);
    const T yaw_in = hold ? s[8] : yaw_cmd_in;
    const T yaw_cmd = (yaw_in != yaw_in) ? T(0) : wrap_angle(yaw_in);            // :119
    const T current_yaw = wrap_angle(s[8]);                                      // :127
    const float o_roll = pid_compute(C.angle[0], ang[0], (float)roll_cmd, (float)s[6], dt);     // :134
    const float o_pitch = pid_compute(C.angle[1], ang[1], (float)pitch_cmd, (float)s[7], dt);
    const float o_yaw = pid_compute(C.angle[2], ang[2], (float)yaw_cmd, (float)current_yaw, dt);
    const T pc = np_clip((T)o_roll, -C.max_roll_rate, C.max_roll_rate);          // :137-139
    const T qc = np_clip((T)o_pitch, -C.max_pitch_rate, C.max_pitch_rate);
    const T rc = np_clip((T)o_yaw, -C.max_yaw_rate, C.max_yaw_rate);
    rate_agent(C, rate, pc, qc, rc, throttle, s[9], s[10], s[11], dt, surf);     // :142-152
}

}  // namespace acb
