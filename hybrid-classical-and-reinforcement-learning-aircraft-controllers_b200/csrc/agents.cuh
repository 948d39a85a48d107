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
    if constexpr (!M::kParity) {
        // throughput build: a - 2 pi floor((a + pi) / (2 pi)) with a two-word 2 pi (as Math<float>::wrap_pi); an angle already
        // in [-pi, pi) comes back bit for bit (k = 0), so "hold the current yaw" still gives an error of exactly zero
        const float k = floorf((a + 3.14159274101257324f) * 0.159154943091895346f);
        float r = fmaf(k, -6.28318548202514648f, a);
        r = fmaf(k, 1.74845553146951715e-07f, r);
        return r;
    } else {
        const T two_pi = T(2) * T(3.141592653589793);
        T m = M::fmod(a + T(3.141592653589793), two_pi);
        if (m != T(0)) {
            if (m < T(0)) m += two_pi;
        } else {
            m = T(0);
        }
        return m - T(3.141592653589793);
    }
}
// The reference's arctan2(sin(a), cos(a)) idiom (waypoint_agent.py:132,147,185,189): the parity build evaluates it as
// written, the throughput build as a - 2 pi round(a / 2 pi) (its arguments are sums of two wrapped angles).
template <typename T> __device__ __forceinline__ T wrap_atan(T a) {
    using M = Math<T>;
    if constexpr (M::kParity) return M::atan2(M::sin(a), M::cos(a));
    else return M::wrap_pi(a);
}

// RateAgent.compute_action.  surf = elevator, aileron, rudder, throttle (ControlSurfaces.to_array order)
template <typename T>
__device__ __forceinline__ void rate_agent(const CtrlParams<T>& C, PidMem (&mem)[3], T p_cmd_in, T q_cmd_in, T r_cmd_in,
                                           T throttle, T p, T q, T r, const PidDt& dt, T (&surf)[4]) {
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
                                               T pitch_cmd_in, T yaw_cmd_in, T throttle, const T (&s)[12], const PidDt& dt,
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

// HSAAgent.compute_action (controllers/hsa_agent.py:119-231).  outer = heading, energy, balance PID memory.
// d = airspeed, altitude, ground_speed, heading of the current state (what get_state() hands the agent).
template <typename T>
__device__ __forceinline__ void hsa_agent(const CtrlParams<T>& C, PidMem (&outer)[3], PidMem (&ang)[3], PidMem (&rate)[3],
                                          T heading_cmd, T speed_cmd, T altitude_cmd, const T (&s)[12], const T (&d)[4], const PidDt& dt,
                                          T (&surf)[4]) {
    using M = Math<T>;
    // :147-160 heading error wrapped, fed to the PID as a virtual setpoint
    const T heading_error = wrap_angle(heading_cmd - d[3]);
    const T virtual_setpoint = d[3] + heading_error;
    T roll_angle = (T)pid_compute(C.heading, outer[0], (float)virtual_setpoint, (float)d[3], dt);
    roll_angle = np_clip(roll_angle, -C.max_bank, C.max_bank);
    // :167-186 TECS specific energies
    const T g = T(9.81), h = d[1], V = d[0];
    const T half_v2 = T(0.5) * M::sq(V), half_vc2 = T(0.5) * M::sq(speed_cmd);
    const T e_specific = M::mad(g, h, half_v2), e_specific_cmd = M::mad(g, altitude_cmd, half_vc2);
    const T e_balance = M::mad(g, h, -half_v2), e_balance_cmd = M::mad(g, altitude_cmd, -half_vc2);
    // :192-197 total energy -> throttle
    const T throttle_adjustment = (T)pid_compute(C.energy, outer[1], (float)e_specific_cmd, (float)e_specific, dt);
    const T throttle = np_clip(C.baseline_throttle + throttle_adjustment, T(0), T(1));
    // :203-213 energy balance -> pitch, load-factor compensation in the turn
    T pitch_angle = (T)pid_compute(C.balance, outer[2], (float)e_balance_cmd, (float)e_balance, dt);
    const T cos_roll = M::cos_bounded(roll_angle);
    if (M::abs(cos_roll) > T(0.01)) {
        const T load_factor = M::fdiv(T(1), cos_roll);
        pitch_angle = M::mad(C.load_factor_gain, load_factor - T(1), pitch_angle);
    }
    pitch_angle = np_clip(pitch_angle, -C.max_pitch_cmd, C.max_pitch_cmd);
    // :216-229 yaw command = current yaw
    attitude_agent(C, ang, rate, roll_angle, pitch_angle, s[8], throttle, s, dt, surf);
}

// WaypointAgent.compute_action (controllers/waypoint_agent.py:81-242) up to the HSA command it issues.
// wp = north, east, down, speed (NaN = None).  hsa = heading, speed, altitude.
template <typename T>
__device__ __forceinline__ void waypoint_guidance(const CtrlParams<T>& C, int guidance, const T (&wp)[4], const T (&s)[12],
                                                  const T (&d)[4], T (&hsa)[3]) {
    using M = Math<T>;
    const T e0 = wp[0] - s[0], e1 = wp[1] - s[1];                        // :244-253
    const T horizontal_distance = M::norm2(e0, e1);                     // :117
    const T airspeed = d[0], heading = d[3];
    const T V = py_max(airspeed, T(10));
    T heading_cmd;
    if (guidance == ACB200_GUIDANCE_LOS) {                               // :121-146
        heading_cmd = M::atan2(e1, e0);
        const T turn_radius = M::kParity ? M::div(M::sq(V), T(9.81) * M::tan(C.los_max_bank)) : M::sq(V) * C.los_inv_g_tan;
        T heading_error = heading_cmd - heading;
        heading_error = wrap_atan(heading_error);
        const T anticipation_dist = M::fdiv(turn_radius * M::abs(heading_error), C.rad90);
        if (horizontal_distance < anticipation_dist && M::abs(heading_error) > C.rad20) {
            const T blend = T(1) - M::fdiv(horizontal_distance, anticipation_dist);
            heading_cmd = M::mad(blend * C.los_lead_angle, np_sign(heading_error), heading_cmd);
            heading_cmd = wrap_atan(heading_cmd);
        }
    } else if (!M::kParity) {
        // pure pursuit, throughput build: the look-ahead point lies on the line to the waypoint (:170-182 scale the unit vector
        // by a positive distance), so its bearing IS the bearing of the waypoint; look-ahead, proximity scaling and turn radius
        // cancel out of atan2 and are not evaluated (the parity build below runs them as written: 1e-7 rad of rounding)
        heading_cmd = M::atan2(e1, e0);
    } else {                                                             // pure pursuit :148-182
        const T turn_radius = M::kParity ? M::div(M::sq(V), T(9.81) * M::tan(C.pp_max_bank)) : M::sq(V) * C.pp_inv_g_tan;
        T lookahead_dist = C.lookahead_time * V;
        const T proximity_dist = C.proximity_scale_distance * turn_radius;
        if (horizontal_distance < proximity_dist) {
            const T scale = M::mad(T(0.4), M::fdiv(horizontal_distance, proximity_dist), T(0.6));
            lookahead_dist = lookahead_dist * scale;
        }
        lookahead_dist = np_clip(lookahead_dist, C.lookahead_min, C.lookahead_max);
        if (horizontal_distance > T(1)) {
            const T dir0 = M::fdiv(e0, horizontal_distance), dir1 = M::fdiv(e1, horizontal_distance);
            const T effective = (horizontal_distance < lookahead_dist) ? horizontal_distance : lookahead_dist;
            heading_cmd = M::atan2(dir1 * effective, dir0 * effective);
        } else {
            heading_cmd = M::atan2(e1, e0);
        }
    }
    heading_cmd = wrap_atan(heading_cmd);  // :189
    // :207-232 altitude and speed commands, slowing down into sharp turns
    T speed_cmd = (wp[3] != wp[3]) ? airspeed : wp[3];
    const T heading_to_wp = M::atan2(e1, e0);
    T heading_error = M::abs(heading_to_wp - heading);
    heading_error = M::abs(wrap_atan(heading_error));
    if (horizontal_distance < C.turn_threshold_distance && heading_error > C.turn_threshold_angle) {
        // speed_cmd * (1 - reduction), reduction = max_speed_reduction * (1 - d / threshold)  (:224-227)
        const T keep = M::mad(-C.max_speed_reduction, T(1) - M::fdiv(horizontal_distance, C.turn_threshold_distance), T(1));
        speed_cmd = speed_cmd * keep;
        speed_cmd = py_max(speed_cmd, C.min_speed);
    }
    hsa[0] = heading_cmd; hsa[1] = speed_cmd; hsa[2] = -wp[2];
}

}  // namespace acb
