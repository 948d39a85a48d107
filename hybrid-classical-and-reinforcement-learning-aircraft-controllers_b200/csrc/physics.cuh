// physics.cuh -- one aircraft's 6-DOF step, state in registers.
//
// Restates Simplified6DOF._dynamics / step / get_state (simulation/simplified_6dof.py:228-331,
// 333-503) for a thread that owns one aircraft.  In the parity build (T = double) the operation
// order is the reference's, op for op; the throughput build (T = float) keeps the same algebra and
// lets nvcc contract it.
#pragma once
#include "math.cuh"
#include "params.cuh"

namespace acb {

// Simplified6DOF._dynamics, simplified_6dof.py:333-503.  c = elevator, aileron, rudder, throttle
// (already clipped by set_controls).  Returns the number of non-finite derivatives it zeroed.
template <typename T>
__device__ __forceinline__ int dynamics(const DynParams<T>& P, const T (&s)[12], const T (&c)[4], T (&out)[12]) {
    using M = Math<T>;
    const T u = s[3], v = s[4], w = s[5];
    const T theta = s[7];
    const T p = s[9], q = s[10], r = s[11];

    // :353-358
    T sin_phi, cos_phi, sin_theta, cos_theta, sin_psi, cos_psi;
    M::sincos(s[6], &sin_phi, &cos_phi);
    M::sincos(theta, &sin_theta, &cos_theta);
    M::sincos(s[8], &sin_psi, &cos_psi);

    // :363-364
    const T airspeed = M::norm3(u, v, w);
    const T safe_airspeed = py_max(airspeed, P.min_airspeed_aero);

    // :367-373
    const T abs_u = M::abs(u);
    const T u_safe = (abs_u > T(1e-6)) ? py_max(abs_u, P.min_u_velocity) * np_sign(u) : P.min_u_velocity;
    const T alpha_raw = M::atan2(w, u_safe);
    const T alpha = np_clip(alpha_raw, -P.max_alpha_rad, P.max_alpha_rad);
    T sin_alpha, cos_alpha;
    M::alpha_sincos(alpha, w, u_safe, (alpha == alpha_raw) && (u_safe > T(0)), &sin_alpha, &cos_alpha);

    // :376-379
    const T beta = M::asin(np_clip(M::div(v, safe_airspeed), T(-1), T(1)));
    const T q_dyn = P.half_rho * M::sq(airspeed);

    // :383-390
    const T elevator_rad = c[0] * P.max_elevator_rad;
    const T cl = P.cl_0 + P.cl_alpha * alpha + P.cl_elevator * elevator_rad;
    const T cd = P.cd_0 + P.cd_alpha2 * M::sq(alpha);
    const T rudder_rad = c[2] * P.max_rudder_rad;
    const T cy = P.cy_rudder * rudder_rad;

    // :392-399
    const T q_S = q_dyn * P.wing_area;
    const T lift = q_S * cl, drag = q_S * cd, side_force = q_S * cy;
    const T fx_aero = -drag * cos_alpha + lift * sin_alpha;
    const T fz_aero = -drag * sin_alpha - lift * cos_alpha;

    // :402-404
    const T thrust_factor = py_max(T(0), T(1) - M::div(airspeed, P.thrust_zero_velocity));
    const T thrust = P.max_thrust * c[3] * thrust_factor;

    // :407-411
    const T fx = fx_aero + thrust + (P.neg_g * sin_theta) * P.mass;
    const T fy = side_force + (P.g * cos_theta * sin_phi) * P.mass;
    const T fz = fz_aero + (P.g * cos_theta * cos_phi) * P.mass;

    // :415-435
    const T aileron_rad = c[1] * P.max_aileron_rad;
    const T two_v = T(2) * safe_airspeed;
    const T half_span_over_V = M::div(P.wing_span, two_v);
    const T half_chord_over_V = M::div(P.chord, two_v);
    const T l_moment = q_S * P.wing_span
                       * (P.cl_aileron * aileron_rad + P.damping_roll * p * half_span_over_V + P.cl_beta * beta);
    const T m_moment = q_S * P.chord
                       * (P.cm_elevator * elevator_rad + P.cm_alpha * alpha + P.damping_pitch * q * half_chord_over_V);
    const T n_moment = q_S * P.wing_span
                       * (P.cn_rudder * rudder_rad + P.damping_yaw * r * half_span_over_V + P.cn_beta * beta);

    // :440-452
    const T sin_phi_sin_theta = sin_phi * sin_theta;
    const T cos_phi_sin_theta = cos_phi * sin_theta;
    out[0] = cos_theta * cos_psi * u + (sin_phi_sin_theta * cos_psi - cos_phi * sin_psi) * v
             + (cos_phi_sin_theta * cos_psi + sin_phi * sin_psi) * w;
    out[1] = cos_theta * sin_psi * u + (sin_phi_sin_theta * sin_psi + cos_phi * cos_psi) * v
             + (cos_phi_sin_theta * sin_psi - sin_phi * cos_psi) * w;
    out[2] = -sin_theta * u + sin_phi * cos_theta * v + cos_phi * cos_theta * w;

    // :455-460, clamp :488-490
    const T vd0 = fx * P.inv_mass - q * w + r * v;
    const T vd1 = fy * P.inv_mass - r * u + p * w;
    const T vd2 = fz * P.inv_mass - p * v + q * u;
    out[3] = np_clip(vd0, -P.max_acceleration, P.max_acceleration);
    out[4] = np_clip(vd1, -P.max_acceleration, P.max_acceleration);
    out[5] = np_clip(vd2, -P.max_acceleration, P.max_acceleration);

    // :463-471
    const T theta_safe = np_clip(theta, -P.max_pitch_rad, P.max_pitch_rad);
    T tan_theta_safe, cos_theta_safe;
    M::tan_cos_safe(theta_safe, sin_theta, cos_theta, theta_safe != theta, &tan_theta_safe, &cos_theta_safe);
    out[6] = p + sin_phi * tan_theta_safe * q + cos_phi * tan_theta_safe * r;
    out[7] = cos_phi * q - sin_phi * r;
    out[8] = M::div(sin_phi * q + cos_phi * r, cos_theta_safe);

    // :474-487
    const T rd0 = M::div(l_moment - P.izz_m_iyy * q * r, P.ixx);
    const T rd1 = M::div(m_moment - P.ixx_m_izz * p * r, P.iyy);
    const T rd2 = M::div(n_moment - P.iyy_m_ixx * p * q, P.izz);
    out[9] = np_clip(rd0, -P.max_angular_acceleration, P.max_angular_acceleration);
    out[10] = np_clip(rd1, -P.max_angular_acceleration, P.max_angular_acceleration);
    out[11] = np_clip(rd2, -P.max_angular_acceleration, P.max_angular_acceleration);

    // :496-501
    int repaired = 0;
#pragma unroll
    for (int i = 0; i < 12; ++i)
        if (!M::finite(out[i])) { out[i] = T(0); ++repaired; }
    return repaired;
}

// Simplified6DOF.step, simplified_6dof.py:228-291: classic RK4 then the state clamps.
// The RK4 sum is accumulated as ((k1 + 2 k2) + 2 k3) + k4, numpy's left-to-right order (:253).
template <typename T>
__device__ __forceinline__ int rk4_step(const DynParams<T>& P, T (&s)[12], T& time, const T (&c)[4], T dt) {
    using M = Math<T>;
    T k[12], acc[12], t[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) { k[i] = T(0); acc[i] = T(0); }
    int repaired = 0;
    const T half_dt = T(0.5) * dt;
    // stage loop kept rolled: one copy of the derivative code in the instruction stream
#pragma unroll 1
    for (int stage = 0; stage < 4; ++stage) {
        const T a = (stage == 3) ? dt : half_dt;
#pragma unroll
        for (int i = 0; i < 12; ++i) t[i] = (stage == 0) ? s[i] : s[i] + a * k[i];
        repaired += dynamics(P, t, c, k);
        const T wgt = (stage == 1 || stage == 2) ? T(2) : T(1);
#pragma unroll
        for (int i = 0; i < 12; ++i) acc[i] = (stage == 0) ? k[i] : acc[i] + wgt * k[i];
    }
    const T dt6 = M::div(dt, T(6));
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = s[i] + dt6 * acc[i];
    time += dt;

    // :258  position: nan -> 0, +-inf -> +-1e4
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const T x = s[i];
        s[i] = (x != x) ? T(0) : (M::finite(x) ? x : (x > T(0) ? T(10000) : T(-10000)));
    }
    // :262
#pragma unroll
    for (int i = 3; i < 6; ++i) s[i] = np_clip(s[i], -P.max_velocity, P.max_velocity);
    // :266-270
    s[6] = M::wrap_pi(s[6]);
    s[7] = np_clip(s[7], -P.max_pitch_rad, P.max_pitch_rad);
    s[8] = M::wrap_pi(s[8]);
    // :273
#pragma unroll
    for (int i = 9; i < 12; ++i) s[i] = np_clip(s[i], -P.max_rate_rad, P.max_rate_rad);
    // :276-283  ground: altitude = -z < 0
    if (-s[2] < T(0)) {
        s[2] = T(0);
        s[5] = py_max(T(0), s[5]);
    }
    // :286-291
#pragma unroll
    for (int i = 0; i < 12; ++i)
        if (!M::finite(s[i])) { s[i] = T(0); ++repaired; }
    return repaired;
}

// Simplified6DOF.get_state + _body_to_ned, simplified_6dof.py:295-331,505-530:
// d = airspeed, altitude, ground_speed, heading.
// (Throughput build: the general Cody-Waite sincos stays here on purpose.  The SFU's sin/cos (abs err 5e-7) and the short-range
// polynomial pairs (1.5e-7) were both tried for these three angles and cost the ground speed / heading of a steeply climbing
// aircraft -- ground speed << airspeed amplifies them -- more than the 1e-5 / 2e-5 bars of tests/test_gpu_parity_f32.py.)
template <typename T> __device__ __forceinline__ void derived_outputs(const T (&s)[12], T (&d)[4]) {
    using M = Math<T>;
    const T u = s[3], v = s[4], w = s[5];
    T sphi, cphi, sth, cth, sps, cps;
    M::sincos(s[6], &sphi, &cphi);
    M::sincos(s[7], &sth, &cth);
    M::sincos(s[8], &sps, &cps);
    const T r00 = cth * cps, r01 = M::mad(sphi * sth, cps, -(cphi * sps)), r02 = M::mad(cphi * sth, cps, sphi * sps);
    const T r10 = cth * sps, r11 = M::mad(sphi * sth, sps, cphi * cps), r12 = M::mad(cphi * sth, sps, -(sphi * cps));
    const T vn = M::row_dot3(r00, r01, r02, u, v, w);
    const T ve = M::row_dot3(r10, r11, r12, u, v, w);
    d[0] = M::norm3(u, v, w);
    d[1] = -s[2];
    d[2] = M::norm2(vn, ve);
    d[3] = M::atan2(ve, vn);
}

// set_controls, simplified_6dof.py:221-226
template <typename T> __device__ __forceinline__ void clip_controls(const T (&in)[4], T (&out)[4]) {
    out[0] = np_clip(in[0], T(-1), T(1));
    out[1] = np_clip(in[1], T(-1), T(1));
    out[2] = np_clip(in[2], T(-1), T(1));
    out[3] = np_clip(in[3], T(0), T(1));
}

// SimulationAircraftBackend.step, simulation_backend.py:95-99: K substeps of h under zero-order hold
template <typename T>
__device__ __forceinline__ int backend_step(const DynParams<T>& P, T (&s)[12], T& time, const T (&c)[4], int K, T h) {
    int repaired = 0;
#pragma unroll 1
    for (int k = 0; k < K; ++k) repaired += rk4_step(P, s, time, c, h);
    return repaired;
}

}  // namespace acb

#include "physics_f32.cuh"   // the throughput build's backend_step(float) overload
