// params.cuh -- launch-constant parameter blocks, converted once on the host into the arithmetic
// type of the build and passed to the kernels as __grid_constant__ arguments (constant bank).
#pragma once
#include "../../include/acb200.h"
#include <math.h>

#include "math.cuh"

namespace acb {

// Derived from acb200_aircraft_params exactly the way Simplified6DOF.__init__ / _dynamics derive
// them (simplified_6dof.py:181-187, 379, 407-409, 455, 474-481): each field is the result of the
// same double operation the reference performs, so hoisting it out of the loop changes no bit.
template <typename T> struct DynParams {
    T mass, inv_mass, g, neg_g, half_rho;
    T wing_area, wing_span, chord;
    T cl_0, cl_alpha, cd_0, cd_alpha2;
    T cl_elevator, cm_elevator, cy_rudder, cn_rudder, cl_aileron;
    T cm_alpha, cn_beta, cl_beta;
    T damping_roll, damping_pitch, damping_yaw;
    T max_thrust, thrust_zero_velocity;
    T min_airspeed_aero, min_u_velocity;
    T max_elevator_rad, max_aileron_rad, max_rudder_rad;
    T max_velocity, max_rate_rad, max_pitch_rad, max_alpha_rad;
    T max_acceleration, max_angular_acceleration;
    T ixx, iyy, izz, izz_m_iyy, ixx_m_izz, iyy_m_ixx;
    // ---- products and reciprocals folded on the host in double, used by the throughput build only
    // (physics_f32.cuh); the parity build keeps the reference's operation order and ignores them ----
    T weight;                                    // mass * g
    T half_rho_S;                                // 0.5 rho S
    T cl_elev_rad, cm_elev_rad, cy_rud_rad, cn_rud_rad, cl_ail_rad;   // coefficient * max deflection [rad]
    T damp_p, damp_q, damp_r;                    // damping * (b or c) / 2
    T inv_vt0, inv_min_airspeed;
    T tan_max_alpha, sin_max_pitch, cos_max_pitch;
    // moments / inertia with 0.5 rho S (b or c) / I folded into every coefficient: L/Ixx = V^2 (ql_beta beta + ql_damp p / V + ql_ail aileron)
    T half_rho_S_m;                              // 0.5 rho S / m
    T ql_beta, ql_damp, ql_ail;                  // 0.5 rho S b / Ixx * {cl_beta, damping_roll b / 2, cl_aileron max_aileron}
    T qm_alpha, qm_damp, qm_elev;                // 0.5 rho S c / Iyy * {cm_alpha, damping_pitch c / 2, cm_elevator max_elevator}
    T qn_beta, qn_damp, qn_rud;                  // 0.5 rho S b / Izz * {cn_beta, damping_yaw b / 2, cn_rudder max_rudder}
    T rk_two, rk_one;                            // RK4 weights 2 and 1 as run-time values (physics_f32x2.cuh)
    T euler_rate_gain;                           // bound of |d(Euler angle)/dt| / max|body rate|: 1 + 2 / cos(max_pitch), +1 %
    T inv_ixx, inv_iyy, inv_izz;
    T b_ixx, c_iyy, b_izz;                       // span / Ixx, chord / Iyy, span / Izz
    T gyro_x, gyro_y, gyro_z;                    // (Izz-Iyy)/Ixx, (Ixx-Izz)/Iyy, (Iyy-Ixx)/Izz
};

template <typename T> struct PidCfg3 { acb200_pid_config axis[3]; };

template <typename T> struct CtrlParams {
    acb200_pid_config rate[3];
    acb200_pid_config angle[3];
    T max_roll_rate, max_pitch_rate, max_yaw_rate, max_roll, max_pitch;
    // HSA / waypoint levels (hsa_agent.py:38-108, waypoint_agent.py:33-74); angles pre-converted with np.radians
    acb200_pid_config heading, energy, balance;
    T max_bank, baseline_throttle, load_factor_gain, max_pitch_cmd;
    T pp_max_bank, los_max_bank, los_lead_angle, turn_threshold_angle, rad90, rad20;
    T lookahead_time, lookahead_min, lookahead_max, proximity_scale_distance;
    T turn_threshold_distance, max_speed_reduction, min_speed;
    T pp_inv_g_tan, los_inv_g_tan;      // 1 / (9.81 tan(max bank)): the turn radius per V^2 (throughput build; the parity build divides)
};

static inline double host_radians(double deg) { return deg * (3.141592653589793 / 180.0); }  // np.radians

template <typename T> static inline DynParams<T> make_dyn_params(const acb200_aircraft_params& a) {
    DynParams<T> d;
    d.mass = (T)a.mass; d.inv_mass = (T)(1.0 / a.mass); d.g = (T)a.gravity; d.neg_g = (T)(-a.gravity);
    d.half_rho = (T)(0.5 * a.air_density);
    d.wing_area = (T)a.wing_area; d.wing_span = (T)a.wing_span; d.chord = (T)a.chord;
    d.cl_0 = (T)a.cl_0; d.cl_alpha = (T)a.cl_alpha; d.cd_0 = (T)a.cd_0; d.cd_alpha2 = (T)a.cd_alpha2;
    d.cl_elevator = (T)a.cl_elevator; d.cm_elevator = (T)a.cm_elevator; d.cy_rudder = (T)a.cy_rudder;
    d.cn_rudder = (T)a.cn_rudder; d.cl_aileron = (T)a.cl_aileron;
    d.cm_alpha = (T)a.cm_alpha; d.cn_beta = (T)a.cn_beta; d.cl_beta = (T)a.cl_beta;
    d.damping_roll = (T)a.damping_roll; d.damping_pitch = (T)a.damping_pitch; d.damping_yaw = (T)a.damping_yaw;
    d.max_thrust = (T)a.max_thrust; d.thrust_zero_velocity = (T)a.thrust_zero_velocity;
    d.min_airspeed_aero = (T)a.min_airspeed_aero; d.min_u_velocity = (T)a.min_u_velocity;
    d.max_elevator_rad = (T)host_radians(a.max_elevator_deflection);
    d.max_aileron_rad = (T)host_radians(a.max_aileron_deflection);
    d.max_rudder_rad = (T)host_radians(a.max_rudder_deflection);
    d.max_velocity = (T)a.max_velocity; d.max_rate_rad = (T)host_radians(a.max_angular_rate);
    d.max_pitch_rad = (T)host_radians(a.max_pitch_angle); d.max_alpha_rad = (T)host_radians(a.max_alpha);
    d.max_acceleration = (T)a.max_acceleration; d.max_angular_acceleration = (T)a.max_angular_acceleration;
    d.ixx = (T)a.inertia_xx; d.iyy = (T)a.inertia_yy; d.izz = (T)a.inertia_zz;
    d.izz_m_iyy = (T)(a.inertia_zz - a.inertia_yy); d.ixx_m_izz = (T)(a.inertia_xx - a.inertia_zz);
    d.iyy_m_ixx = (T)(a.inertia_yy - a.inertia_xx);
    d.weight = (T)(a.mass * a.gravity);
    d.half_rho_S = (T)(0.5 * a.air_density * a.wing_area);
    d.cl_elev_rad = (T)(a.cl_elevator * host_radians(a.max_elevator_deflection));
    d.cm_elev_rad = (T)(a.cm_elevator * host_radians(a.max_elevator_deflection));
    d.cy_rud_rad = (T)(a.cy_rudder * host_radians(a.max_rudder_deflection));
    d.cn_rud_rad = (T)(a.cn_rudder * host_radians(a.max_rudder_deflection));
    d.cl_ail_rad = (T)(a.cl_aileron * host_radians(a.max_aileron_deflection));
    d.damp_p = (T)(a.damping_roll * a.wing_span * 0.5);
    d.damp_q = (T)(a.damping_pitch * a.chord * 0.5);
    d.damp_r = (T)(a.damping_yaw * a.wing_span * 0.5);
    d.inv_vt0 = (T)(1.0 / a.thrust_zero_velocity);
    d.inv_min_airspeed = (T)(1.0 / a.min_airspeed_aero);
    d.tan_max_alpha = (T)tan(host_radians(a.max_alpha));
    d.sin_max_pitch = (T)sin(host_radians(a.max_pitch_angle));
    d.cos_max_pitch = (T)cos(host_radians(a.max_pitch_angle));
    {
        const double hs = 0.5 * a.air_density * a.wing_area;
        const double ql = hs * a.wing_span / a.inertia_xx, qm = hs * a.chord / a.inertia_yy, qn = hs * a.wing_span / a.inertia_zz;
        d.half_rho_S_m = (T)(hs / a.mass);
        d.ql_beta = (T)(ql * a.cl_beta); d.ql_damp = (T)(ql * a.damping_roll * a.wing_span * 0.5);
        d.ql_ail = (T)(ql * a.cl_aileron * host_radians(a.max_aileron_deflection));
        d.qm_alpha = (T)(qm * a.cm_alpha); d.qm_damp = (T)(qm * a.damping_pitch * a.chord * 0.5);
        d.qm_elev = (T)(qm * a.cm_elevator * host_radians(a.max_elevator_deflection));
        d.qn_beta = (T)(qn * a.cn_beta); d.qn_damp = (T)(qn * a.damping_yaw * a.wing_span * 0.5);
        d.qn_rud = (T)(qn * a.cn_rudder * host_radians(a.max_rudder_deflection));
    }
    d.rk_two = T(2); d.rk_one = T(1);
    d.euler_rate_gain = (T)(1.01 * (1.0 + 2.0 / fabs(cos(host_radians(a.max_pitch_angle)))));
    d.inv_ixx = (T)(1.0 / a.inertia_xx); d.inv_iyy = (T)(1.0 / a.inertia_yy); d.inv_izz = (T)(1.0 / a.inertia_zz);
    d.b_ixx = (T)(a.wing_span / a.inertia_xx); d.c_iyy = (T)(a.chord / a.inertia_yy); d.b_izz = (T)(a.wing_span / a.inertia_zz);
    d.gyro_x = (T)((a.inertia_zz - a.inertia_yy) / a.inertia_xx);
    d.gyro_y = (T)((a.inertia_xx - a.inertia_zz) / a.inertia_yy);
    d.gyro_z = (T)((a.inertia_yy - a.inertia_xx) / a.inertia_zz);
    return d;
}

template <typename T> static inline CtrlParams<T> make_ctrl_params(const acb200_controller_params& c) {
    CtrlParams<T> o;
    for (int i = 0; i < 3; ++i) { o.rate[i] = c.rate[i]; o.angle[i] = c.angle[i]; }
    o.max_roll_rate = (T)c.max_roll_rate; o.max_pitch_rate = (T)c.max_pitch_rate; o.max_yaw_rate = (T)c.max_yaw_rate;
    o.max_roll = (T)c.max_roll; o.max_pitch = (T)c.max_pitch;
    o.heading = c.heading; o.energy = c.energy; o.balance = c.balance;
    o.max_bank = (T)c.max_bank; o.baseline_throttle = (T)c.baseline_throttle; o.load_factor_gain = (T)c.load_factor_gain;
    o.max_pitch_cmd = (T)c.max_pitch_cmd;
    o.pp_max_bank = (T)host_radians(c.max_bank_angle_deg); o.los_max_bank = (T)host_radians(c.los_max_bank_deg);
    o.los_lead_angle = (T)host_radians(c.los_lead_angle_deg); o.turn_threshold_angle = (T)host_radians(c.turn_threshold_angle_deg);
    o.rad90 = (T)host_radians(90.0); o.rad20 = (T)host_radians(20.0);
    o.lookahead_time = (T)c.lookahead_time; o.lookahead_min = (T)c.lookahead_min; o.lookahead_max = (T)c.lookahead_max;
    o.proximity_scale_distance = (T)c.proximity_scale_distance; o.turn_threshold_distance = (T)c.turn_threshold_distance;
    o.max_speed_reduction = (T)c.max_speed_reduction; o.min_speed = (T)c.min_speed;
    o.pp_inv_g_tan = (T)(1.0 / (9.81 * tan(host_radians(c.max_bank_angle_deg))));
    o.los_inv_g_tan = (T)(1.0 / (9.81 * tan(host_radians(c.los_max_bank_deg))));
    return o;
}

}  // namespace acb
