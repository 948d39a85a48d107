/*
 * acb_oracle.c -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY
 * (see acb_oracle.h for the rules on who may load this).  Parity status: PINNED against
 * tests/golden/ (generated from the unmodified reference by tests/golden/make_golden.py).
 *
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off -fno-builtin -fopenmp; no -march => no FMA,
 * like the reference's own -O3 build of its PID, CMakeLists.txt:15-16).
 */
#include "acb_oracle.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_PI 3.141592653589793

/* ------------------------------------------------------------------------------------------
 * numpy / Python scalar semantics used by the reference
 * ---------------------------------------------------------------------------------------- */

/* np.clip(x, lo, hi) on a scalar == minimum(maximum(x, lo), hi); NaN propagates. */
static inline double np_clip(double x, double lo, double hi) {
    if (x != x) return x;
    if (x < lo) x = lo;
    if (x > hi) x = hi;
    return x;
}
/* Python builtin max(a, b): a unless b > a (so a NaN first argument is kept). */
static inline double py_max(double a, double b) { return (b > a) ? b : a; }
/* np.sign */
static inline double np_sign(double x) { return (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : ((x != x) ? x : 0.0)); }
/* np.radians(x) == x * (pi/180) */
static inline double np_radians(double deg) { return deg * (ORC_PI / 180.0); }
/* numpy scalar x**2 goes through libm pow (probed in this container: identical to pow(x,2.0) on
 * 100k samples, differs from x*x on 0.09 %).  -fno-builtin keeps gcc from folding it to x*x. */
static inline double np_sq(double x) { return pow(x, 2.0); }
/* np.linalg.norm of a contiguous 3-vector = sqrt(ddot(x,x)); OpenBLAS's ddot tail loop is fused:
 * probed identical to sqrt(fma(w,w,fma(v,v,u*u))) on 100k samples (plain sum differs on 11 %). */
static inline double np_norm3(double u, double v, double w) { return sqrt(fma(w, w, fma(v, v, u * u))); }
static inline double np_norm2(double a, double b) { return sqrt(fma(b, b, a * a)); }
/* one row of (3x3 ndarray) @ (3-vector) through OpenBLAS dgemv: probed identical to
 * fma(r2,x2, fma(r0,x0, r1*x1)) on 20k samples for every row. */
static inline double np_row_dot3(double r0, double r1, double r2, double x0, double x1, double x2) {
    return fma(r2, x2, fma(r0, x0, r1 * x1));
}
/* np.nan_to_num(x, nan=0, posinf=a, neginf=b) */
static inline double np_nan_to_num(double x, double pinf, double ninf) {
    if (x != x) return 0.0;
    if (isinf(x)) return x > 0 ? pinf : ninf;
    return x;
}
/* Python/numpy floored modulo for floats (npy_divmod): fmod, then move into the divisor's sign. */
static inline double np_mod(double a, double b) {
    double m = fmod(a, b);
    if (m != 0.0) {
        if ((b < 0.0) != (m < 0.0)) m += b;
    } else {
        m = copysign(0.0, b);
    }
    return m;
}

/* controllers/utils/pid_utils.py:49-65 */
double orc_wrap_angle(double a) { return np_mod(a + ORC_PI, 2 * ORC_PI) - ORC_PI; }

/* ------------------------------------------------------------------------------------------
 * parameters
 * ---------------------------------------------------------------------------------------- */

/* simulation/simplified_6dof.py:45-117 defaults */
void orc_default_aircraft_params(orc_aircraft_params* p) {
    p->mass = 8.0; p->inertia_xx = 0.4; p->inertia_yy = 0.6; p->inertia_zz = 0.7;
    p->wing_area = 0.5; p->wing_span = 2.0; p->chord = 0.25;
    p->cl_0 = 0.4; p->cl_alpha = 5.0; p->cd_0 = 0.025; p->cd_alpha2 = 0.04;
    p->cl_elevator = 0.6; p->cm_elevator = -2.0; p->cy_rudder = 0.5; p->cn_rudder = -0.40; p->cl_aileron = 0.50;
    p->cm_alpha = -0.15; p->cn_beta = 0.04; p->cl_beta = -0.02;
    p->damping_roll = -0.8; p->damping_pitch = -2.0; p->damping_yaw = -0.6;
    p->max_thrust = 50.0; p->air_density = 1.225; p->gravity = 9.81;
    p->min_airspeed_aero = 10.0; p->min_u_velocity = 0.1;
    p->max_elevator_deflection = 30.0; p->max_aileron_deflection = 30.0; p->max_rudder_deflection = 30.0;
    p->thrust_zero_velocity = 50.0;
    p->max_velocity = 100.0; p->max_angular_rate = 360.0; p->max_pitch_angle = 85.0; p->max_alpha = 30.0;
    p->max_acceleration = 50.0; p->max_angular_acceleration = 1000.0;
    p->max_timestep = 1.0; p->min_timestep = 1e-6;
}

/* controllers/utils/pid_utils.py:16-46 (create_pid_config): doubles narrowed to float by pybind11 */
static orc_pid_config make_pid_config(double kp, double ki, double kd, double i_limit, double omin, double omax) {
    orc_pid_config c;
    c.kp = (float)kp; c.ki = (float)ki; c.kd = (float)kd;
    c.integral_min = (float)(-i_limit); c.integral_max = (float)i_limit;
    c.output_min = (float)omin; c.output_max = (float)omax;
    c.derivative_filter_alpha = 0.1f; /* cpp/include/pid_controller.h:48 */
    return c;
}

/* controllers/types.py:294-350 defaults, rate_agent.py:41-55, attitude_agent.py:43-76,
 * hsa_agent.py:38-108 with config_loader.py:40-75 defaults */
void orc_default_controller_params(orc_controller_params* c) {
    const double il = 25.0; /* types.py:291 */
    c->max_roll_rate = np_radians(180.0); c->max_pitch_rate = np_radians(180.0); c->max_yaw_rate = np_radians(160.0);
    c->max_roll = np_radians(30.0); c->max_pitch = np_radians(30.0);
    c->rate[0] = make_pid_config(1.3, 0.4, 0.012, il, -1.0, 1.0);
    c->rate[1] = make_pid_config(0.6, 0.2, 0.008, il, -1.0, 1.0);
    c->rate[2] = make_pid_config(1.6, 0.15, 0.01, il, -1.0, 1.0);
    c->angle[0] = make_pid_config(8.0, 2.0, 0.3, il, -c->max_roll_rate, c->max_roll_rate);
    c->angle[1] = make_pid_config(6.0, 1.5, 0.2, il, -c->max_pitch_rate, c->max_pitch_rate);
    c->angle[2] = make_pid_config(1.6, 0.15, 0.01, il, -c->max_yaw_rate, c->max_yaw_rate);
    /* HSAConfig / TECSConfig defaults (config_loader.py:40-75) */
    c->max_bank_rad = np_radians(25.0);
    c->heading = make_pid_config(1.0, 0.05, 0.2, il, -c->max_bank_rad, c->max_bank_rad);
    c->energy = make_pid_config(0.12, 0.03, 0.03, 10.0, -0.5, 0.5);                                  /* hsa_agent.py:74-84 */
    c->balance = make_pid_config(0.06, 0.005, 0.03, 5.0, -np_radians(15.0), np_radians(15.0));      /* hsa_agent.py:86-100 */
    c->baseline_throttle = 0.1; c->load_factor_gain = 0.05; c->max_pitch_cmd_rad = np_radians(10.0); /* :104-110 */
    /* GuidanceConfig defaults, config_loader.py:40-64 */
    c->max_bank_angle_deg = 25.0;
    c->lookahead_time = 1.2; c->lookahead_min = 12.0; c->lookahead_max = 30.0; c->proximity_scale_distance = 2.0;
    c->los_max_bank_deg = 20.0; c->los_lead_angle_deg = 30.0;
    c->turn_threshold_angle_deg = 60.0; c->turn_threshold_distance = 60.0; c->max_speed_reduction = 0.15; c->min_speed = 14.0;
}

/* ------------------------------------------------------------------------------------------
 * physics
 * ---------------------------------------------------------------------------------------- */

/* Simplified6DOF._dynamics, simulation/simplified_6dof.py:333-503.  ctrl = elevator, aileron,
 * rudder, throttle (already clipped by set_controls). */
void orc_dynamics(const orc_aircraft_params* P, const double s[12], const double ctrl[4], double out[12]) {
    const double u = s[3], v = s[4], w = s[5];
    const double phi = s[6], theta = s[7], psi = s[8];
    const double p = s[9], q = s[10], r = s[11];
    const double elevator = ctrl[0], aileron = ctrl[1], rudder = ctrl[2], throttle = ctrl[3];

    /* :353-358 */
    const double sin_phi = sin(phi), cos_phi = cos(phi);
    const double sin_theta = sin(theta), cos_theta = cos(theta);
    const double sin_psi = sin(psi), cos_psi = cos(psi);

    /* :363-364 */
    const double airspeed = np_norm3(u, v, w);
    const double safe_airspeed = py_max(airspeed, P->min_airspeed_aero);

    /* :367-370 */
    const double min_u = P->min_u_velocity;
    const double u_safe = (fabs(u) > 1e-6) ? py_max(fabs(u), min_u) * np_sign(u) : min_u;
    const double max_alpha = np_radians(P->max_alpha);
    double alpha = atan2(w, u_safe);
    alpha = np_clip(alpha, -max_alpha, max_alpha);
    const double sin_alpha = sin(alpha), cos_alpha = cos(alpha);

    /* :376 */
    const double beta = asin(np_clip(v / safe_airspeed, -1.0, 1.0));
    /* :379 */
    const double q_dyn = 0.5 * P->air_density * np_sq(airspeed);

    /* :383-390 */
    const double elevator_rad = elevator * np_radians(P->max_elevator_deflection);
    const double cl = P->cl_0 + P->cl_alpha * alpha + P->cl_elevator * elevator_rad;
    const double cd = P->cd_0 + P->cd_alpha2 * np_sq(alpha);
    const double rudder_rad = rudder * np_radians(P->max_rudder_deflection);
    const double cy = P->cy_rudder * rudder_rad;

    /* :392-399 */
    const double q_S = q_dyn * P->wing_area;
    const double lift = q_S * cl, drag = q_S * cd, side_force = q_S * cy;
    const double fx_aero = -drag * cos_alpha + lift * sin_alpha;
    const double fy_aero = side_force;
    const double fz_aero = -drag * sin_alpha - lift * cos_alpha;

    /* :402-404 */
    const double thrust_factor = py_max(0.0, 1.0 - airspeed / P->thrust_zero_velocity);
    const double thrust = P->max_thrust * throttle * thrust_factor;

    /* :407-411 */
    const double g = P->gravity, mass = P->mass;
    const double fx = fx_aero + thrust + (-g * sin_theta) * mass;
    const double fy = fy_aero + (g * cos_theta * sin_phi) * mass;
    const double fz = fz_aero + (g * cos_theta * cos_phi) * mass;

    /* :415-435 */
    const double aileron_rad = aileron * np_radians(P->max_aileron_deflection);
    const double half_span_over_V = P->wing_span / (2 * safe_airspeed);
    const double half_chord_over_V = P->chord / (2 * safe_airspeed);
    const double l_moment = q_S * P->wing_span * (P->cl_aileron * aileron_rad
                                                  + P->damping_roll * p * half_span_over_V
                                                  + P->cl_beta * beta);
    const double m_moment = q_S * P->chord * (P->cm_elevator * elevator_rad
                                              + P->cm_alpha * alpha
                                              + P->damping_pitch * q * half_chord_over_V);
    const double n_moment = q_S * P->wing_span * (P->cn_rudder * rudder_rad
                                                  + P->damping_yaw * r * half_span_over_V
                                                  + P->cn_beta * beta);

    /* :440-452 */
    const double sin_phi_sin_theta = sin_phi * sin_theta;
    const double cos_phi_sin_theta = cos_phi * sin_theta;
    out[0] = cos_theta * cos_psi * u
             + (sin_phi_sin_theta * cos_psi - cos_phi * sin_psi) * v
             + (cos_phi_sin_theta * cos_psi + sin_phi * sin_psi) * w;
    out[1] = cos_theta * sin_psi * u
             + (sin_phi_sin_theta * sin_psi + cos_phi * cos_psi) * v
             + (cos_phi_sin_theta * sin_psi - sin_phi * cos_psi) * w;
    out[2] = -sin_theta * u + sin_phi * cos_theta * v + cos_phi * cos_theta * w;

    /* :455-460 */
    const double inv_mass = 1.0 / mass;
    double vd0 = fx * inv_mass - q * w + r * v;
    double vd1 = fy * inv_mass - r * u + p * w;
    double vd2 = fz * inv_mass - p * v + q * u;

    /* :463-471 */
    const double max_pitch = np_radians(P->max_pitch_angle);
    const double theta_safe = np_clip(theta, -max_pitch, max_pitch);
    const double cos_theta_safe = cos(theta_safe);
    const double tan_theta_safe = tan(theta_safe);
    out[6] = p + sin_phi * tan_theta_safe * q + cos_phi * tan_theta_safe * r;
    out[7] = cos_phi * q - sin_phi * r;
    out[8] = (sin_phi * q + cos_phi * r) / cos_theta_safe;

    /* :474-482 */
    const double Ixx = P->inertia_xx, Iyy = P->inertia_yy, Izz = P->inertia_zz;
    double rd0 = (l_moment - (Izz - Iyy) * q * r) / Ixx;
    double rd1 = (m_moment - (Ixx - Izz) * p * r) / Iyy;
    double rd2 = (n_moment - (Iyy - Ixx) * p * q) / Izz;

    /* :485-490 */
    const double maa = P->max_angular_acceleration, ma = P->max_acceleration;
    out[9] = np_clip(rd0, -maa, maa); out[10] = np_clip(rd1, -maa, maa); out[11] = np_clip(rd2, -maa, maa);
    out[3] = np_clip(vd0, -ma, ma); out[4] = np_clip(vd1, -ma, ma); out[5] = np_clip(vd2, -ma, ma);

    /* :496-501 */
    for (int i = 0; i < 12; ++i)
        if (!isfinite(out[i])) out[i] = 0.0;
}

/* Simplified6DOF.step, simulation/simplified_6dof.py:228-293 (the get_state at :293 is orc_derived). */
void orc_rk4_step(const orc_aircraft_params* P, double s[12], double* time, const double ctrl[4], double dt) {
    double k1[12], k2[12], k3[12], k4[12], t[12];
    /* :248-251 */
    orc_dynamics(P, s, ctrl, k1);
    for (int i = 0; i < 12; ++i) t[i] = s[i] + 0.5 * dt * k1[i];
    orc_dynamics(P, t, ctrl, k2);
    for (int i = 0; i < 12; ++i) t[i] = s[i] + 0.5 * dt * k2[i];
    orc_dynamics(P, t, ctrl, k3);
    for (int i = 0; i < 12; ++i) t[i] = s[i] + dt * k3[i];
    orc_dynamics(P, t, ctrl, k4);
    /* :253-254 */
    const double dt6 = dt / 6.0;
    for (int i = 0; i < 12; ++i) s[i] = s[i] + dt6 * (k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i]);
    *time += dt;

    /* :258 */
    for (int i = 0; i < 3; ++i) s[i] = np_nan_to_num(s[i], 10000.0, -10000.0);
    /* :261-262 */
    for (int i = 3; i < 6; ++i) s[i] = np_clip(s[i], -P->max_velocity, P->max_velocity);
    /* :266-270 */
    const double max_pitch = np_radians(P->max_pitch_angle);
    s[6] = atan2(sin(s[6]), cos(s[6]));
    s[7] = np_clip(s[7], -max_pitch, max_pitch);
    s[8] = atan2(sin(s[8]), cos(s[8]));
    /* :273 */
    const double max_rate = np_radians(P->max_angular_rate);
    for (int i = 9; i < 12; ++i) s[i] = np_clip(s[i], -max_rate, max_rate);
    /* :276-283 */
    const double altitude = -s[2];
    if (altitude < 0) {
        s[2] = 0.0;
        s[5] = py_max(0.0, s[5]);
    }
    /* :286-291 */
    for (int i = 0; i < 12; ++i)
        if (!isfinite(s[i])) s[i] = 0.0;
}

/* Simplified6DOF.get_state + _body_to_ned, simulation/simplified_6dof.py:295-331, 505-530 */
void orc_derived(const double s[12], double out[4]) {
    const double u = s[3], v = s[4], w = s[5];
    const double phi = s[6], theta = s[7], psi = s[8];
    const double sphi = sin(phi), cphi = cos(phi), sth = sin(theta), cth = cos(theta), sps = sin(psi), cps = cos(psi);
    /* :518-528, each entry with the reference's left-to-right products */
    const double r00 = cth * cps, r01 = sphi * sth * cps - cphi * sps, r02 = cphi * sth * cps + sphi * sps;
    const double r10 = cth * sps, r11 = sphi * sth * sps + cphi * cps, r12 = cphi * sth * sps - sphi * cps;
    const double vn = np_row_dot3(r00, r01, r02, u, v, w);
    const double ve = np_row_dot3(r10, r11, r12, u, v, w);
    out[0] = np_norm3(u, v, w);   /* :307 */
    out[1] = -s[2];               /* :310 */
    out[2] = np_norm2(vn, ve);    /* :319 */
    out[3] = atan2(ve, vn);       /* :317 */
}

/* set_controls, simulation/simplified_6dof.py:214-226 */
static inline void clip_controls(const double in[4], double out[4]) {
    out[0] = np_clip(in[0], -1.0, 1.0);
    out[1] = np_clip(in[1], -1.0, 1.0);
    out[2] = np_clip(in[2], -1.0, 1.0);
    out[3] = np_clip(in[3], 0.0, 1.0);
}

/* SimulationAircraftBackend.step, simulation/simulation_backend.py:82-101; dt check :241-245 */
static int substep_count(const orc_aircraft_params* P, double dt, double dt_physics, double* h) {
    int K = (int)(dt / dt_physics);
    if (K < 1) K = 1;
    *h = dt / K;
    if (*h <= P->min_timestep || *h > P->max_timestep) return -1;
    return K;
}

static void backend_step(const orc_aircraft_params* P, double s[12], double* time, const double ctrl[4],
                         int K, double h) {
    for (int k = 0; k < K; ++k) orc_rk4_step(P, s, time, ctrl, h);
}

/* ------------------------------------------------------------------------------------------
 * PID core (float, no FMA)  cpp/src/pid_controller.cpp:24-60,80-82
 * ---------------------------------------------------------------------------------------- */
static inline float pid_clamp(float value, float min_val, float max_val) {
    /* std::max(min_val, std::min(value, max_val)):  std::min(a,b) = (b<a)?b:a ; std::max(a,b) = (a<b)?b:a */
    float m = (max_val < value) ? max_val : value;
    return (min_val < m) ? m : min_val;
}

float orc_pid_compute(const orc_pid_config* c, orc_pid_state* st, float setpoint, float measurement, float dt) {
    st->error = setpoint - measurement;
    float p_term = c->kp * st->error;
    st->integral += st->error * dt;
    st->integral = pid_clamp(st->integral, c->integral_min, c->integral_max);
    float i_term = c->ki * st->integral;
    if (dt > 1e-6f) st->derivative = (st->error - st->error_prev) / dt;
    else st->derivative = 0.0f;
    st->derivative_filtered = c->derivative_filter_alpha * st->derivative
                              + (1.0f - c->derivative_filter_alpha) * st->derivative_filtered;
    float d_term = c->kd * st->derivative_filtered;
    st->output = p_term + i_term + d_term;
    st->output = pid_clamp(st->output, c->output_min, c->output_max);
    st->error_prev = st->error;
    return st->output;
}

/* pid[18][n] slot helpers: only integral, error_prev, derivative_filtered influence later outputs. */
static inline void pid_load(const float* pid, int64_t n, int64_t i, int loop, int axis, orc_pid_state* st) {
    const int base = (loop * 3 + axis) * 3;
    memset(st, 0, sizeof(*st));
    st->integral = pid[(int64_t)(base + 0) * n + i];
    st->error_prev = pid[(int64_t)(base + 1) * n + i];
    st->derivative_filtered = pid[(int64_t)(base + 2) * n + i];
}
static inline void pid_store(float* pid, int64_t n, int64_t i, int loop, int axis, const orc_pid_state* st) {
    const int base = (loop * 3 + axis) * 3;
    pid[(int64_t)(base + 0) * n + i] = st->integral;
    pid[(int64_t)(base + 1) * n + i] = st->error_prev;
    pid[(int64_t)(base + 2) * n + i] = st->derivative_filtered;
}

/* RateAgent.compute_action, controllers/rate_agent.py:65-124.  surf = elevator, aileron, rudder, throttle */
static void rate_agent(const orc_controller_params* C, orc_pid_state st[3], const double rate_cmd[3], double throttle,
                       const double s[12], double dt, double surf[4]) {
    /* :92-94 */
    const double p_cmd = np_clip(rate_cmd[0], -C->max_roll_rate, C->max_roll_rate);
    const double q_cmd = np_clip(rate_cmd[1], -C->max_pitch_rate, C->max_pitch_rate);
    const double r_cmd = np_clip(rate_cmd[2], -C->max_yaw_rate, C->max_yaw_rate);
    /* :96-104: acb.Vector3 narrows to float (cpp/bindings/bindings.cpp:89-100), dt narrows too */
    const float fdt = (float)dt;
    const float o_roll = orc_pid_compute(&C->rate[0], &st[0], (float)p_cmd, (float)s[9], fdt);
    const float o_pitch = orc_pid_compute(&C->rate[1], &st[1], (float)q_cmd, (float)s[10], fdt);
    const float o_yaw = orc_pid_compute(&C->rate[2], &st[2], (float)r_cmd, (float)s[11], fdt);
    /* :111-122 */
    surf[1] = np_clip((double)o_roll, -1.0, 1.0);
    surf[0] = np_clip(-(double)o_pitch, -1.0, 1.0);
    surf[2] = np_clip(-(double)o_yaw, -1.0, 1.0);
    surf[3] = np_clip(throttle, 0.0, 1.0);
}

/* AttitudeAgent.compute_action, controllers/attitude_agent.py:86-154.  att_cmd = roll, pitch, yaw (NaN = None) */
static void attitude_agent(const orc_controller_params* C, orc_pid_state ang[3], orc_pid_state rate[3],
                           const double att_cmd[3], double throttle, const double s[12], double dt, double surf[4]) {
    /* :116-119 */
    const double roll_cmd = np_clip(att_cmd[0], -C->max_roll, C->max_roll);
    const double pitch_cmd = np_clip(att_cmd[1], -C->max_pitch, C->max_pitch);
    /* NaN = yaw_angle None -> 0.0 (:119); +inf = "hold current yaw", the command HSAAgent issues (hsa_agent.py:214-216) */
    const double yaw_in = isinf(att_cmd[2]) ? s[8] : att_cmd[2];
    const double yaw_cmd = (yaw_in != yaw_in) ? 0.0 : orc_wrap_angle(yaw_in);
    /* :125-129 */
    const double current_yaw = orc_wrap_angle(s[8]);
    /* :134 */
    const float fdt = (float)dt;
    const float o_roll = orc_pid_compute(&C->angle[0], &ang[0], (float)roll_cmd, (float)s[6], fdt);
    const float o_pitch = orc_pid_compute(&C->angle[1], &ang[1], (float)pitch_cmd, (float)s[7], fdt);
    const float o_yaw = orc_pid_compute(&C->angle[2], &ang[2], (float)yaw_cmd, (float)current_yaw, fdt);
    /* :137-139 */
    double rc[3];
    rc[0] = np_clip((double)o_roll, -C->max_roll_rate, C->max_roll_rate);
    rc[1] = np_clip((double)o_pitch, -C->max_pitch_rate, C->max_pitch_rate);
    rc[2] = np_clip((double)o_yaw, -C->max_yaw_rate, C->max_yaw_rate);
    /* :142-152 */
    rate_agent(C, rate, rc, throttle, s, dt, surf);
}

/* HSAAgent.compute_action, controllers/hsa_agent.py:119-231.  outer = heading, energy, balance PIDs.
 * d = airspeed, altitude, ground_speed, heading of the CURRENT state (what get_state() hands the agent). */
static void hsa_agent(const orc_controller_params* C, orc_pid_state outer[3], orc_pid_state ang[3], orc_pid_state rate[3],
                      double heading_cmd, double speed_cmd, double altitude_cmd, const double s[12], const double d[4],
                      double dt, double surf[4]) {
    const float fdt = (float)dt;
    /* :147-160 */
    const double heading_error = orc_wrap_angle(heading_cmd - d[3]);
    const double virtual_setpoint = d[3] + heading_error;
    double roll_angle = (double)orc_pid_compute(&C->heading, &outer[0], (float)virtual_setpoint, (float)d[3], fdt);
    roll_angle = np_clip(roll_angle, -C->max_bank_rad, C->max_bank_rad);
    /* :167-186 */
    const double g = 9.81, h = d[1], V = d[0];
    const double E_specific = g * h + 0.5 * np_sq(V);
    const double E_specific_cmd = g * altitude_cmd + 0.5 * np_sq(speed_cmd);
    const double E_balance = g * h - 0.5 * np_sq(V);
    const double E_balance_cmd = g * altitude_cmd - 0.5 * np_sq(speed_cmd);
    /* :192-197 */
    const double throttle_adjustment = (double)orc_pid_compute(&C->energy, &outer[1], (float)E_specific_cmd, (float)E_specific, fdt);
    double throttle = C->baseline_throttle + throttle_adjustment;
    throttle = np_clip(throttle, 0.0, 1.0);
    /* :203-213 */
    double pitch_angle = (double)orc_pid_compute(&C->balance, &outer[2], (float)E_balance_cmd, (float)E_balance, fdt);
    const double cos_roll = cos(roll_angle);
    if (fabs(cos_roll) > 0.01) {
        const double load_factor = 1.0 / cos_roll;
        pitch_angle += C->load_factor_gain * (load_factor - 1.0);
    }
    pitch_angle = np_clip(pitch_angle, -C->max_pitch_cmd_rad, C->max_pitch_cmd_rad);
    /* :216-229: yaw command = current yaw */
    const double att[3] = {roll_angle, pitch_angle, s[8]};
    attitude_agent(C, ang, rate, att, throttle, s, dt, surf);
}

/* WaypointAgent.compute_action, controllers/waypoint_agent.py:81-242 -> (heading, speed, altitude) for the HSA level.
 * wp = north, east, down, speed (NaN = None). */
static void waypoint_guidance(const orc_controller_params* C, int guidance, const double wp[4], const double s[12],
                              const double d[4], double hsa_cmd[3]) {
    const double e0 = wp[0] - s[0], e1 = wp[1] - s[1];                           /* :244-253 */
    const double horizontal_distance = np_norm2(e0, e1);                          /* :117 */
    const double airspeed = d[0], heading = d[3];
    double heading_cmd;
    if (guidance == 0) { /* LOS :121-146 */
        heading_cmd = atan2(e1, e0);
        const double V = py_max(airspeed, 10.0);
        const double max_bank = np_radians(C->los_max_bank_deg);
        const double turn_radius = np_sq(V) / (9.81 * tan(max_bank));
        double heading_error = heading_cmd - heading;
        heading_error = atan2(sin(heading_error), cos(heading_error));
        const double anticipation_dist = turn_radius * fabs(heading_error) / np_radians(90.0);
        const double lead_angle_rad = np_radians(C->los_lead_angle_deg);
        if (horizontal_distance < anticipation_dist && fabs(heading_error) > np_radians(20.0)) {
            const double blend = 1.0 - (horizontal_distance / anticipation_dist);
            heading_cmd = heading_cmd + blend * lead_angle_rad * np_sign(heading_error);
            heading_cmd = atan2(sin(heading_cmd), cos(heading_cmd));
        }
    } else { /* pure pursuit :148-182 */
        const double V = py_max(airspeed, 10.0);
        const double max_bank = np_radians(C->max_bank_angle_deg);
        const double turn_radius = np_sq(V) / (9.81 * tan(max_bank));
        double lookahead_dist = C->lookahead_time * V;
        const double proximity_dist = C->proximity_scale_distance * turn_radius;
        if (horizontal_distance < proximity_dist) {
            const double scale = 0.6 + 0.4 * (horizontal_distance / proximity_dist);
            lookahead_dist *= scale;
        }
        lookahead_dist = np_clip(lookahead_dist, C->lookahead_min, C->lookahead_max);
        if (horizontal_distance > 1.0) {
            const double dir0 = e0 / horizontal_distance, dir1 = e1 / horizontal_distance;
            const double effective = (horizontal_distance < lookahead_dist) ? horizontal_distance : lookahead_dist; /* min(a,b) */
            heading_cmd = atan2(dir1 * effective, dir0 * effective);
        } else {
            heading_cmd = atan2(e1, e0);
        }
    }
    heading_cmd = atan2(sin(heading_cmd), cos(heading_cmd));                      /* :189 */
    /* :207-232 */
    const double altitude_cmd = -wp[2];
    double speed_cmd = (wp[3] != wp[3]) ? airspeed : wp[3];
    const double heading_to_wp = atan2(e1, e0);
    double heading_error = fabs(heading_to_wp - heading);
    heading_error = fabs(atan2(sin(heading_error), cos(heading_error)));
    const double threshold_angle = np_radians(C->turn_threshold_angle_deg);
    if (horizontal_distance < C->turn_threshold_distance && heading_error > threshold_angle) {
        const double reduction = C->max_speed_reduction * (1.0 - horizontal_distance / C->turn_threshold_distance);
        speed_cmd = speed_cmd * (1.0 - reduction);
        speed_cmd = py_max(speed_cmd, C->min_speed);
    }
    hsa_cmd[0] = heading_cmd; hsa_cmd[1] = speed_cmd; hsa_cmd[2] = altitude_cmd;
}

/* ------------------------------------------------------------------------------------------
 * batched drivers: physics and cascade
 * ---------------------------------------------------------------------------------------- */
static int resolve_threads(int nthreads) {
#ifdef _OPENMP
    return nthreads > 0 ? nthreads : omp_get_max_threads();
#else
    (void)nthreads;
    return 1;
#endif
}

int orc_step_surface(const orc_aircraft_params* P, int64_t n, double* state, const double* ctrl_in,
                     double* ctrl_held, double* derived, double dt, double dt_physics, int nthreads) {
    double h;
    const int K = substep_count(P, dt, dt_physics, &h);
    if (K < 0) return -1;
    const int nt = resolve_threads(nthreads);
    (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        double s[12], time, cin[4], c[4], d[4];
        for (int f = 0; f < 12; ++f) s[f] = state[(int64_t)f * n + i];
        time = state[12 * n + i];
        for (int f = 0; f < 4; ++f) cin[f] = ctrl_in[(int64_t)f * n + i];
        clip_controls(cin, c);
        backend_step(P, s, &time, c, K, h);
        for (int f = 0; f < 12; ++f) state[(int64_t)f * n + i] = s[f];
        state[12 * n + i] = time;
        if (ctrl_held) for (int f = 0; f < 4; ++f) ctrl_held[(int64_t)f * n + i] = c[f];
        if (derived) {
            orc_derived(s, d);
            for (int f = 0; f < 4; ++f) derived[(int64_t)f * n + i] = d[f];
        }
    }
    return K;
}

int orc_step_cascade(const orc_aircraft_params* P, const orc_controller_params* C, int level, int64_t n,
                     double* state, const double* cmd, float* pid, double* surfaces_out, double* derived,
                     double dt_ctrl, double dt, double dt_physics, int nthreads) {
    double h;
    const int K = substep_count(P, dt, dt_physics, &h);
    if (K < 0) return -1;
    if (level != 3 && level != 4) return -2;
    const int nt = resolve_threads(nthreads);
    (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        double s[12], time, surf[4], c[4], d[4], cm[4];
        orc_pid_state rate[3], ang[3];
        for (int f = 0; f < 12; ++f) s[f] = state[(int64_t)f * n + i];
        time = state[12 * n + i];
        for (int f = 0; f < 4; ++f) cm[f] = cmd[(int64_t)f * n + i];
        for (int a = 0; a < 3; ++a) { pid_load(pid, n, i, 0, a, &rate[a]); pid_load(pid, n, i, 1, a, &ang[a]); }
        if (level == 4) rate_agent(C, rate, cm, cm[3], s, dt_ctrl, surf);
        else attitude_agent(C, ang, rate, cm, cm[3], s, dt_ctrl, surf);
        for (int a = 0; a < 3; ++a) { pid_store(pid, n, i, 0, a, &rate[a]); if (level == 3) pid_store(pid, n, i, 1, a, &ang[a]); }
        clip_controls(surf, c);
        backend_step(P, s, &time, c, K, h);
        for (int f = 0; f < 12; ++f) state[(int64_t)f * n + i] = s[f];
        state[12 * n + i] = time;
        if (surfaces_out) for (int f = 0; f < 4; ++f) surfaces_out[(int64_t)f * n + i] = surf[f];
        if (derived) {
            orc_derived(s, d);
            for (int f = 0; f < 4; ++f) derived[(int64_t)f * n + i] = d[f];
        }
    }
    return K;
}

static inline void outer_load(const float* p, int64_t n, int64_t i, orc_pid_state outer[3]) {
    for (int k = 0; k < 3; ++k) {
        memset(&outer[k], 0, sizeof outer[k]);
        outer[k].integral = p[(int64_t)(k * 3 + 0) * n + i];
        outer[k].error_prev = p[(int64_t)(k * 3 + 1) * n + i];
        outer[k].derivative_filtered = p[(int64_t)(k * 3 + 2) * n + i];
    }
}
static inline void outer_store(float* p, int64_t n, int64_t i, const orc_pid_state outer[3]) {
    for (int k = 0; k < 3; ++k) {
        p[(int64_t)(k * 3 + 0) * n + i] = outer[k].integral;
        p[(int64_t)(k * 3 + 1) * n + i] = outer[k].error_prev;
        p[(int64_t)(k * 3 + 2) * n + i] = outer[k].derivative_filtered;
    }
}

/* one aircraft, levels 1-2: agent on the current state, then set_controls + step */
static void guided_step_one(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, int64_t i, double* state,
                            const double hsa_cmd[3], float* pid, float* pid_outer, double* surfaces_out, double* derived,
                            double dt_ctrl, int K, double h) {
    double s[12], time, surf[4], c[4], d[4];
    orc_pid_state rate[3], ang[3], outer[3];
    for (int f = 0; f < 12; ++f) s[f] = state[(int64_t)f * n + i];
    time = state[12 * n + i];
    for (int a = 0; a < 3; ++a) { pid_load(pid, n, i, 0, a, &rate[a]); pid_load(pid, n, i, 1, a, &ang[a]); }
    outer_load(pid_outer, n, i, outer);
    orc_derived(s, d);
    hsa_agent(C, outer, ang, rate, hsa_cmd[0], hsa_cmd[1], hsa_cmd[2], s, d, dt_ctrl, surf);
    for (int a = 0; a < 3; ++a) { pid_store(pid, n, i, 0, a, &rate[a]); pid_store(pid, n, i, 1, a, &ang[a]); }
    outer_store(pid_outer, n, i, outer);
    clip_controls(surf, c);
    backend_step(P, s, &time, c, K, h);
    for (int f = 0; f < 12; ++f) state[(int64_t)f * n + i] = s[f];
    state[12 * n + i] = time;
    if (surfaces_out) for (int f = 0; f < 4; ++f) surfaces_out[(int64_t)f * n + i] = surf[f];
    if (derived) {
        orc_derived(s, d);
        for (int f = 0; f < 4; ++f) derived[(int64_t)f * n + i] = d[f];
    }
}

int orc_step_hsa(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, double* state, const double* cmd,
                 float* pid, float* pid_outer, double* surfaces_out, double* derived, double dt_ctrl, double dt,
                 double dt_physics, int nthreads) {
    double h;
    const int K = substep_count(P, dt, dt_physics, &h);
    if (K < 0) return -1;
    const int nt = resolve_threads(nthreads);
    (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        const double hc[3] = {cmd[i], cmd[n + i], cmd[2 * n + i]};
        guided_step_one(P, C, n, i, state, hc, pid, pid_outer, surfaces_out, derived, dt_ctrl, K, h);
    }
    return K;
}

/* mission_stats (nullable, double[16], accumulated): the fleet statistics acb200_mission.stats holds (include/acb200.h) --
 * [0] aircraft-steps flown, [1] waypoints reached (MissionPlanner.waypoints_reached summed over the fleet,
 * mission_planner.py:151-153), [2] laps closed, [3] missions completed (:163-165), [4] sum |cross-track error| (horizontal
 * distance from the leg previous -> active waypoint), [5] its squares, [6] sum |altitude - waypoint altitude|, [7] sum 3-D
 * distance to the active waypoint (:188-205), [8] sum airspeed, [9] steps below 5 m; from the state the guidance sees. */
int orc_step_waypoint_stats(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, double* state,
                            const double* waypoints, int n_wp, int32_t* wp_index, int guidance, double acceptance_radius, int loop,
                            float* pid, float* pid_outer, double* surfaces_out, double* derived, double* hsa_cmd_out,
                            double dt_ctrl, double dt, double dt_physics, int nthreads, double* mission_stats) {
    double h;
    const int K = substep_count(P, dt, dt_physics, &h);
    if (K < 0) return -1;
    if (n_wp < 1) return -2;
    const int nt = resolve_threads(nthreads);
    (void)nt;
    double ms[16];
    for (int k = 0; k < 16; ++k) ms[k] = 0.0;
#pragma omp parallel for num_threads(nt) schedule(static) reduction(+ : ms[:16])
    for (int64_t i = 0; i < n; ++i) {
        int32_t idx = wp_index[i];
        if (idx >= n_wp) continue;                               /* mission complete: the demo stops stepping (:196-203) */
        /* MissionPlanner.update, mission_planner.py:128-171: 3-D distance < acceptance radius -> next waypoint */
        const double* wp = waypoints + (int64_t)idx * 4;
        const double dist = np_norm3(wp[0] - state[i], wp[1] - state[n + i], wp[2] - state[2 * n + i]);
        int advanced = 0, lapped = 0;
        if (dist < acceptance_radius) {
            idx += 1;
            advanced = 1;
            if (idx >= n_wp && loop) { idx = 0; lapped = 1; }
            wp_index[i] = idx;
            if (idx >= n_wp) { ms[1] += 1.0; ms[3] += 1.0; continue; }
            wp = waypoints + (int64_t)idx * 4;
        }
        double s[12], d[4], hc[3];
        for (int f = 0; f < 12; ++f) s[f] = state[(int64_t)f * n + i];
        orc_derived(s, d);
        if (mission_stats) {
            ms[0] += 1.0; ms[1] += advanced; ms[2] += lapped;
            int prev = idx - 1;
            if (prev < 0) prev = loop ? n_wp - 1 : -1;
            double ct = 0.0;
            if (prev >= 0) {
                const double ax = waypoints[prev * 4 + 0], ay = waypoints[prev * 4 + 1];
                const double bx = wp[0] - ax, by = wp[1] - ay, px = s[0] - ax, py = s[1] - ay;
                const double len = sqrt(bx * bx + by * by);
                if (len > 1e-6) ct = fabs((px * by - py * bx) / len);
            }
            ms[4] += ct; ms[5] += ct * ct;
            ms[6] += fabs(d[1] - (-wp[2]));
            ms[7] += np_norm3(wp[0] - s[0], wp[1] - s[1], wp[2] - s[2]);
            ms[8] += d[0];
            ms[9] += (d[1] < 5.0) ? 1.0 : 0.0;
        }
        waypoint_guidance(C, guidance, wp, s, d, hc);
        if (hsa_cmd_out) for (int f = 0; f < 3; ++f) hsa_cmd_out[(int64_t)f * n + i] = hc[f];
        guided_step_one(P, C, n, i, state, hc, pid, pid_outer, surfaces_out, derived, dt_ctrl, K, h);
    }
    if (mission_stats) for (int k = 0; k < 16; ++k) mission_stats[k] += ms[k];
    return K;
}

int orc_step_waypoint(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, double* state,
                      const double* waypoints, int n_wp, int32_t* wp_index, int guidance, double acceptance_radius, int loop,
                      float* pid, float* pid_outer, double* surfaces_out, double* derived, double* hsa_cmd_out,
                      double dt_ctrl, double dt, double dt_physics, int nthreads) {
    return orc_step_waypoint_stats(P, C, n, state, waypoints, n_wp, wp_index, guidance, acceptance_radius, loop, pid, pid_outer,
                                   surfaces_out, derived, hsa_cmd_out, dt_ctrl, dt, dt_physics, nthreads, NULL);
}

/* ------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al., SC'11) -- the counter-based generator the CUDA kernels use for
 * resets.  The reference draws from three MT19937 streams per env (rate_env.py:64,73-77), which a
 * counter-based generator cannot replay; parity for resets is distributional (tests/test_reset_sampler.py).
 * ---------------------------------------------------------------------------------------- */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* 53-bit uniform in [0,1) from two words, the construction numpy's random_sample uses. */
static inline double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}
/* RandomState.uniform(lo, hi) = lo + (hi-lo)*random_sample() */
static inline double uniform_lh(double lo, double hi, double u) { return lo + (hi - lo) * u; }

/* generators.py:41-45: difficulty_scale is the Python float 0.3 / 0.5 / 0.7.  The fleet keeps it as float32 (so that the
 * curriculum is one small tensor write); (double)0.3f = 0.30000001192..., 4e-8 away.  The value the reference multiplies
 * with is recovered as the nearest multiple of 1e-6 (correctly rounded quotient = the literal's double). */
double orc_difficulty_scale(float difficulty) { return rint((double)difficulty * 1e6) / 1e6; }

typedef struct { uint32_t key[2]; uint32_t ctr[4]; } rng_t;

static inline void rng_init(rng_t* g, uint64_t seed, int64_t global_id, uint32_t episode) {
    g->key[0] = (uint32_t)seed; g->key[1] = (uint32_t)(seed >> 32);
    g->ctr[0] = (uint32_t)(uint64_t)global_id; g->ctr[1] = (uint32_t)((uint64_t)global_id >> 32);
    g->ctr[2] = episode; g->ctr[3] = 0;
}
static inline void rng_block(rng_t* g, uint32_t block, uint32_t out[4]) {
    g->ctr[3] = block;
    orc_philox4x32_10(g->ctr, g->key, out);
}

/* ------------------------------------------------------------------------------------------
 * NoisySensorInterface.update for a fleet -- interfaces/sensor.py:200-243.
 *   measured position/velocity/attitude = truth + sigma * z           (:209-217)
 *   measured angular_rate = truth + sigma_gyro * z + gyro_bias        (:218-222, the bias BEFORE this update's walk)
 *   measured airspeed / altitude = truth + sigma * z                  (:223-228)
 *   gyro_bias += 1e-4 * z;  accel_bias += 1e-3 * z                    (:230-231)
 * in that draw order (20 normals per update).  The reference's z come from numpy's PCG64 ziggurat (sensor.py:189-190),
 * which cannot be replayed by a counter-based generator; here z = Box-Muller on Philox4x32-10 words keyed
 * (seed; global id, tick, 'SEN'+block), the stream csrc/sensor.cu uses, so kernel-vs-oracle parity is tight and
 * parity with the reference is distributional (tests/test_sensor.py against tests/golden/sensor_stats.json).
 * ---------------------------------------------------------------------------------------- */
#define ORC_SENSOR_STREAM 0x53454E00u

static void normal_pair(uint32_t w0, uint32_t w1, double* z0, double* z1) {
    const double two_pi = 6.283185307179586476925286766559;
    const double u1 = ((double)w0 + 0.5) * (1.0 / 4294967296.0);
    const double u2 = ((double)w1 + 0.5) * (1.0 / 4294967296.0);
    const double r = sqrt(-2.0 * log(u1));
    *z0 = r * cos(two_pi * u2);
    *z1 = r * sin(two_pi * u2);
}

void orc_sensor_normals(uint64_t seed, int64_t global_id, uint32_t tick, double z[20]) {
    rng_t g;
    rng_init(&g, seed, global_id, tick);
    for (uint32_t k = 0; k < 5; ++k) {
        uint32_t w[4];
        rng_block(&g, ORC_SENSOR_STREAM + k, w);
        normal_pair(w[0], w[1], &z[4 * k + 0], &z[4 * k + 1]);
        normal_pair(w[2], w[3], &z[4 * k + 2], &z[4 * k + 3]);
    }
}

static int sensor_sigma_ok(const orc_sensor_config* S) {
    const double sd[9] = {S->imu_gyro_stddev, S->imu_accel_stddev, S->gps_position_stddev, S->gps_velocity_stddev,
                          S->airspeed_stddev, S->altitude_stddev, S->attitude_stddev, S->gyro_bias_walk_stddev,
                          S->accel_bias_walk_stddev};
    for (int k = 0; k < 9; ++k) if (!(sd[k] >= 0.0) || isinf(sd[k])) return 0;   /* Generator.normal: "scale < 0" */
    return 1;
}

/* one aircraft, given its 20 standard normals in the reference's draw order */
static void sensor_apply_one(const orc_sensor_config* S, double s[13], double d[4], double b[6], const double z[20]) {
    for (int a = 0; a < 3; ++a) {
        s[a] = s[a] + S->gps_position_stddev * z[a];
        s[3 + a] = s[3 + a] + S->gps_velocity_stddev * z[3 + a];
        s[6 + a] = s[6 + a] + S->attitude_stddev * z[6 + a];
        s[9 + a] = (s[9 + a] + S->imu_gyro_stddev * z[9 + a]) + b[a];
    }
    d[0] = d[0] + S->airspeed_stddev * z[12];
    d[1] = d[1] + S->altitude_stddev * z[13];
    for (int a = 0; a < 3; ++a) {
        b[a] = b[a] + S->gyro_bias_walk_stddev * z[14 + a];
        b[3 + a] = b[3 + a] + S->accel_bias_walk_stddev * z[17 + a];
    }
}

/* z_ext NULL: z = the Philox/Box-Muller stream of (seed, id_offset + i, tick).  z_ext [20][n]: caller-supplied standard
 * normals (the tests pass numpy's PCG64 draws to compare with the reference bit for bit). */
int orc_sensor_update(const orc_sensor_config* S, int64_t n, const double* state, const double* derived, double* bias,
                      uint32_t tick, const double* z_ext, double* m_state, double* m_derived) {
    if (!sensor_sigma_ok(S)) return -1;
    for (int64_t i = 0; i < n; ++i) {
        double s[13], d[4];
        for (int f = 0; f < 13; ++f) s[f] = state[(int64_t)f * n + i];
        for (int f = 0; f < 4; ++f) d[f] = derived[(int64_t)f * n + i];
        if (S->enabled) {
            double z[20], b[6];
            if (z_ext) for (int k = 0; k < 20; ++k) z[k] = z_ext[(int64_t)k * n + i];
            else orc_sensor_normals(S->seed, S->id_offset + i, tick, z);
            for (int f = 0; f < 6; ++f) b[f] = bias[(int64_t)f * n + i];
            sensor_apply_one(S, s, d, b, z);
            for (int f = 0; f < 6; ++f) bias[(int64_t)f * n + i] = b[f];
        }
        for (int f = 0; f < 13; ++f) m_state[(int64_t)f * n + i] = s[f];
        for (int f = 0; f < 4; ++f) m_derived[(int64_t)f * n + i] = d[f];
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * environment
 * ---------------------------------------------------------------------------------------- */
#define OBS 18

typedef struct {
    double s[12], time;
    double derived[4];
    orc_pid_state rate[3], ang[3];
    double att_cmd[4];
    double rate_cmd[3];
    float prev_action[4];
    double prev_err[3], sign_changes[3], settle_timer;
    double sched[7];
    double env_time;
    int32_t step_count;
    uint8_t is_settled, command_type;
    float difficulty;
    uint32_t episode_idx;
    double ep_return;
} env_t;

static void env_gather(const orc_env_buffers* B, int64_t i, env_t* e) {
    const int64_t n = B->n;
    for (int f = 0; f < 12; ++f) e->s[f] = B->state[(int64_t)f * n + i];
    e->time = B->state[12 * n + i];
    for (int f = 0; f < 4; ++f) e->derived[f] = B->derived[(int64_t)f * n + i];
    for (int a = 0; a < 3; ++a) { pid_load(B->pid, n, i, 0, a, &e->rate[a]); pid_load(B->pid, n, i, 1, a, &e->ang[a]); }
    for (int f = 0; f < 4; ++f) e->att_cmd[f] = B->att_cmd ? B->att_cmd[(int64_t)f * n + i] : 0.0;
    for (int f = 0; f < 3; ++f) e->rate_cmd[f] = B->rate_cmd[(int64_t)f * n + i];
    for (int f = 0; f < 4; ++f) e->prev_action[f] = B->prev_action[(int64_t)f * n + i];
    for (int f = 0; f < 3; ++f) { e->prev_err[f] = B->rew_mem[(int64_t)f * n + i]; e->sign_changes[f] = B->rew_mem[(int64_t)(3 + f) * n + i]; }
    e->settle_timer = B->rew_mem[6 * n + i];
    for (int f = 0; f < 7; ++f) e->sched[f] = B->sched[(int64_t)f * n + i];
    e->env_time = B->env_time[i];
    e->step_count = B->step_count[i];
    e->is_settled = B->is_settled[i];
    e->command_type = B->command_type[i];
    e->difficulty = B->difficulty[i];
    e->episode_idx = B->episode_idx[i];
    e->ep_return = B->ep_return[i];
}

static void env_scatter(orc_env_buffers* B, int64_t i, const env_t* e) {
    const int64_t n = B->n;
    for (int f = 0; f < 12; ++f) B->state[(int64_t)f * n + i] = e->s[f];
    B->state[12 * n + i] = e->time;
    for (int f = 0; f < 4; ++f) B->derived[(int64_t)f * n + i] = e->derived[f];
    for (int a = 0; a < 3; ++a) { pid_store(B->pid, n, i, 0, a, &e->rate[a]); pid_store(B->pid, n, i, 1, a, &e->ang[a]); }
    for (int f = 0; f < 3; ++f) B->rate_cmd[(int64_t)f * n + i] = e->rate_cmd[f];
    for (int f = 0; f < 4; ++f) B->prev_action[(int64_t)f * n + i] = e->prev_action[f];
    for (int f = 0; f < 3; ++f) { B->rew_mem[(int64_t)f * n + i] = e->prev_err[f]; B->rew_mem[(int64_t)(3 + f) * n + i] = e->sign_changes[f]; }
    B->rew_mem[6 * n + i] = e->settle_timer;
    for (int f = 0; f < 7; ++f) B->sched[(int64_t)f * n + i] = e->sched[f];
    B->env_time[i] = e->env_time;
    B->step_count[i] = e->step_count;
    B->is_settled[i] = e->is_settled;
    B->episode_idx[i] = e->episode_idx;
    B->ep_return[i] = e->ep_return;
}

/* RateControlEnv._get_observation, rate_env.py:374-408 */
static void env_observe(const env_t* e, float* obs) {
    obs[0] = (float)e->s[9]; obs[1] = (float)e->s[10]; obs[2] = (float)e->s[11];
    obs[3] = (float)e->rate_cmd[0]; obs[4] = (float)e->rate_cmd[1]; obs[5] = (float)e->rate_cmd[2];
    obs[6] = (float)(e->rate_cmd[0] - e->s[9]);
    obs[7] = (float)(e->rate_cmd[1] - e->s[10]);
    obs[8] = (float)(e->rate_cmd[2] - e->s[11]);
    obs[9] = (float)e->derived[0]; obs[10] = (float)e->derived[1];
    obs[11] = (float)e->s[6]; obs[12] = (float)e->s[7]; obs[13] = (float)e->s[8];
    obs[14] = e->prev_action[0]; obs[15] = e->prev_action[1]; obs[16] = e->prev_action[2]; obs[17] = e->prev_action[3];
}

/* the six ordered draws of `rng.choice(3, size=k, replace=False)` are the 6 permutations' prefixes */
static const uint8_t PERM3[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};

/* RateControlEnv.reset, rate_env.py:151-210 with FlightEnvelopeSampler.sample (generators.py:190-214)
 * and _generate_new_command (rate_env.py:302-340, generators.py:47-132).  Draw plan (Philox block b of
 * counter (id_lo, id_hi, episode, b)):
 *   b0: airspeed U(15,30), altitude U(50,200)      b1: roll, pitch U(+-15deg)
 *   b2: yaw U(0,2pi), p U(+-0.1)                    b3: q, r U(+-0.1)
 *   b4: w0 -> number of axes, w1 -> permutation, w2 bits 0..2 -> signs, w3 spare
 *   b5: magnitude axis slot 0, slot 1               b6: magnitude slot 2, sine frequency
 */
static void env_reset_one(const orc_env_config* E, int64_t local_i, env_t* e) {
    rng_t g;
    uint32_t w[4];
    e->episode_idx += 1;
    rng_init(&g, E->seed, E->id_offset + local_i, e->episode_idx);

    const double att_lo = np_radians(-15.0), att_hi = np_radians(15.0);
    rng_block(&g, 0, w);
    const double airspeed = uniform_lh(15.0, 30.0, u53(w[0], w[1]));
    const double altitude = uniform_lh(50.0, 200.0, u53(w[2], w[3]));
    rng_block(&g, 1, w);
    const double roll = uniform_lh(att_lo, att_hi, u53(w[0], w[1]));
    const double pitch = uniform_lh(att_lo, att_hi, u53(w[2], w[3]));
    rng_block(&g, 2, w);
    const double yaw = uniform_lh(0.0, 2 * ORC_PI, u53(w[0], w[1]));
    const double p = uniform_lh(-0.1, 0.1, u53(w[2], w[3]));
    rng_block(&g, 3, w);
    const double q = uniform_lh(-0.1, 0.1, u53(w[0], w[1]));
    const double r = uniform_lh(-0.1, 0.1, u53(w[2], w[3]));

    /* rate_env.py:174-185 + simplified_6dof.py:206-212 */
    e->s[0] = 0.0; e->s[1] = 0.0; e->s[2] = -altitude;
    e->s[3] = airspeed; e->s[4] = 0.0; e->s[5] = 0.0;
    e->s[6] = roll; e->s[7] = pitch; e->s[8] = yaw;
    e->s[9] = p; e->s[10] = q; e->s[11] = r;
    e->time = 0.0;
    orc_derived(e->s, e->derived);

    /* command generation */
    const double max_rates[3] = {np_radians(180.0), np_radians(180.0), np_radians(160.0)}; /* generators.py:16-18 */
    const double scale = orc_difficulty_scale(e->difficulty);
    uint32_t sel[4];
    rng_block(&g, 4, sel);
    double mag[3], ufreq;
    rng_block(&g, 5, w);
    mag[0] = u53(w[0], w[1]); mag[1] = u53(w[2], w[3]);
    rng_block(&g, 6, w);
    mag[2] = u53(w[0], w[1]); ufreq = u53(w[2], w[3]);
    const uint8_t* perm = PERM3[(uint32_t)(((uint64_t)sel[1] * 6u) >> 32)];
    for (int f = 0; f < 7; ++f) e->sched[f] = 0.0;
    e->rate_cmd[0] = e->rate_cmd[1] = e->rate_cmd[2] = 0.0;

    if (e->command_type == ORC_CMD_STEP || e->command_type == ORC_CMD_RAMP) {
        /* generators.py:47-72 / 74-98: n_axes in {1,2,3}, axes without replacement, U(0.3,1)*scale, random sign */
        const int n_axes = 1 + (int)(((uint64_t)sel[0] * 3u) >> 32);
        double cmd[3] = {0.0, 0.0, 0.0};
        for (int k = 0; k < n_axes; ++k) {
            const int axis = perm[k];
            const double magnitude = uniform_lh(0.3, 1.0, mag[k]) * scale;
            const double sign = ((sel[2] >> k) & 1u) ? 1.0 : -1.0;
            cmd[axis] = sign * magnitude * max_rates[axis];
        }
        if (e->command_type == ORC_CMD_STEP) {
            for (int a = 0; a < 3; ++a) e->rate_cmd[a] = cmd[a];
        } else {
            for (int a = 0; a < 3; ++a) e->sched[a] = cmd[a]; /* ramp end; start = 0 (rate_env.py:312-321) */
        }
    } else if (e->command_type == ORC_CMD_SINE) {
        /* generators.py:100-132: f U(0.1,2), 1-2 axes, amp = U(0.3,1)*0.5*scale*max */
        const int n_axes = 1 + (int)(((uint64_t)sel[0] * 2u) >> 32);
        for (int k = 0; k < n_axes; ++k) {
            const int axis = perm[k];
            const double amp = uniform_lh(0.3, 1.0, mag[k]) * 0.5 * scale;
            e->sched[axis] = amp * max_rates[axis];
        }
        e->sched[3] = uniform_lh(0.1, 2.0, ufreq);
    }
    /* ORC_CMD_RANDOM: starts from zero (rate_env.py:333-336) */

    /* rate_env.py:191-198 */
    e->env_time = 0.0;
    e->step_count = 0;
    e->prev_action[0] = 0.0f; e->prev_action[1] = 0.0f; e->prev_action[2] = 0.0f; e->prev_action[3] = 0.5f;
    for (int a = 0; a < 3; ++a) { e->prev_err[a] = 0.0; e->sign_changes[a] = 0.0; }
    e->settle_timer = 0.0;
    e->is_settled = 0;
    e->ep_return = 0.0;
    /* agents are reset with the env (pid_demonstrations.py:49, residual_rate_env.py:89) */
    memset(e->rate, 0, sizeof(e->rate));
    memset(e->ang, 0, sizeof(e->ang));
}

/* RateControlEnv._update_command, rate_env.py:342-372 (+ generators.py:145-164 for the random walk) */
void orc_walk_normals(uint64_t seed, int64_t global_id, uint32_t episode, int32_t step_count, double z[3]) {
    rng_t g;
    uint32_t w[4];
    rng_init(&g, seed, global_id, episode);
    rng_block(&g, 16u + (uint32_t)step_count, w);
    /* Box-Muller on (0,1] x [0,1) uniforms built from 32-bit words */
    const double u0 = ((double)w[0] + 1.0) / 4294967296.0, u1 = (double)w[1] / 4294967296.0;
    const double u2 = ((double)w[2] + 1.0) / 4294967296.0, u3 = (double)w[3] / 4294967296.0;
    const double r0 = sqrt(-2.0 * log(u0)), r1 = sqrt(-2.0 * log(u2));
    z[0] = r0 * cos(2 * ORC_PI * u1); z[1] = r0 * sin(2 * ORC_PI * u1); z[2] = r1 * cos(2 * ORC_PI * u3);
}

static void env_update_command(const orc_env_config* E, int64_t local_i, env_t* e, const double* walk_z) {
    if (e->command_type == ORC_CMD_RAMP) {
        const double t = e->env_time - 0.0, duration = 3.0;
        if (t < duration) {
            const double a = t / duration;
            for (int k = 0; k < 3; ++k) e->rate_cmd[k] = (1 - a) * 0.0 + a * e->sched[k];
        } else {
            for (int k = 0; k < 3; ++k) e->rate_cmd[k] = e->sched[k];
        }
    } else if (e->command_type == ORC_CMD_RANDOM) {
        /* rate_env.py:358-365 with generators.py:145-164:
         *   delta = randn(3) * diffusion * sqrt(dt);  delta *= difficulty_scale * max_rates;  cmd += delta;  clip */
        const double max_rates[3] = {np_radians(180.0), np_radians(180.0), np_radians(160.0)};
        double z[3];
        if (walk_z) { for (int k = 0; k < 3; ++k) z[k] = walk_z[k]; }
        else orc_walk_normals(E->seed, E->id_offset + local_i, e->episode_idx, e->step_count, z);
        const double sdt = sqrt(E->dt);
        const double scale = orc_difficulty_scale(e->difficulty);
        for (int k = 0; k < 3; ++k) {
            double delta = z[k] * 0.1 * sdt;
            delta *= scale * max_rates[k];
            e->rate_cmd[k] = np_clip(e->rate_cmd[k] + delta, -max_rates[k], max_rates[k]);
        }
    } else if (e->command_type == ORC_CMD_SINE) {
        const double t = e->env_time, freq = e->sched[3];
        const double sn = sin(2 * ORC_PI * freq * t);
        for (int k = 0; k < 3; ++k) e->rate_cmd[k] = e->sched[k] * sn;
    }
}

int orc_env_reset(const orc_aircraft_params* P, const orc_env_config* E, orc_env_buffers* B,
                  const uint8_t* mask, int nthreads) {
    (void)P;
    const int nt = resolve_threads(nthreads);
    (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (int64_t i = 0; i < B->n; ++i) {
        if (mask && !mask[i]) continue;
        env_t e;
        env_gather(B, i, &e);
        env_reset_one(E, i, &e);
        env_scatter(B, i, &e);
        env_observe(&e, B->obs + i * OBS);
    }
    return 0;
}

int orc_env_step(const orc_aircraft_params* P, const orc_controller_params* C, const orc_env_config* E,
                 orc_env_buffers* B, const float* action_in, int nthreads) {
    double h;
    const int K = substep_count(P, E->dt, E->dt_physics, &h);
    if (K < 0) return -1;
    const int nt = resolve_threads(nthreads);
    (void)nt;
    double stats[16];
    for (int k = 0; k < 16; ++k) stats[k] = 0.0;
    const int64_t n = B->n;

#pragma omp parallel for num_threads(nt) schedule(static) reduction(+ : stats[:16])
    for (int64_t i = 0; i < n; ++i) {
        env_t e;
        env_gather(B, i, &e);

        /* ---- controller in front of the env ---- */
        float act[4];
        double residual_bonus = 0.0;
        if (E->controller == ORC_CTRL_EXTERNAL) {
            for (int k = 0; k < 4; ++k) act[k] = action_in[i * 4 + k];
        } else if (E->controller == ORC_CTRL_RATE_PID || E->controller == ORC_CTRL_RESIDUAL) {
            /* pid_demonstrations.py:55-78 / residual_rate_env.py:112-135 */
            double surf[4];
            rate_agent(C, e.rate, e.rate_cmd, E->pid_throttle, e.s, E->dt, surf);
            act[0] = (float)surf[1]; act[1] = (float)surf[0]; act[2] = (float)surf[2]; act[3] = (float)surf[3];
            if (E->controller == ORC_CTRL_RESIDUAL) {
                /* residual_rate_env.py:138-155: float32 arithmetic throughout */
                const float scale = (float)E->residual_scale;
                float res[4];
                for (int k = 0; k < 4; ++k) res[k] = action_in[i * 4 + k];
                for (int k = 0; k < 4; ++k) act[k] = act[k] + res[k] * scale;
                for (int k = 0; k < 3; ++k) act[k] = (act[k] < -1.0f) ? -1.0f : ((act[k] > 1.0f) ? 1.0f : act[k]);
                act[3] = (act[3] < 0.0f) ? 0.0f : ((act[3] > 1.0f) ? 1.0f : act[3]);
                const float rm = (res[0] * res[0] + res[1] * res[1]) + res[2] * res[2];
                /* 0.05 * (1.0 - rm/3.0): Python floats are weak under NEP 50 -> float32 */
                residual_bonus = (double)(0.05f * (1.0f - rm / 3.0f));
            }
        } else { /* ORC_CTRL_ATTITUDE_PID: AttitudeAgent in front of the env, same composition */
            double surf[4];
            attitude_agent(C, e.ang, e.rate, e.att_cmd, e.att_cmd[3], e.s, E->dt, surf);
            act[0] = (float)surf[1]; act[1] = (float)surf[0]; act[2] = (float)surf[2]; act[3] = (float)surf[3];
        }

        /* ---- RateControlEnv.step, rate_env.py:212-300 ---- */
        /* :225 clip to the float32 Box */
        for (int k = 0; k < 3; ++k) act[k] = (act[k] != act[k]) ? act[k] : ((act[k] < -1.0f) ? -1.0f : ((act[k] > 1.0f) ? 1.0f : act[k]));
        act[3] = (act[3] != act[3]) ? act[3] : ((act[3] < 0.0f) ? 0.0f : ((act[3] > 1.0f) ? 1.0f : act[3]));
        /* :228-237 */
        double cin[4] = {(double)act[1], (double)act[0], (double)act[2], (double)act[3]}, c[4];
        clip_controls(cin, c);
        backend_step(P, e.s, &e.time, c, K, h);
        orc_derived(e.s, e.derived);
        /* :241-245 */
        e.env_time += E->dt;
        e.step_count += 1;
        double wz[3];
        if (B->walk_z) for (int k = 0; k < 3; ++k) wz[k] = B->walk_z[(int64_t)k * n + i];
        env_update_command(E, i, &e, B->walk_z ? wz : NULL);
        /* :248-252 */
        const double err[3] = {e.rate_cmd[0] - e.s[9], e.rate_cmd[1] - e.s[10], e.rate_cmd[2] - e.s[11]};
        const double airspeed = e.derived[0], altitude = e.derived[1], roll = e.s[6], pitch = e.s[7];

        /* RateTrackingReward.compute, rewards.py:48-137 */
        const double tracking_error = (np_sq(err[0]) + np_sq(err[1]) + np_sq(err[2])) / 3.0;
        const double r_tracking = -0.5 * tracking_error;
        double r_smoothness;
        if (e.step_count == 1) {
            /* prev_action is the float64 array of rate_env.py:193 -> float64 arithmetic */
            double cc = 0.0;
            for (int k = 0; k < 3; ++k) { const double d = (double)act[k] - (double)e.prev_action[k]; cc += d * d; }
            r_smoothness = -0.01 * cc;
        } else {
            /* both float32 (rate_env.py:282) -> float32 arithmetic, weak Python scalar (NumPy >= 2) */
            float cc = 0.0f;
            for (int k = 0; k < 3; ++k) { const float d = act[k] - e.prev_action[k]; cc += d * d; }
            r_smoothness = (double)(-0.01f * cc);
        }
        const double roll_stability = exp(-fabs(roll) / np_radians(45.0));
        const double pitch_stability = exp(-fabs(pitch) / np_radians(30.0));
        const double airspeed_stability = np_clip((airspeed - 8.0) / 12.0, 0.0, 1.0);
        const double altitude_stability = np_clip((altitude - 10.0) / 90.0, 0.0, 1.0);
        const double stability_score = (roll_stability + pitch_stability + airspeed_stability + altitude_stability) / 4.0;
        const double r_stability = 0.3 * stability_score;
        double osc = 0.0;
        for (int k = 0; k < 3; ++k) {
            const int change = (np_sign(err[k]) != np_sign(e.prev_err[k])) && (fabs(e.prev_err[k]) > 0.01);
            e.sign_changes[k] = 0.9 * e.sign_changes[k] + (change ? 1.0 : 0.0);
            osc += e.sign_changes[k];
            e.prev_err[k] = err[k];
        }
        const double r_oscillation = -0.1 * osc;
        const double r_survival = 1.0;
        double reward = r_tracking + r_smoothness + r_stability + r_oscillation + r_survival;

        /* SettlingTimeBonus.compute, rewards.py:168-221 */
        double settle_reward = 0.0;
        {
            int settled = 1;
            for (int k = 0; k < 3; ++k)
                settled = settled && (fabs(err[k]) < py_max(fabs(e.rate_cmd[k]) * 0.05, 0.05));
            if (settled) {
                e.settle_timer += E->dt;
                if (e.settle_timer >= 0.2) { e.is_settled = 1; settle_reward = 2.0 * E->dt; }
            } else {
                e.settle_timer = 0.0;
                e.is_settled = 0;
            }
        }
        reward += settle_reward;                                   /* :273 */
        for (int k = 0; k < 4; ++k) e.prev_action[k] = act[k];       /* :282 */

        /* _check_termination :437-460, truncation :286, crash penalty :289-294 */
        const int terminated = (altitude < 5.0) || (fabs(roll) > np_radians(120.0)) || (fabs(pitch) > np_radians(80.0)) || (airspeed < 8.0);
        const int truncated = e.step_count >= E->max_steps;
        double crash_penalty = 0.0;
        if (terminated && !truncated) { crash_penalty = -100.0; reward += crash_penalty; }
        reward += residual_bonus;                                    /* residual_rate_env.py:153-155 */
        e.ep_return += reward;

        float* obs = B->obs + i * OBS;
        env_observe(&e, obs);
        B->reward[i] = reward;
        B->terminated[i] = (uint8_t)terminated;
        B->truncated[i] = (uint8_t)truncated;
        if (B->components) {
            const double comp[8] = {r_tracking, r_smoothness, r_stability, r_oscillation, r_survival, settle_reward, crash_penalty, residual_bonus};
            for (int k = 0; k < 8; ++k) B->components[(int64_t)k * n + i] = comp[k];
        }
        stats[6] += r_tracking; stats[7] += r_smoothness; stats[8] += r_stability; stats[9] += r_oscillation;
        stats[10] += r_survival; stats[11] += settle_reward; stats[12] += crash_penalty; stats[13] += residual_bonus;
        stats[15] += 1.0;

        if (terminated || truncated) {
            stats[0] += 1.0; stats[1] += e.ep_return; stats[2] += e.ep_return * e.ep_return; stats[3] += (double)e.step_count;
            stats[4] += (terminated && !truncated) ? 1.0 : 0.0; stats[5] += truncated ? 1.0 : 0.0;
            if (E->auto_reset) {
                /* SB3 VecEnv: keep the last observation, reset, return the reset observation */
                if (B->terminal_obs) memcpy(B->terminal_obs + i * OBS, obs, OBS * sizeof(float));
                env_reset_one(E, i, &e);
                env_observe(&e, obs);
            }
        }
        env_scatter(B, i, &e);
    }
    if (B->stats) for (int k = 0; k < 16; ++k) B->stats[k] += stats[k];
    return K;
}
