/*
 * acb_oracle.h -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the parity checker for the CUDA kernels.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it; the product never does.
 *
 * Reference = lhwright13/Hybrid-Classical-and-Reinforcement-Learning-Aircraft-Controllers,
 * paths relative to the reference root.  Every function cites the lines it restates.
 *
 * Parity status: PINNED.  tests/golden/ holds trajectories produced by the unmodified reference
 * (numpy physics + the reference C++ PID compiled into oracle/_ref) with tests/golden/make_golden.py;
 * tests/test_oracle_golden.py checks this restatement against them.
 *
 * Arithmetic: physics / reward in IEEE double with the reference's operation order, PID in IEEE
 * float, no FMA contraction anywhere (-ffp-contract=off), except the three places where the
 * reference itself executes fused ops through BLAS (norm of a 3-vector, 3x3 mat-vec), which are
 * written with explicit fma().
 *
 * Layout: structure of arrays, field-major: X[field][n] with n contiguous.
 */
#ifndef ACB_ORACLE_H
#define ACB_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* simulation/simplified_6dof.py:31-117 (AircraftParams), field order preserved. */
typedef struct {
    double mass, inertia_xx, inertia_yy, inertia_zz;
    double wing_area, wing_span, chord;
    double cl_0, cl_alpha, cd_0, cd_alpha2;
    double cl_elevator, cm_elevator, cy_rudder, cn_rudder, cl_aileron;
    double cm_alpha, cn_beta, cl_beta;
    double damping_roll, damping_pitch, damping_yaw;
    double max_thrust, air_density, gravity;
    double min_airspeed_aero, min_u_velocity;
    double max_elevator_deflection, max_aileron_deflection, max_rudder_deflection; /* degrees */
    double thrust_zero_velocity;
    double max_velocity, max_angular_rate, max_pitch_angle, max_alpha;             /* deg/s, deg */
    double max_acceleration, max_angular_acceleration;
    double max_timestep, min_timestep;
} orc_aircraft_params;                                                              /* 39 doubles */

/* cpp/include/pid_controller.h:22-49 (PIDGains + PIDConfig) */
typedef struct {
    float kp, ki, kd;
    float output_min, output_max, integral_min, integral_max, derivative_filter_alpha;
} orc_pid_config;

/* cpp/include/pid_controller.h:123-133 (PIDController private state) */
typedef struct {
    float error, error_prev, integral, derivative, derivative_filtered, output;
} orc_pid_state;

/* controllers/types.py:294-350 (ControllerConfig) narrowed the way create_pid_config does
 * (controllers/utils/pid_utils.py:16-46) + the agents' cached limits. */
typedef struct {
    orc_pid_config rate[3];   /* roll_rate_gains, pitch_rate_gains, yaw_gains; out [-1,1]     rate_agent.py:41-50   */
    orc_pid_config angle[3];  /* roll_angle_gains, pitch_angle_gains, yaw_gains; out +-max rate  attitude_agent.py:43-64 */
    double max_roll_rate, max_pitch_rate, max_yaw_rate; /* rad/s   rate_agent.py:53-55 */
    double max_roll, max_pitch;                         /* rad     attitude_agent.py:70-71 */
    /* hsa_agent.py:38-108 + config_loader.py:40-109 */
    orc_pid_config heading, energy, balance;
    double max_bank_rad, baseline_throttle, load_factor_gain, max_pitch_cmd_rad;
    /* GuidanceConfig, config_loader.py:40-64 (waypoint_agent.py:33-74); angles in degrees as in the config */
    double max_bank_angle_deg;       /* hsa.max_bank_angle, used by pure pursuit  waypoint_agent.py:53,150 */
    double lookahead_time, lookahead_min, lookahead_max, proximity_scale_distance;
    double los_max_bank_deg, los_lead_angle_deg;
    double turn_threshold_angle_deg, turn_threshold_distance, max_speed_reduction, min_speed;
} orc_controller_params;

void orc_default_aircraft_params(orc_aircraft_params* p);
void orc_default_controller_params(orc_controller_params* c);

/* ---- scalar kernels (one aircraft) ---- */
void   orc_dynamics(const orc_aircraft_params* P, const double s[12], const double ctrl[4], double out[12]);
void   orc_rk4_step(const orc_aircraft_params* P, double s[12], double* time, const double ctrl[4], double dt);
void   orc_derived(const double s[12], double out[4]); /* airspeed, altitude, ground_speed, heading */
float  orc_pid_compute(const orc_pid_config* c, orc_pid_state* st, float sp, float meas, float dt);
double orc_wrap_angle(double a);

/* ---- batched SoA drivers (OpenMP over aircraft; nthreads<=0 -> all cores) ---- */

/* SimulationAircraftBackend.set_controls + step (simulation_backend.py:82-101):
 * state[13][n] (12 states + time), ctrl_in[4][n] in ControlSurfaces.to_array order
 * (elevator, aileron, rudder, throttle), clipped copy written to ctrl_held[4][n] if non-null,
 * derived[4][n] (airspeed, altitude, ground_speed, heading) if non-null.
 * K = max(1,int(dt/dt_physics)), h = dt/K.  Returns K, or <0 on bad dt (ValueError in the reference). */
int orc_step_surface(const orc_aircraft_params* P, int64_t n, double* state, const double* ctrl_in,
                     double* ctrl_held, double* derived, double dt, double dt_physics, int nthreads);

/* level 4 = RateAgent, 3 = AttitudeAgent (controllers/types.py:9-24).
 * cmd[4][n]: level 4: roll_rate,pitch_rate,yaw_rate,throttle; level 3: roll,pitch,yaw(NaN=None),throttle.
 * pid[18][n] float: ((loop*3+axis)*3 + {integral,error_prev,derivative_filtered}), loop 0=rate 1=angle.
 * Computes surfaces from the CURRENT state (agent.compute_action(cmd,state,dt_ctrl)), then
 * backend.set_controls + backend.step(dt).  surfaces_out[4][n] (to_array order) if non-null. */
int orc_step_cascade(const orc_aircraft_params* P, const orc_controller_params* C, int level, int64_t n,
                     double* state, const double* cmd, float* pid, double* surfaces_out, double* derived,
                     double dt_ctrl, double dt, double dt_physics, int nthreads);

/* level 2 = HSAAgent (hsa_agent.py:119-231): cmd[3][n] = heading, speed, altitude.
 * pid_outer[9][n] float: (k*3 + {integral,error_prev,derivative_filtered}), k 0=heading 1=energy 2=balance. */
int orc_step_hsa(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, double* state, const double* cmd,
                 float* pid, float* pid_outer, double* surfaces_out, double* derived, double dt_ctrl, double dt,
                 double dt_physics, int nthreads);

/* level 1 = MissionPlanner.update + WaypointAgent.compute_action + sim.step, the loop of
 * examples/03_waypoint_square_demo.py:147-209.  waypoints[n_wp][4] = north, east, down, speed (NaN = keep the current
 * airspeed), one mission shared by the fleet; wp_index[n] = current waypoint (n_wp = mission complete: the aircraft is
 * left untouched, as the demo stops stepping); loop != 0 wraps to waypoint 0 instead.  guidance: 0 LOS, 1 pure pursuit. */
int orc_step_waypoint(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, double* state,
                      const double* waypoints, int n_wp, int32_t* wp_index, int guidance, double acceptance_radius, int loop,
                      float* pid, float* pid_outer, double* surfaces_out, double* derived, double* hsa_cmd_out,
                      double dt_ctrl, double dt, double dt_physics, int nthreads);

/* the same with the fleet's mission statistics accumulated into mission_stats[16] (nullable; see acb_oracle.c) */
int orc_step_waypoint_stats(const orc_aircraft_params* P, const orc_controller_params* C, int64_t n, double* state,
                            const double* waypoints, int n_wp, int32_t* wp_index, int guidance, double acceptance_radius, int loop,
                            float* pid, float* pid_outer, double* surfaces_out, double* derived, double* hsa_cmd_out,
                            double dt_ctrl, double dt, double dt_physics, int nthreads, double* mission_stats);

/* ---- rate-tracking environment (learned_controllers/envs/rate_env.py) ---- */

enum { ORC_CTRL_EXTERNAL = 0, ORC_CTRL_RATE_PID = 1, ORC_CTRL_ATTITUDE_PID = 2, ORC_CTRL_RESIDUAL = 3 };
enum { ORC_CMD_STEP = 0, ORC_CMD_RAMP = 1, ORC_CMD_SINE = 2, ORC_CMD_RANDOM = 3 };

typedef struct {
    int64_t n;
    /* physics */
    double* state;        /* [13][n] */
    double* derived;      /* [4][n]  */
    /* controller memory */
    float*  pid;          /* [18][n] */
    double* att_cmd;      /* [4][n] roll,pitch,yaw(NaN=None),throttle  (ORC_CTRL_ATTITUDE_PID) */
    /* env memory */
    double* rate_cmd;     /* [3][n]  rate_env.py:92 */
    float*  prev_action;  /* [4][n]  rate_env.py:93,193,282 (env order: aileron, elevator, rudder, throttle) */
    double* rew_mem;      /* [7][n]  prev_errors[3], sign_changes[3], settle_timer   rewards.py:44-46,187 */
    double* sched;        /* [7][n]  ramp end[3] | sine amplitudes[3], sine frequency   rate_env.py:96-97 */
    double* env_time;     /* [n]     rate_env.py:87,241 */
    int32_t* step_count;  /* [n]     rate_env.py:89 */
    uint8_t* is_settled;  /* [n]     rewards.py:188 */
    uint8_t* command_type;/* [n]     rate_env.py:63 */
    float*  difficulty;   /* [n]     difficulty_scale, generators.py:41-45 */
    uint32_t* episode_idx;/* [n]     Philox counter word */
    double* ep_return;    /* [n] */
    /* per-step outputs */
    float*  obs;          /* [n][18] row-major  rate_env.py:374-408 */
    double* reward;       /* [n] */
    uint8_t* terminated;  /* [n] */
    uint8_t* truncated;   /* [n] */
    float*  terminal_obs; /* [n][18] written for finished aircraft only (SB3 terminal_observation) */
    double* components;   /* [8][n] tracking,smoothness,stability,oscillation,survival,settle_bonus,crash_penalty,residual_bonus; nullable */
    double* stats;        /* [16] accumulated over calls; nullable */
    /* ORACLE ONLY (no counterpart in acb200_env): standard normals [3][n] that replace this step's Philox/Box-Muller
     * draw of the random-walk command, so that the reference's own MT19937 `randn(3)` values (generators.py:159) can
     * be replayed through the restatement (tests/golden/rate_env_random.npz); NULL = the Philox stream. */
    const double* walk_z;
} orc_env_buffers;

typedef struct {
    double dt;              /* 0.02  rate_env.py:42 */
    double dt_physics;      /* 0.001 rate_env.py:69 */
    int32_t max_steps;      /* int(episode_length/dt) rate_env.py:88 */
    int32_t controller;     /* ORC_CTRL_* */
    int32_t auto_reset;     /* 1: SB3 VecEnv semantics (reset finished envs, return reset obs) */
    double pid_throttle;    /* 0.6: residual_rate_env.py:122, pid_demonstrations.py:68 */
    double residual_scale;  /* 0.3: residual_rate_env.py:33 */
    uint64_t seed;          /* Philox key = (seed, global aircraft id) */
    int64_t id_offset;      /* global id of local aircraft 0 (multi-GPU sharding) */
} orc_env_config;

/* RateControlEnv.reset for every aircraft with mask[i]!=0 (mask NULL = all).  Initial state and
 * command are drawn with Philox4x32-10 (rate_env.py:151-210, generators.py:47-72,190-214). */
int orc_env_reset(const orc_aircraft_params* P, const orc_env_config* E, orc_env_buffers* B,
                  const uint8_t* mask, int nthreads);

/* RateControlEnv.step (rate_env.py:212-300) for all n aircraft.  action[n][4] float, env order. */
int orc_env_step(const orc_aircraft_params* P, const orc_controller_params* C, const orc_env_config* E,
                 orc_env_buffers* B, const float* action, int nthreads);

/* The three standard normals the random-walk command of (seed, global id, episode, step_count) consumes
 * (block 16 + step_count of the aircraft's Philox stream, Box-Muller); exposed for the distribution test. */
void orc_walk_normals(uint64_t seed, int64_t global_id, uint32_t episode, int32_t step_count, double z[3]);
/* difficulty_scale as the reference holds it (a Python float 0.3 / 0.5 / 0.7, generators.py:41-45) recovered from the
 * float32 the fleet stores: the nearest multiple of 1e-6. */
double orc_difficulty_scale(float difficulty);

/* Philox4x32-10 block, exposed for the known-answer test. */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

/* NoisySensorInterface (interfaces/sensor.py:137-263); same field order as acb200_sensor_config */
typedef struct {
    int32_t enabled, _pad;
    double imu_gyro_stddev, imu_accel_stddev, gps_position_stddev, gps_velocity_stddev, airspeed_stddev, altitude_stddev,
        attitude_stddev, gyro_bias_walk_stddev, accel_bias_walk_stddev;
    uint64_t seed;
    int64_t id_offset;
} orc_sensor_config;
void orc_sensor_normals(uint64_t seed, int64_t global_id, uint32_t tick, double z[20]);
/* state [13][n], derived [4][n], bias [6][n] (in/out), z_ext [20][n] or NULL (Philox stream), m_state [13][n],
 * m_derived [4][n]; -1 for a bad stddev */
int orc_sensor_update(const orc_sensor_config* S, int64_t n, const double* state, const double* derived, double* bias,
                      uint32_t tick, const double* z_ext, double* m_state, double* m_derived);

#ifdef __cplusplus
}
#endif
#endif
