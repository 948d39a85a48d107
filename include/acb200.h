/*
 * acb200.h -- C ABI of the B200-native batched flight-control path.
 *
 * One shared library (libacb200.so, built from
 * hybrid-classical-and-reinforcement-learning-aircraft-controllers_b200/csrc) replaces, for N
 * independent aircraft at once, the reference's per-aircraft Python/C++ hot path:
 *
 *   Simplified6DOF.step / _dynamics / get_state      simulation/simplified_6dof.py:228-331,333-503
 *   SimulationAircraftBackend.step / set_controls     simulation/simulation_backend.py:82-112
 *   PIDController / MultiAxisPIDController::compute   cpp/src/pid_controller.cpp:24-69,100-111
 *   RateAgent / AttitudeAgent.compute_action          controllers/rate_agent.py:65-124, attitude_agent.py:86-154
 *   RateControlEnv.reset / step (+ rewards, sampler)  learned_controllers/envs/rate_env.py:151-300,
 *                                                     envs/rewards.py, data/generators.py
 *   ResidualRateControlEnv.step                       learned_controllers/envs/residual_rate_env.py:99-157
 *
 * (paths relative to the reference repository root).  Plain pointers and sizes only; no torch, no
 * C++ types.  Every buffer is CALLER-OWNED device memory (e.g. torch CUDA tensors: pass
 * tensor.data_ptr()); the library allocates nothing per aircraft and never frees caller memory.
 * All work is enqueued on the context's stream and is asynchronous; no hidden device synchronise
 * (the *_host entry points, which move host buffers, synchronise that stream before returning).
 * A context is bound to one device and is not thread-safe.
 *
 * Layout: structure of arrays, field-major X[field][n], n contiguous.  `real` is double when
 * precision == 64 (parity build: fp64 physics/reward, fp32 PID, no FMA contraction) and float when
 * precision == 32 (throughput build).
 *
 * Errors: every function returns 0 or a negative ACB200_E* code and never throws or exits;
 * acb200_last_error() gives the message.  The Python adapters map ACB200_EVALUE to ValueError
 * (the reference's bad-dt / bad-parameter exceptions, simplified_6dof.py:119-143,241-245).
 */
#ifndef ACB200_H
#define ACB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ACB200_VERSION "0.1.0"

enum {
    ACB200_OK = 0,
    ACB200_EVALUE = -1,   /* invalid argument value (reference: ValueError) */
    ACB200_ECUDA = -2,    /* CUDA runtime error */
    ACB200_ENULL = -3,    /* required pointer is NULL */
    ACB200_ENOSUP = -4    /* combination not supported */
};

/* control levels, controllers/types.py:9-24 (ControlMode) */
enum { ACB200_LEVEL_WAYPOINT = 1, ACB200_LEVEL_HSA = 2, ACB200_LEVEL_ATTITUDE = 3, ACB200_LEVEL_RATE = 4, ACB200_LEVEL_SURFACE = 5 };
/* who drives the env's action */
enum { ACB200_CTRL_EXTERNAL = 0, ACB200_CTRL_RATE_PID = 1, ACB200_CTRL_ATTITUDE_PID = 2, ACB200_CTRL_RESIDUAL = 3 };
/* RateControlEnv command_type, rate_env.py:302-340 */
enum { ACB200_CMD_STEP = 0, ACB200_CMD_RAMP = 1, ACB200_CMD_SINE = 2, ACB200_CMD_RANDOM = 3 };

#define ACB200_STATE_FIELDS 13   /* N,E,D, u,v,w, roll,pitch,yaw, p,q,r, time   simplified_6dof.py:172,179 */
#define ACB200_CTRL_FIELDS 4     /* elevator, aileron, rudder, throttle          controllers/types.py:190-194 */
#define ACB200_DERIVED_FIELDS 4  /* airspeed, altitude, ground_speed, heading    simplified_6dof.py:306-319 */
#define ACB200_PID_FIELDS 18     /* ((loop*3+axis)*3 + {integral, error_prev, derivative_filtered}); loop 0 rate, 1 angle */
#define ACB200_OBS_DIM 18        /* rate_env.py:374-408 */
#define ACB200_STATS_FIELDS 16

/* simulation/simplified_6dof.py:31-117 (AircraftParams), same field order; always double. */
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
    double max_velocity, max_angular_rate, max_pitch_angle, max_alpha;             /* m/s, deg/s, deg, deg */
    double max_acceleration, max_angular_acceleration;
    double max_timestep, min_timestep;
} acb200_aircraft_params;

/* cpp/include/pid_controller.h:22-49 (PIDGains + PIDConfig) */
typedef struct {
    float kp, ki, kd;
    float output_min, output_max, integral_min, integral_max, derivative_filter_alpha;
} acb200_pid_config;

/* cpp/include/pid_controller.h:123-133 (PIDController state) */
typedef struct {
    float error, error_prev, integral, derivative, derivative_filtered, output;
} acb200_pid_state;

/* controllers/types.py:294-350 (ControllerConfig) after create_pid_config (pid_utils.py:16-46) */
typedef struct {
    acb200_pid_config rate[3];    /* roll_rate_gains, pitch_rate_gains, yaw_gains; output [-1,1]   rate_agent.py:41-50 */
    acb200_pid_config angle[3];   /* roll_angle_gains, pitch_angle_gains, yaw_gains; output +-max rate  attitude_agent.py:43-64 */
    double max_roll_rate, max_pitch_rate, max_yaw_rate;   /* rad/s */
    double max_roll, max_pitch;                           /* rad   */
    /* HSAAgent, controllers/hsa_agent.py:38-108 with HSAConfig / TECSConfig (config_loader.py:66-83) */
    acb200_pid_config heading;    /* out +-max_bank   */
    acb200_pid_config energy;     /* out +-0.5 throttle */
    acb200_pid_config balance;    /* out +-max_pitch_command */
    double max_bank, baseline_throttle, load_factor_gain, max_pitch_cmd;   /* rad, -, -, rad (10 deg, hsa_agent.py:108) */
    /* WaypointAgent, controllers/waypoint_agent.py:33-74 with GuidanceConfig (config_loader.py:40-64); degrees as configured */
    double max_bank_angle_deg;
    double lookahead_time, lookahead_min, lookahead_max, proximity_scale_distance;
    double los_max_bank_deg, los_lead_angle_deg;
    double turn_threshold_angle_deg, turn_threshold_distance, max_speed_reduction, min_speed;
} acb200_controller_params;

/* guidance_type of WaypointAgent (waypoint_agent.py:121-186) */
enum { ACB200_GUIDANCE_LOS = 0, ACB200_GUIDANCE_PP = 1 };
#define ACB200_OUTER_PID_FIELDS 9   /* (k*3 + {integral, error_prev, derivative_filtered}); k 0 heading, 1 energy, 2 balance */

/* MissionPlanner (controllers/mission_planner.py:128-205) for the whole fleet: one waypoint list, one index per aircraft. */
typedef struct {
    const void* waypoints;     /* real[n_waypoints][4]: north, east, down, speed (NaN = None: keep the current airspeed) */
    int32_t n_waypoints;
    int32_t guidance;          /* ACB200_GUIDANCE_* */
    int32_t* wp_index;         /* [n] current waypoint; == n_waypoints: mission complete, the aircraft is no longer stepped */
    int32_t loop;              /* != 0: after the last waypoint continue with waypoint 0 */
    int32_t _pad;
    double acceptance_radius;  /* 3-D distance, mission_planner.py:173-186 */
    double* stats;             /* double[ACB200_MISSION_STATS_FIELDS], device, accumulated over calls (warp shuffle -> one atomic per
                                  block and field); nullable.  [0] aircraft-steps flown, [1] waypoints reached (what
                                  MissionPlanner.waypoints_reached counts, summed over the fleet), [2] laps closed (looping missions),
                                  [3] missions completed, [4] sum |cross-track error| (horizontal distance from the leg previous ->
                                  active waypoint) and [5] its squares, [6] sum |altitude - waypoint altitude|, [7] sum 3-D distance to
                                  the active waypoint (MissionPlanner.get_distance_to_current_waypoint), [8] sum airspeed, [9] steps
                                  below 5 m (the env's crash floor); [10..15] reserved.  Contributions are taken from the state the
                                  guidance sees, before the physics step.  One all-reduce per report interval sums it over ranks. */
} acb200_mission;
#define ACB200_MISSION_STATS_FIELDS 16

/* The aircraft fleet: what N SimulationAircraftBackend objects hold. */
typedef struct {
    int64_t n;
    int32_t precision;     /* 64 or 32 */
    int32_t _pad;
    void* state;           /* real[13][n] */
    void* controls;        /* real[4][n]  controls held by the simulator (clipped)   simplified_6dof.py:221-226; nullable */
    void* derived;         /* real[4][n]  airspeed, altitude, ground_speed, heading = what get_state() derives from the state
                            * (simplified_6dof.py:295-331).  Every entry point that writes the state refreshes it; the HSA and
                            * waypoint steps READ it as the agents' measurements instead of deriving it again, so a caller that
                            * writes `state` itself must go through acb200_fleet_reset(mask, init_state) -- or pass NULL here,
                            * then they derive it from the state on every step.  nullable */
    float* pid;            /* float[18][n] PID memory of the rate and angle loops; required for levels 3,4 */
} acb200_fleet;

/* What N RateControlEnv objects hold (plus the fleet they fly). */
typedef struct {
    acb200_fleet fleet;
    void* att_cmd;          /* real[4][n]   roll,pitch,yaw,throttle for ACB200_CTRL_ATTITUDE_PID; yaw NaN = None (->0), +inf = hold current yaw */
    void* rate_cmd;         /* real[3][n]   rate_env.py:92 */
    float* prev_action;     /* float[4][n]  aileron, elevator, rudder, throttle   rate_env.py:93,193,282 */
    void* rew_mem;          /* real[7][n]   prev_errors[3], sign_changes[3], settle_timer   rewards.py:44-46,187 */
    void* sched;            /* real[7][n]   ramp end[3] | sine amplitudes[3], sine frequency   rate_env.py:96-97 */
    void* env_time;         /* real[n]      rate_env.py:87,241 */
    int32_t* step_count;    /* [n]          rate_env.py:89 */
    uint8_t* is_settled;    /* [n]          rewards.py:188 */
    uint8_t* command_type;  /* [n]          ACB200_CMD_*; the curriculum is a write to this and `difficulty` */
    float* difficulty;      /* float[n]     difficulty_scale 0.3/0.5/0.7   generators.py:41-45 */
    uint32_t* episode_idx;  /* [n]          Philox counter word, +1 per reset */
    void* ep_return;        /* real[n]      running episode return (SB3 Monitor "r") */
    /* outputs of reset/step */
    float* obs;             /* float[n][18] row-major */
    void* reward;           /* real[n] */
    uint8_t* terminated;    /* [n] */
    uint8_t* truncated;     /* [n] */
    float* terminal_obs;    /* float[n][18] written for finished aircraft only (SB3 terminal_observation); nullable */
    void* components;       /* real[8][n]  tracking, smoothness, stability, oscillation, survival, settle_bonus,
                                            crash_penalty, residual_bonus (info["reward_components"]); nullable */
    double* stats;          /* double[16] accumulated: episodes, sum return, sum return^2, sum length, crashes, truncations,
                               sum of the 8 components, non-finite repairs, aircraft-steps; nullable */
} acb200_env;

typedef struct {
    double dt;              /* 0.02   rate_env.py:42 */
    double dt_physics;      /* 0.001  rate_env.py:69 */
    int32_t max_steps;      /* int(episode_length/dt)  rate_env.py:88 */
    int32_t controller;     /* ACB200_CTRL_* */
    int32_t auto_reset;     /* 1 = SB3 VecEnv semantics */
    int32_t _pad;
    double pid_throttle;    /* 0.6  residual_rate_env.py:122 */
    double residual_scale;  /* 0.3  residual_rate_env.py:33 */
    uint64_t seed;          /* Philox key */
    int64_t id_offset;      /* global id of local aircraft 0: results do not depend on how the fleet is sharded */
} acb200_env_config;

typedef struct acb200_ctx acb200_ctx;

/* ---- context ---- */
const char* acb200_version(void);
int acb200_create(int device, acb200_ctx** out);
void acb200_destroy(acb200_ctx* ctx);
const char* acb200_last_error(const acb200_ctx* ctx);
int acb200_set_stream(acb200_ctx* ctx, void* cuda_stream);          /* cudaStream_t; NULL = legacy default stream */
int acb200_synchronize(acb200_ctx* ctx);

/* ---- parameters (AircraftParams.__post_init__ validation -> ACB200_EVALUE) ---- */
void acb200_default_aircraft_params(acb200_aircraft_params* p);       /* 'rc_plane'  simulation_backend.py:74-76 */
void acb200_cessna_aircraft_params(acb200_aircraft_params* p);        /* 'cessna'    simulation_backend.py:64-73 */
void acb200_default_controller_params(acb200_controller_params* c);
int acb200_set_aircraft_params(acb200_ctx* ctx, const acb200_aircraft_params* p);
int acb200_set_controller_params(acb200_ctx* ctx, const acb200_controller_params* c);
/* AircraftParams.__post_init__ (simplified_6dof.py:119-143) alone: 0 or ACB200_EVALUE with the reference's message in
 * acb200_last_error; the context's parameters are not touched. */
int acb200_validate_aircraft_params(acb200_ctx* ctx, const acb200_aircraft_params* p);
int acb200_get_aircraft_params(const acb200_ctx* ctx, acb200_aircraft_params* out);
int acb200_get_controller_params(const acb200_ctx* ctx, acb200_controller_params* out);
/* Parameter groups: group g of n_groups owns aircraft [bounds[g], bounds[g+1]) (bounds[0] = 0,
 * non-decreasing, every bound but the last a multiple of 128) and flies with params[g] -- what the reference is with one
 * AircraftParams per env instance (simulation_backend.py:55-80).  Every env step entry point (acb200_env_step, _range,
 * _host) then launches one kernel per group it touches, each with its group's parameters in the constant bank; the
 * context's own parameters (acb200_set_aircraft_params) are neither used by those calls nor changed.  All groups are
 * validated first; on ACB200_EVALUE nothing changes.  n_groups = 0 removes the groups.  The fleet the groups describe is
 * the one whose n equals bounds[n_groups]; a step on a fleet of another size fails with ACB200_EVALUE. */
int acb200_set_param_groups(acb200_ctx* ctx, const acb200_aircraft_params* params, const int64_t* bounds, int32_t n_groups);

/* Per-AIRCRAFT parameters: every aircraft flies its own AircraftParams -- what n reference env
 * instances each constructed with its own params are (simulation_backend.py:55-80).  The table is caller-owned device
 * memory of acb200_param_table_words() * n values of `real` (float for precision 32, double for 64), field-major
 * table[word][n]; acb200_build_param_table validates params[0..count) (AircraftParams.__post_init__; for precision 32
 * also the throughput build's ranges), folds each exactly as the constant-bank block is folded and writes rows
 * [first, first + count) (synchronises the context's stream; a failed validation writes nothing).  After
 * acb200_set_param_table every env step entry point (acb200_env_step, _range, _host) runs the table variant of the kernel
 * (one aircraft per thread, block loaded once per env step into registers) for a fleet with fleet.n == n and the same
 * precision, anything else fails with ACB200_EVALUE; table NULL returns to the context's parameters.  Parameter groups and
 * a table cannot be active together.  The constant-bank kernels stay the default and the faster path. */
int acb200_param_table_words(void);
int acb200_build_param_table(acb200_ctx* ctx, const acb200_aircraft_params* params, int64_t count, int64_t first, int64_t n,
                             int32_t precision, void* table);
int acb200_set_param_table(acb200_ctx* ctx, const void* table, int64_t n, int32_t precision);

/* ---- physics ---- */
/* Simplified6DOF.reset for aircraft with mask[i] != 0 (mask NULL = all).  init_state real[13][n] or
 * NULL for the default level flight at 100 m / 20 m/s (simplified_6dof.py:197-204).  Zeroes their PID
 * memory and held controls, refreshes derived. */
int acb200_fleet_reset(acb200_ctx* ctx, const acb200_fleet* f, const uint8_t* mask, const void* init_state);

/* backend.set_controls(ctrl_in) then backend.step(dt): K = max(1,int(dt/dt_physics)) RK4 substeps of
 * h = dt/K.  ctrl_in real[4][n].  h outside (min_timestep, max_timestep] -> ACB200_EVALUE. */
int acb200_step_surface(acb200_ctx* ctx, const acb200_fleet* f, const void* ctrl_in, double dt, double dt_physics);

/* surfaces = agent(level).compute_action(cmd, state, dt_ctrl); backend.set_controls; backend.step(dt).
 * level 4: cmd real[4][n] = roll_rate, pitch_rate, yaw_rate, throttle
 * level 3: cmd real[4][n] = roll, pitch, yaw (NaN = None, +inf = hold current yaw), throttle
 * surfaces_out real[4][n] nullable. */
int acb200_step_cascade(acb200_ctx* ctx, const acb200_fleet* f, int level, const void* cmd, double dt_ctrl,
                        double dt, double dt_physics, void* surfaces_out);

/* level 2: surfaces = HSAAgent.compute_action(cmd, state, dt_ctrl) (hsa_agent.py:119-231); set_controls; step(dt).
 * cmd real[3][n] = heading, speed, altitude.  pid_outer float[9][n].  f->pid holds the inner attitude/rate loops. */
int acb200_step_hsa(acb200_ctx* ctx, const acb200_fleet* f, float* pid_outer, const void* cmd, double dt_ctrl, double dt,
                    double dt_physics, void* surfaces_out);
/* level 1: the loop body of examples/03_waypoint_square_demo.py:147-209 for every aircraft: MissionPlanner.update,
 * WaypointAgent.compute_action (waypoint_agent.py:81-242), set_controls, step(dt).  hsa_cmd_out real[3][n] nullable. */
int acb200_step_waypoint(acb200_ctx* ctx, const acb200_fleet* f, float* pid_outer, const acb200_mission* m, double dt_ctrl,
                         double dt, double dt_physics, void* surfaces_out, void* hsa_cmd_out);

/* ---- rate-tracking environment ---- */
int acb200_env_reset(acb200_ctx* ctx, const acb200_env* e, const acb200_env_config* cfg, const uint8_t* mask);
/* action float[n][4] (aileron, elevator, rudder, throttle); may be NULL unless controller is EXTERNAL/RESIDUAL.
 * precision 32: one launch of the two-aircraft-per-thread kernel on packed FP32 instructions
 * (csrc/kernels_pair_f32.cuh) when n and first are even and the buffers 8-byte aligned, else (and for precision 64) of the
 * one-aircraft-per-thread kernel; the two compute the same bits.  Environment: ACB200_ENV_PAIR=0 forces the latter. */
int acb200_env_step(acb200_ctx* ctx, const acb200_env* e, const acb200_env_config* cfg, const float* action);
/* The same step for aircraft [first, first + count) only (first a multiple of 128), with the parameters the context holds
 * NOW: per-group parameter randomisation is  for each group: acb200_set_aircraft_params(...); acb200_env_step_range(...).
 * The reference has one global AircraftParams per env instance (simulation_backend.py:55-80); a group here is what N
 * env instances constructed with the same params are there.  action indexes the whole fleet (float[n][4]). */
int acb200_env_step_range(acb200_ctx* ctx, const acb200_env* e, const acb200_env_config* cfg, const float* action,
                          int64_t first, int64_t count);
/* Same step with HOST buffers (pinned for speed), what SubprocVecEnv.step_wait hands to the trainer
 * (learned_controllers/utils/training_utils.py:60-69): copies action_host (float[n][4], nullable) and att_cmd_host
 * (real[4][n], nullable: the per-step attitude command of ACB200_CTRL_ATTITUDE_PID) in, obs / reward / terminated /
 * truncated out, pipelined in chunks over three internal streams so that the two PCIe directions and the kernel
 * overlap, and returns when the host buffers are complete.  reward_host is float[n].  action_dev: caller-owned
 * device scratch of n*5 floats (needed when action_host or reward_host is given).  Every output pointer may be NULL: that
 * output is then not copied (the device buffers of `e` still hold it) -- with the controller in the kernel a trainer that
 * only wants reward + flags moves 6 B per aircraft instead of 78.  acb200_env_gather_done returns the finished aircraft. */
int acb200_env_step_host(acb200_ctx* ctx, const acb200_env* e, const acb200_env_config* cfg,
                         const float* action_host, const void* att_cmd_host, float* action_dev, float* obs_host,
                         float* reward_host, uint8_t* terminated_host, uint8_t* truncated_host);
/* The aircraft that finished an episode in the LAST step (terminated | truncated), as a compact list in ascending id
 * order: ids_out int64[capacity], their terminal observations (terminal_obs_out float[capacity][18], nullable; what SB3's
 * VecEnv reports as infos[i]["terminal_observation"]) and the observations of the episodes that started (reset_obs_out
 * float[capacity][18], nullable); *count_out = how many finished (may exceed capacity: only the first `capacity` are
 * written).  All pointers device memory; three small launches on the context's stream; about 0.2 % of the fleet per step
 * at 500-step episodes, so a trainer whose controller runs in the kernel never moves the (n, 18) matrix to the host. */
int acb200_env_gather_done(acb200_ctx* ctx, const acb200_env* e, int64_t capacity, int64_t* ids_out, float* terminal_obs_out,
                           float* reset_obs_out, int64_t* count_out);
/* acb200_env_gather_done without any allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same results (bit for bit) as acb200_env_gather_done; the per-tile scratch is a file-scope
 * device array of 64 KiB (csrc/envtools.cu) that every GPU loading the library holds until exit, which limits the fleet to
 * acb200_noalloc_max_aircraft() = 16 777 216 aircraft -- a larger fleet.n returns ACB200_EVALUE with a message naming the
 * limit.  Calls no cudaMalloc / cudaFree and no allocating library routine.  The array is shared by every call on a device
 * and keeps the last call's contents (each call overwrites what it reads, nothing needs clearing); calls that OVERLAP on
 * the GPU -- two streams, or two contexts on one device -- race on it: issue them on one stream or order them with events. */
int acb200_env_gather_done_noalloc(acb200_ctx* ctx, const acb200_env* e, int64_t capacity, int64_t* ids_out,
                                   float* terminal_obs_out, float* reset_obs_out, int64_t* count_out);
int64_t acb200_noalloc_max_aircraft(void);
/* number of chunks acb200_env_step_host cuts the fleet into; 0 = automatic (about 128k aircraft per chunk) */
int acb200_set_host_chunks(acb200_ctx* ctx, int chunks);

/* ---- the one collective of the path ----------------------------------------------------------------
 * The aircraft are independent, so a fleet sharded over G GPUs (one process and one context per GPU) exchanges nothing per
 * step; what crosses GPUs is the sum of the 16-double statistics vectors (acb200_env.stats, acb200_mission.stats) once per
 * report interval: ncclAllReduce(SUM, float64) over NVLink / NVSwitch, latency-bound (128 bytes).  NCCL is resolved at run
 * time (dlopen: $ACB200_NCCL_LIB, else the copy already loaded in the process, else the loader's path); without it these
 * calls return ACB200_ENOSUP and everything else works.
 *   rank 0:      acb200_comm_unique_id(ctx, id)            128 bytes, carried to the other ranks by the caller's own plumbing
 *   every rank:  acb200_comm_init(ctx, world, rank, id)    collective (ncclCommInitRank)
 *   every rank:  acb200_episode_stats(ctx, stats, out, 16, 1)   out = sum over ranks, enqueued on the context's stream
 * do_allreduce = 0 copies stats to out (no communicator needed).  stats / out are device pointers; out may equal stats. */
#define ACB200_COMM_ID_BYTES 128
int acb200_comm_unique_id(acb200_ctx* ctx, void* id_out);
int acb200_comm_init(acb200_ctx* ctx, int world, int rank, const void* id);
int acb200_comm_destroy(acb200_ctx* ctx);
int acb200_comm_info(const acb200_ctx* ctx, int* world, int* rank, int* nccl_version);
int acb200_episode_stats(acb200_ctx* ctx, const double* stats, double* out, int32_t count, int32_t do_allreduce);

/* ---- sensor model ------------------------------------------------------------------------
 * NoisySensorInterface, interfaces/sensor.py:137-263: Gaussian noise on position, velocity, attitude, angular rate (plus a
 * slowly drifting gyro bias), airspeed and altitude; gyro / accelerometer biases take one random-walk step per update.
 * The reference draws from numpy's PCG64 Generator (sensor.py:189-190), which a counter-based generator cannot replay:
 * parity is exact in the arithmetic given the normals (oracle: orc_sensor_update) and distributional against the reference. */
typedef struct {
    int32_t enabled;                 /* 0: measured = truth, biases untouched   sensor.py:204-206 */
    int32_t _pad;
    double imu_gyro_stddev;          /* rad/s  0.01   sensor.py:176-182 */
    double imu_accel_stddev;         /* m/s^2  0.1    (configured and reported; AircraftState carries no acceleration) */
    double gps_position_stddev;      /* m      1.0  */
    double gps_velocity_stddev;      /* m/s    0.1  */
    double airspeed_stddev;          /* m/s    0.5  */
    double altitude_stddev;          /* m      0.5  */
    double attitude_stddev;          /* rad    0.01 */
    double gyro_bias_walk_stddev;    /* 1e-4 per update   sensor.py:230 */
    double accel_bias_walk_stddev;   /* 1e-3 per update   sensor.py:231 */
    uint64_t seed;                   /* Philox key */
    int64_t id_offset;               /* global id of local aircraft 0 */
} acb200_sensor_config;

void acb200_default_sensor_config(acb200_sensor_config* c);
/* sensor.update(true_state) for every aircraft: reads f->state / f->derived, writes measured_state real[13][n] (time copied)
 * and measured_derived real[4][n] (airspeed, altitude noisy; ground_speed, heading copied).  bias real[6][n] = gyro_bias[3],
 * accel_bias[3], read and advanced.  tick: the update counter (Philox counter word) -- the caller passes 0,1,2,...; the same
 * (seed, global id, tick) always yields the same noise.  Negative or non-finite stddev -> ACB200_EVALUE. */
int acb200_sensor_update(acb200_ctx* ctx, const acb200_fleet* f, const acb200_sensor_config* cfg, void* bias, uint32_t tick,
                         void* measured_state, void* measured_derived);
/* sensor.reset(): zero the biases of aircraft with mask[i] != 0 (NULL = all)   sensor.py:245-251 */
int acb200_sensor_reset(acb200_ctx* ctx, int64_t n, int32_t precision, void* bias, const uint8_t* mask);

/* ---- telemetry sampling ----------------------------------------------------------------------
 * What TelemetryLogger.log_state / log_surfaces keep per logged aircraft (visualization/logger.py:72-120), for k sampled
 * aircraft at once: out real[21][k] = state rows 0..12 | derived rows 0..3 | held controls rows 0..3 of aircraft
 * ids[j] (device int64[k]; an id outside [0, n) logs NaN).  The caller advances `out` through its own device ring and
 * writes the file (host mirror: telemetry.TelemetryRecorderB200, the reference's JSON / HDF5 layouts). */
#define ACB200_TELEMETRY_FIELDS 21
int acb200_telemetry_record(acb200_ctx* ctx, const acb200_fleet* f, const int64_t* ids, int64_t k, void* out);

/* ---- learned rate policy, actor forward (row a20) -------------------------------------------------------------
 * Replaces: model.predict(obs, state=lstm_state, deterministic=True) at controllers/learned_rate_agent.py:220-227
 * with the network of learned_controllers/networks/lstm_policy.py:13-164 (LSTMFeaturesExtractor + sb3-contrib's
 * stateful actor LSTM + pi-MLP [128,64] + action_net; or SimpleMLPPolicy's pi-MLP [256,128,64] + action_net).
 * One fused tcgen05 kernel per call: bf16 operands, fp32 accumulation in tensor memory, gates in fp32.
 * All weight pointers are HOST fp32 arrays in torch's own layouts (Linear.weight [out][in]; nn.LSTM weight_ih
 * [4*256][in], weight_hh [4*256][256], gate order i,f,g,o); they are packed once, at creation. */
typedef struct acb200_policy acb200_policy;
enum { ACB200_POLICY_LSTM = 0, ACB200_POLICY_MLP = 1 };
#define ACB200_POLICY_OBS 18
#define ACB200_POLICY_ACT 4
#define ACB200_POLICY_HIDDEN 256
#define ACB200_POLICY_TILE 128 /* aircraft per thread block; recurrent state buffers are sized in whole tiles */

typedef struct acb200_lstm_policy_weights {
    const float *embed_w, *embed_b;                      /* features_extractor.embedding.0     [128][18], [128] */
    const float *l0_w_ih, *l0_b_ih, *l0_b_hh;            /* features_extractor.lstm  *_l0  [1024][128], [1024] x2 */
    const float *l1_w_ih, *l1_b_ih, *l1_b_hh;            /* features_extractor.lstm  *_l1  [1024][256], [1024] x2 */
    const float *proj_w, *proj_b;                        /* features_extractor.output_proj.0  [128][256], [128] */
    const float *actor_w_ih, *actor_w_hh;                /* lstm_actor *_l0     [1024][128], [1024][256] */
    const float *actor_b_ih, *actor_b_hh;                /*                     [1024] x2 */
    const float *pi0_w, *pi0_b, *pi1_w, *pi1_b;          /* mlp_extractor.policy_net.0 [128][256], .2 [64][128] */
    const float *action_w, *action_b;                    /* action_net [4][64], [4] */
} acb200_lstm_policy_weights;
/* (the extractor LSTM's weight_hh_l0/l1 multiply h0 = 0 on every call, lstm_policy.py:86, and are not needed) */

typedef struct acb200_mlp_policy_weights {
    const float *pi0_w, *pi0_b, *pi1_w, *pi1_b, *pi2_w, *pi2_b; /* policy_net.0 [256][18], .2 [128][256], .4 [64][128] */
    const float *action_w, *action_b;                           /* action_net [4][64], [4] */
} acb200_mlp_policy_weights;

int acb200_policy_create_lstm(acb200_ctx* ctx, const acb200_lstm_policy_weights* w, acb200_policy** out);
int acb200_policy_create_mlp(acb200_ctx* ctx, const acb200_mlp_policy_weights* w, acb200_policy** out);
void acb200_policy_destroy(acb200_policy* p);
/* bytes of the recurrent-state buffers for n aircraft (0, 0 for the MLP policy): h as bf16 tiles, c as f32 tiles */
int acb200_policy_state_bytes(const acb200_policy* p, int64_t n, size_t* h_bytes, size_t* c_bytes);
/* obs f32[n][18] (row-major, what acb200_env_step writes); episode_start / episode_start_b u8[n] or NULL: the state is
 * zeroed where EITHER is set (sb3-contrib's (1 - episode_start) * state; two arrays so that the env's terminated and
 * truncated outputs can be passed as they are); h_state / c_state device buffers of acb200_policy_state_bytes (updated
 * in place; NULL for the MLP policy); action_out f32[n][4] clipped to the action space (rate_env.py:128-138). */
int acb200_policy_forward(acb200_ctx* ctx, acb200_policy* p, int64_t n, const float* obs, const uint8_t* episode_start,
                          const uint8_t* episode_start_b, void* h_state, void* c_state, float* action_out);
/* `steps` closed-loop steps of the whole fleet in one call: action = policy(env.obs, episode_start = env.terminated |
 * env.truncated); acb200_env_step(action); ... (BASELINE config C4: policy forward fused with env step and reward; the loop
 * of learned_controllers/eval_rate.py:89-108 for every aircraft).  cfg->controller must be EXTERNAL or RESIDUAL.  `action`
 * float[n][4] is the caller's buffer the actions travel through (it holds the last step's afterwards).  The fleet is cut into
 * `parts` ranges (0 = automatic: 2 from 65 536 aircraft, else 1; at most 3) that advance independently on internal streams,
 * so that one part's env step runs beside another part's policy forward on the same SMs; results are identical to the
 * sequence of acb200_policy_forward + acb200_env_step calls.  Asynchronous (the context's stream continues after every part)
 * and capturable in a CUDA graph after one warm-up call. */
int acb200_policy_env_rollout(acb200_ctx* ctx, acb200_policy* p, const acb200_env* env, const acb200_env_config* cfg, void* h_state,
                              void* c_state, float* action, int32_t steps, int32_t parts);
/* recurrent state <-> the (n, 256) f32 arrays of SB3's lstm_states; all device pointers */
int acb200_policy_state_export(acb200_ctx* ctx, const acb200_policy* p, int64_t n, const void* h_state, const void* c_state,
                               float* h_out, float* c_out);
int acb200_policy_state_import(acb200_ctx* ctx, const acb200_policy* p, int64_t n, const float* h_in, const float* c_in,
                               void* h_state, void* c_state);
/* algorithmic FLOPs (2*M*N*K over the layers as executed) and packed weight bytes per 128-aircraft tile */
int acb200_policy_info(const acb200_policy* p, double* flops_per_aircraft, size_t* weight_bytes, int* smem_bytes);

/* ---- rollout post-processing on the device ------------------------------------------------
 * Replaces stable-baselines3 RolloutBuffer.compute_returns_and_advantage (third party; run by the PPO / RecurrentPPO the
 * reference builds at learned_controllers/utils/training_utils.py:112-160, train_overnight.py:151-167) on device-resident
 * rollouts: rewards, values, advantages, returns f32 [T][n] row-major; episode_starts u8 [T][n] (1 where step t began an
 * episode); last_values f32 [n], dones u8 [n] for the step after the last one.  float32 arithmetic in numpy's order:
 * bit-identical to the host implementation. */
int acb200_gae(acb200_ctx* ctx, int32_t T, int64_t n, const float* rewards, const float* values, const uint8_t* episode_starts,
               const float* last_values, const uint8_t* dones, double gamma, double gae_lambda, float* advantages, float* returns);

/* ---- host-side PID core (the same fp32 arithmetic as the kernels; mirrors PIDController::compute/reset) ---- */
float acb200_pid_compute_f32(const acb200_pid_config* cfg, acb200_pid_state* st, float setpoint, float measurement, float dt);
void acb200_pid_reset(acb200_pid_state* st);

/* ---- introspection ---- */
int acb200_kernel_info(int precision, const char* kernel, int* regs, int* local_bytes, int* max_threads);
int64_t acb200_launch_count(const acb200_ctx* ctx);
/* measured non-tensor FP32 throughput of this device (FMA = 2 FLOP), the denominator of the fused step's compute roofline */
int acb200_measure_fp32_peak(acb200_ctx* ctx, double* tflops);   /* kernels launched through this context so far */

#ifdef __cplusplus
}
#endif
#endif /* ACB200_H */
