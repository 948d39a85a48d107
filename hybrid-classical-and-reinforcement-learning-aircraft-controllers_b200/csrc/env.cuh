// env.cuh -- one aircraft's RateControlEnv step / reset (learned_controllers/envs/rate_env.py:151-300,
// envs/rewards.py:48-137,168-221, data/generators.py:47-132,190-214), state in registers.
#pragma once
#include "agents.cuh"
#include "philox.cuh"
#include "physics.cuh"

namespace acb {

template <typename T> struct EnvRegs {
    T s[12], time, d[4];
    PidMem rate[3], ang[3];
    T rate_cmd[3];
    float prev_action[4];
    T prev_err[3], sign_changes[3], settle_timer;
    T sched[7];
    T env_time;
    int32_t step_count;
    uint32_t episode_idx;
    uint8_t is_settled, command_type;
    float difficulty;
    T ep_return;
};

template <typename T> struct EnvConst {
    T dt, h, pid_throttle;
    float dt_f32, residual_scale;
    PidDt pid_dt;          // dt_f32 with the reciprocal the PIDs divide by (pid.cuh)
    int32_t K, max_steps, controller, auto_reset;
    int32_t settle_steps;   // throughput build: consecutive settled steps before the bonus starts
    uint64_t seed;
    int64_t id_offset;
};

// generators.py:41-45: difficulty_scale is the Python float 0.3 / 0.5 / 0.7.  The fleet stores it as float32 (the curriculum
// is one small tensor write); (double)0.3f is 4e-8 away from 0.3, so the value the reference multiplies with is recovered as
// the nearest multiple of 1e-6 (a correctly rounded quotient = the literal's double).  Reset / random-walk path only.
__device__ __forceinline__ double difficulty_scale(float difficulty) {
    return __ddiv_rn(::rint(__dmul_rn((double)difficulty, 1e6)), 1e6);
}

__device__ __constant__ const uint8_t kPerm3[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};

// RateControlEnv._get_observation, rate_env.py:374-408
template <typename T> __device__ __forceinline__ void env_observe(const EnvRegs<T>& e, float (&o)[18]) {
    o[0] = (float)e.s[9]; o[1] = (float)e.s[10]; o[2] = (float)e.s[11];
    o[3] = (float)e.rate_cmd[0]; o[4] = (float)e.rate_cmd[1]; o[5] = (float)e.rate_cmd[2];
    o[6] = (float)(e.rate_cmd[0] - e.s[9]); o[7] = (float)(e.rate_cmd[1] - e.s[10]); o[8] = (float)(e.rate_cmd[2] - e.s[11]);
    o[9] = (float)e.d[0]; o[10] = (float)e.d[1];
    o[11] = (float)e.s[6]; o[12] = (float)e.s[7]; o[13] = (float)e.s[8];
    o[14] = e.prev_action[0]; o[15] = e.prev_action[1]; o[16] = e.prev_action[2]; o[17] = e.prev_action[3];
}

// What a reset draws: RateControlEnv.reset with FlightEnvelopeSampler.sample and _generate_new_command.
template <typename T> struct ResetDraw {
    T s[12], d[4], rate_cmd[3], sched[7];
};

// Draw plan (Philox block b of counter (id_lo, id_hi, episode, b)):
//   b0 airspeed U(15,30), altitude U(50,200) | b1 roll, pitch U(+-15 deg) | b2 yaw U(0,2pi), p U(+-0.1)
//   b3 q, r U(+-0.1) | b4 w0 -> #axes, w1 -> permutation, w2 bits -> signs | b5 mag0, mag1 | b6 mag2, sine f
// All sampling arithmetic is double in both builds (a reset is rare), then narrowed to T.
// Not inlined: it is the cold path of the step kernel, and keeping it out of line keeps its registers
// (Philox rounds, double arithmetic) out of the hot loop's allocation.
template <typename T>
__device__ __noinline__ void env_reset_draw(uint64_t seed, uint64_t global_id, uint32_t episode_idx, float difficulty,
                                            uint8_t command_type, ResetDraw<T>* out) {
    const double PI = 3.141592653589793;
    const Philox g(seed, global_id, episode_idx);
    uint32_t w[4];
    const double att_lo = -15.0 * (PI / 180.0), att_hi = 15.0 * (PI / 180.0);
    g.block(0, w);
    const double airspeed = uniform_lh(15.0, 30.0, u53(w[0], w[1]));
    const double altitude = uniform_lh(50.0, 200.0, u53(w[2], w[3]));
    g.block(1, w);
    const double roll = uniform_lh(att_lo, att_hi, u53(w[0], w[1]));
    const double pitch = uniform_lh(att_lo, att_hi, u53(w[2], w[3]));
    g.block(2, w);
    const double yaw = uniform_lh(0.0, 2 * PI, u53(w[0], w[1]));
    const double p = uniform_lh(-0.1, 0.1, u53(w[2], w[3]));
    g.block(3, w);
    const double q = uniform_lh(-0.1, 0.1, u53(w[0], w[1]));
    const double r = uniform_lh(-0.1, 0.1, u53(w[2], w[3]));

    T s[12];
    s[0] = T(0); s[1] = T(0); s[2] = (T)(-altitude);
    s[3] = (T)airspeed; s[4] = T(0); s[5] = T(0);
    s[6] = (T)roll; s[7] = (T)pitch; s[8] = (T)yaw;
    s[9] = (T)p; s[10] = (T)q; s[11] = (T)r;
    T d[4];
    derived_outputs(s, d);
#pragma unroll
    for (int k = 0; k < 12; ++k) out->s[k] = s[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) out->d[k] = d[k];

    const double max_rates[3] = {180.0 * (PI / 180.0), 180.0 * (PI / 180.0), 160.0 * (PI / 180.0)};
    const double scale = difficulty_scale(difficulty);
    uint32_t sel[4];
    g.block(4, sel);
    double mag[3];
    g.block(5, w);
    mag[0] = u53(w[0], w[1]); mag[1] = u53(w[2], w[3]);
    g.block(6, w);
    mag[2] = u53(w[0], w[1]);
    const double ufreq = u53(w[2], w[3]);
    const int perm = (int)__umulhi(sel[1], 6u);
    T sched[7], rate_cmd[3] = {T(0), T(0), T(0)};
#pragma unroll
    for (int f = 0; f < 7; ++f) sched[f] = T(0);
    double cmd[3] = {0.0, 0.0, 0.0};
    if (command_type == ACB200_CMD_STEP || command_type == ACB200_CMD_RAMP) {
        const int n_axes = 1 + (int)__umulhi(sel[0], 3u);
        for (int k = 0; k < n_axes; ++k) {
            const int axis = kPerm3[perm][k];
            const double magnitude = __dmul_rn(uniform_lh(0.3, 1.0, mag[k]), scale);
            const double sign = ((sel[2] >> k) & 1u) ? 1.0 : -1.0;
            cmd[axis] = __dmul_rn(__dmul_rn(sign, magnitude), max_rates[axis]);
        }
        if (command_type == ACB200_CMD_STEP) {
            for (int a = 0; a < 3; ++a) rate_cmd[a] = (T)cmd[a];
        } else {
            for (int a = 0; a < 3; ++a) sched[a] = (T)cmd[a];
        }
    } else if (command_type == ACB200_CMD_SINE) {
        const int n_axes = 1 + (int)__umulhi(sel[0], 2u);
        for (int k = 0; k < n_axes; ++k) {
            const int axis = kPerm3[perm][k];
            const double amp = __dmul_rn(__dmul_rn(uniform_lh(0.3, 1.0, mag[k]), 0.5), scale);
            cmd[axis] = __dmul_rn(amp, max_rates[axis]);
        }
        for (int a = 0; a < 3; ++a) sched[a] = (T)cmd[a];
        sched[3] = (T)uniform_lh(0.1, 2.0, ufreq);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) out->rate_cmd[a] = rate_cmd[a];
#pragma unroll
    for (int f = 0; f < 7; ++f) out->sched[f] = sched[f];
}

// RateControlEnv.reset (rate_env.py:151-210): a fresh episode for one aircraft.
template <typename T> __device__ __forceinline__ void env_reset_one(const EnvConst<T>& E, int64_t local_i, EnvRegs<T>& e) {
    e.episode_idx += 1;
    ResetDraw<T> dr;
    env_reset_draw<T>(E.seed, (uint64_t)(E.id_offset + local_i), e.episode_idx, e.difficulty, e.command_type, &dr);
#pragma unroll
    for (int k = 0; k < 12; ++k) e.s[k] = dr.s[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) e.d[k] = dr.d[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) e.rate_cmd[k] = dr.rate_cmd[k];
#pragma unroll
    for (int k = 0; k < 7; ++k) e.sched[k] = dr.sched[k];
    e.time = T(0);
    e.env_time = T(0);
    e.step_count = 0;
    e.prev_action[0] = 0.0f; e.prev_action[1] = 0.0f; e.prev_action[2] = 0.0f; e.prev_action[3] = 0.5f;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        e.prev_err[a] = T(0); e.sign_changes[a] = T(0);
        e.rate[a] = PidMem{0.0f, 0.0f, 0.0f}; e.ang[a] = PidMem{0.0f, 0.0f, 0.0f};
    }
    e.settle_timer = T(0);
    e.is_settled = 0;
    e.ep_return = T(0);
}

// RateControlEnv._update_command, rate_env.py:342-372
template <typename T> __device__ __forceinline__ void env_update_command(const EnvConst<T>& E, int64_t local_i, EnvRegs<T>& e) {
    using M = Math<T>;
    const T PI = T(3.141592653589793);
    if (e.command_type == ACB200_CMD_RAMP) {
        const T t = e.env_time, duration = T(3);
        if (t < duration) {
            const T a = M::div(t, duration);
#pragma unroll
            for (int k = 0; k < 3; ++k) e.rate_cmd[k] = M::mad(a, e.sched[k], M::mul(T(1) - a, T(0)));
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) e.rate_cmd[k] = e.sched[k];
        }
    } else if (e.command_type == ACB200_CMD_RANDOM) {
        // generators.py:145-164: delta = randn(3) * 0.1 * sqrt(dt) * scale * max_rates, then clip
        const double DPI = 3.141592653589793;
        const double max_rates[3] = {180.0 * (DPI / 180.0), 180.0 * (DPI / 180.0), 160.0 * (DPI / 180.0)};
        const Philox g(E.seed, (uint64_t)(E.id_offset + local_i), e.episode_idx);
        uint32_t w[4];
        g.block(16u + (uint32_t)e.step_count, w);
        const double u0 = ((double)w[0] + 1.0) / 4294967296.0, u1 = (double)w[1] / 4294967296.0;
        const double u2 = ((double)w[2] + 1.0) / 4294967296.0, u3 = (double)w[3] / 4294967296.0;
        const double r0 = ::sqrt(-2.0 * ::log(u0)), r1 = ::sqrt(-2.0 * ::log(u2));
        const double z[3] = {r0 * ::cos(2 * DPI * u1), r0 * ::sin(2 * DPI * u1), r1 * ::cos(2 * DPI * u3)};
        const double sdt = ::sqrt((double)E.dt);
        const double scale = difficulty_scale(e.difficulty);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double delta = __dmul_rn(__dmul_rn(z[k], 0.1), sdt);
            delta = __dmul_rn(delta, __dmul_rn(scale, max_rates[k]));
            e.rate_cmd[k] = (T)np_clip((double)e.rate_cmd[k] + delta, -max_rates[k], max_rates[k]);
        }
    } else if (e.command_type == ACB200_CMD_SINE) {
        const T sn = M::sin(T(2) * PI * e.sched[3] * e.env_time);
#pragma unroll
        for (int k = 0; k < 3; ++k) e.rate_cmd[k] = e.sched[k] * sn;
    }
}

// Who drives the action of RateControlEnv.step (rate_env.py:212-233): the caller (EXTERNAL), the PID
// compositions the reference itself uses (pid_demonstrations.py:55-78: RATE_PID; AttitudeAgent in front:
// ATTITUDE_PID) or PID + residual (residual_rate_env.py:112-155).
// act_in = the caller's action / residual, env order aileron, elevator, rudder, throttle.
// Returns the clipped float32 action in act and the surfaces the simulator holds in c.
template <typename T>
__device__ __forceinline__ void env_action(const CtrlParams<T>& C, const EnvConst<T>& E, const T (&s)[12], PidMem (&rate)[3],
                                           PidMem (&ang)[3], const T (&rate_cmd)[3], const T (&att_cmd)[4],
                                           const float (&act_in)[4], float (&act)[4], T (&c)[4], T& residual_bonus) {
    residual_bonus = T(0);
    if (E.controller == ACB200_CTRL_EXTERNAL) {
#pragma unroll
        for (int k = 0; k < 4; ++k) act[k] = act_in[k];
    } else if (E.controller == ACB200_CTRL_ATTITUDE_PID) {
        T surf[4];
        attitude_agent(C, ang, rate, att_cmd[0], att_cmd[1], att_cmd[2], att_cmd[3], s, E.pid_dt, surf);
        act[0] = (float)surf[1]; act[1] = (float)surf[0]; act[2] = (float)surf[2]; act[3] = (float)surf[3];
    } else {
        // pid_demonstrations.py:55-78 / residual_rate_env.py:112-135
        T surf[4];
        rate_agent(C, rate, rate_cmd[0], rate_cmd[1], rate_cmd[2], E.pid_throttle, s[9], s[10], s[11], E.pid_dt, surf);
        act[0] = (float)surf[1]; act[1] = (float)surf[0]; act[2] = (float)surf[2]; act[3] = (float)surf[3];
        if (E.controller == ACB200_CTRL_RESIDUAL) {
            // residual_rate_env.py:138-155, float32 arithmetic (individually rounded, as numpy does it)
#pragma unroll
            for (int k = 0; k < 4; ++k) act[k] = __fadd_rn(act[k], __fmul_rn(act_in[k], E.residual_scale));
#pragma unroll
            for (int k = 0; k < 3; ++k) act[k] = (act[k] < -1.0f) ? -1.0f : ((act[k] > 1.0f) ? 1.0f : act[k]);
            act[3] = (act[3] < 0.0f) ? 0.0f : ((act[3] > 1.0f) ? 1.0f : act[3]);
            const float rm = __fadd_rn(__fadd_rn(__fmul_rn(act_in[0], act_in[0]), __fmul_rn(act_in[1], act_in[1])),
                                       __fmul_rn(act_in[2], act_in[2]));
            residual_bonus = (T)__fmul_rn(0.05f, __fsub_rn(1.0f, __fdiv_rn(rm, 3.0f)));
        }
    }
    // :225 clip to the float32 Box (NaN propagates like np.clip)
#pragma unroll
    for (int k = 0; k < 3; ++k) act[k] = np_clip(act[k], -1.0f, 1.0f);
    act[3] = np_clip(act[3], 0.0f, 1.0f);
    // :228-233 and set_controls (simplified_6dof.py:221-226)
    const T cin[4] = {(T)act[1], (T)act[0], (T)act[2], (T)act[3]};
    clip_controls(cin, c);
}

// The rest of RateControlEnv.step after sim.step (rate_env.py:241-300): time, command update, reward,
// settle bonus, termination.  e.s / e.d hold the post-step state and derived outputs.
// comp[8] = tracking, smoothness, stability, oscillation, survival, settle_bonus, crash_penalty, residual_bonus.
template <typename T>
__device__ __forceinline__ void env_after_physics(const EnvConst<T>& E, EnvRegs<T>& e, const float (&act)[4], T residual_bonus,
                                                  T (&comp)[8], T& reward_out, bool& terminated_out, bool& truncated_out,
                                                  int64_t local_i) {
    using M = Math<T>;
    // :241-245
    e.step_count += 1;
    if (M::kParity) e.env_time += E.dt;
    else e.env_time = (T)e.step_count * E.dt;
    env_update_command(E, local_i, e);
    // :248-252
    const T err[3] = {e.rate_cmd[0] - e.s[9], e.rate_cmd[1] - e.s[10], e.rate_cmd[2] - e.s[11]};
    const T airspeed = e.d[0], altitude = e.d[1], roll = e.s[6], pitch = e.s[7];

    // RateTrackingReward.compute, rewards.py:48-137
    const T tracking_error = M::div_by(M::mad(err[2], err[2], M::mad(err[1], err[1], M::sq(err[0]))), T(3));
    const T r_tracking = M::mul(T(-0.5), tracking_error);
    T r_smoothness;
    if (e.step_count == 1 && M::kParity) {
        // first step: prev_action is the float64 array of rate_env.py:193 -> float64 arithmetic
        T cc = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { const T d = (T)act[k] - (T)e.prev_action[k]; cc = cc + d * d; }
        r_smoothness = T(-0.01) * cc;
    } else {
        // float32 - float32 (rate_env.py:282) with a weak Python scalar (NumPy >= 2): float32 arithmetic
        float cc = 0.0f;
#pragma unroll
        for (int k = 0; k < 3; ++k) { const float d = __fsub_rn(act[k], e.prev_action[k]); cc = __fadd_rn(cc, __fmul_rn(d, d)); }
        r_smoothness = (T)__fmul_rn(-0.01f, cc);
    }
    const T RAD45 = T(45.0 * (3.141592653589793 / 180.0)), RAD30 = T(30.0 * (3.141592653589793 / 180.0));
    const T roll_stability = M::exp_neg(M::div_by(-M::abs(roll), RAD45));
    const T pitch_stability = M::exp_neg(M::div_by(-M::abs(pitch), RAD30));
    const T airspeed_stability = np_clip(M::div_by(airspeed - T(8), T(12)), T(0), T(1));
    const T altitude_stability = np_clip(M::div_by(altitude - T(10), T(90)), T(0), T(1));
    const T stability_score = M::div_by(roll_stability + pitch_stability + airspeed_stability + altitude_stability, T(4));
    const T r_stability = M::mul(T(0.3), stability_score);
    T osc = T(0);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // (np.sign(err) != np.sign(prev_err)) and abs(prev_err) > 0.01, rewards.py:104-107.  The second condition makes prev_err
        // a non-zero number, so its sign is +-1 and the first one says "err does not have that strict sign" (true for a zero
        // or NaN err as well, as np.sign gives 0 / NaN there): err mirrored onto prev_err's side must not be positive.
        const T mirrored = (e.prev_err[k] < T(0)) ? -err[k] : err[k];
        const bool change = !(mirrored > T(0)) && (M::abs(e.prev_err[k]) > T(0.01));
        e.sign_changes[k] = M::mad(T(0.9), e.sign_changes[k], change ? T(1) : T(0));
        osc = osc + e.sign_changes[k];
        e.prev_err[k] = err[k];
    }
    const T r_oscillation = M::mul(T(-0.1), osc);
    const T r_survival = T(1);
    T reward = r_tracking + r_smoothness + r_stability + r_oscillation + r_survival;

    // SettlingTimeBonus.compute, rewards.py:168-221
    T settle_reward = T(0);
    {
        bool settled = true;
#pragma unroll
        for (int k = 0; k < 3; ++k) settled = settled && (M::abs(err[k]) < py_max(M::abs(e.rate_cmd[k]) * T(0.05), T(0.05)));
        if (settled) {
            bool enough;
            if (M::kParity) { e.settle_timer += E.dt; enough = e.settle_timer >= T(0.2); }
            else { e.settle_timer += T(1); enough = e.settle_timer >= (T)E.settle_steps; }   // see DESIGN.md: fp64 timer rule
            if (enough) { e.is_settled = 1; settle_reward = M::mul(T(2), E.dt); }
        } else {
            e.settle_timer = T(0);
            e.is_settled = 0;
        }
    }
    reward += settle_reward;
#pragma unroll
    for (int k = 0; k < 4; ++k) e.prev_action[k] = act[k];

    // _check_termination :437-460, truncation :286, crash penalty :289-294
    const T RAD120 = T(120.0 * (3.141592653589793 / 180.0)), RAD80 = T(80.0 * (3.141592653589793 / 180.0));
    const bool terminated = (altitude < T(5)) || (M::abs(roll) > RAD120) || (M::abs(pitch) > RAD80) || (airspeed < T(8));
    const bool truncated = e.step_count >= E.max_steps;
    T crash_penalty = T(0);
    if (terminated && !truncated) { crash_penalty = T(-100); reward += crash_penalty; }
    reward += residual_bonus;
    e.ep_return += reward;

    comp[0] = r_tracking; comp[1] = r_smoothness; comp[2] = r_stability; comp[3] = r_oscillation; comp[4] = r_survival;
    comp[5] = settle_reward; comp[6] = crash_penalty; comp[7] = residual_bonus;
    reward_out = reward;
    terminated_out = terminated;
    truncated_out = truncated;
}

}  // namespace acb
