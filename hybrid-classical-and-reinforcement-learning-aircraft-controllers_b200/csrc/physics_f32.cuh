// physics_f32.cuh -- THROUGHPUT build of one aircraft's 6-DOF substep (float), written for the
// Blackwell SM's issue budget rather than for the reference's operation order.
//
// Same model as physics.cuh (Simplified6DOF._dynamics / step, simulation/simplified_6dof.py:228-293,
// 333-503), restructured so that one derivative evaluation costs ~190 issue slots instead of ~580:
//   * everything that depends only on the held controls (zero-order hold over the K substeps,
//     simulation_backend.py:95-99) or only on parameters is folded once per control step / on the host;
//   * no IEEE division or square root in the loop: MUFU.RCP / MUFU.RSQ (<= 1 ulp) times a multiply;
//     divisions by mass / inertia / V_T0 / min airspeed are multiplications by host-side reciprocals;
//   * alpha = clip(atan2(w, u_safe)) needs atan only on |t| <= tan(max_alpha): one odd polynomial, no
//     octant logic; sin/cos(alpha) = (t, 1) / sqrt(1 + t^2);
//   * clamps are FMNMX pairs (fminf/fmaxf), not compare+select chains;
//   * the per-derivative NaN scrub of the reference (:496-501) is dropped: with finite state and
//     controls every derivative is finite by construction (V_safe >= 10, |u_safe| >= 0.1,
//     cos(theta_safe) >= cos 85 deg, clamped accelerations); the per-substep state scrub (:286-291)
//     stays, as ONE fused test with a cold repair path.
// The arithmetic differs from the fp64 oracle by float rounding only; tolerances are stated and
// tested in tests/test_gpu_parity_f32.py (per step <= 5e-6, trajectories within 6x of the oracle's own
// one-float-ulp sensitivity).
#pragma once
#include "math.cuh"
#include "params.cuh"

namespace acb {

#ifndef ACB_F32_UNROLL_STAGES
#define ACB_F32_UNROLL_STAGES 0
#endif
#ifndef ACB_F32_MUFU_TRIG
#define ACB_F32_MUFU_TRIG 1      // 1: sin.approx/cos.approx (MUFU, abs err ~5e-7) for all three Euler angles; 0: polynomials for roll/pitch
#endif

// Trigonometry of the throughput build.  Default (ACB_F32_MUFU_TRIG=1): the SFU's sin.approx / cos.approx for all
// three Euler angles, 4 issue slots per angle instead of ~27 for a Cody-Waite + polynomial sincos.  Its absolute
// error (~5e-7 on |x| <= 2 pi) perturbs the gravity projection by 5e-6 m/s^2 -- one float ulp of the force sums it
// is added to -- and the ground track by 1.5e-5 m/s; measured on B200 the per-step and trajectory error
// quantiles of the three variants (library-grade polynomial everywhere / polynomial for roll+pitch / MUFU
// everywhere) are identical to two digits up to the 99.99th percentile (profiles/r01_f32_trig_variants.md).
// ACB_F32_MUFU_TRIG=0 keeps yaw on MUFU (it enters only the position kinematics, :440-452) and uses for
//   pitch  (clamped to +-85 deg every substep, :268; a stage moves it by < 0.01 rad) one polynomial pair
//          without range reduction;
//   roll   (wrapped to [-pi, pi] every substep, :266) one fold about +-pi/2, then the same pair.
// Polynomials: least-squares Chebyshev fits in x^2 on |x| <= 1.6, abs err <= 1.5e-7 (tools/fit_trig_f32.py).
struct FastF32 {
    static __device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ void sincos_mufu(float x, float* s, float* c) {
        asm("sin.approx.ftz.f32 %0, %1;" : "=f"(*s) : "f"(x));
        asm("cos.approx.ftz.f32 %0, %1;" : "=f"(*c) : "f"(x));
    }
    // |x| <= 1.6
    static __device__ __forceinline__ void sincos_halfpi(float x, float* s, float* c) {
        const float z = x * x;
        float ps = fmaf(z, 2.630042899909313e-06f, -0.00019821259775198996f);
        ps = fmaf(ps, z, 0.008333231322467327f);
        ps = fmaf(ps, z, -0.1666666567325592f);
        *s = fmaf(x * z, ps, x);
        float pc = fmaf(z, -2.624827857289347e-07f, 2.4772412871243432e-05f);
        pc = fmaf(pc, z, -0.0013888622634112835f);
        pc = fmaf(pc, z, 0.041666656732559204f);
        pc = fmaf(pc, z, -0.5f);
        *c = fmaf(pc, z, 1.0f);
    }
    // |x| <= 3 pi / 2: sin(x) = sin(pi sign(x) - x), cos(x) = -cos(pi sign(x) - x) beyond +-pi/2
    static __device__ __forceinline__ void sincos_pi(float x, float* s, float* c) {
        const bool fold = fabsf(x) > 1.57079637050628662109375f;
        const float xr = fold ? (copysignf(3.14159274101257324f, x) - x) + copysignf(-8.74227765734758577e-08f, x) : x;
        float cc;
        sincos_halfpi(xr, s, &cc);
        *c = fold ? -cc : cc;
    }
    static __device__ __forceinline__ void sincos_pitch(float x, float* s, float* c) {
#if ACB_F32_MUFU_TRIG
        sincos_mufu(x, s, c);
#else
        sincos_halfpi(x, s, c);
#endif
    }
    static __device__ __forceinline__ void sincos_roll(float x, float* s, float* c) {
#if ACB_F32_MUFU_TRIG
        sincos_mufu(x, s, c);
#else
        sincos_pi(x, s, c);
#endif
    }
    static __device__ __forceinline__ void sincos_yaw(float x, float* s, float* c) { sincos_mufu(x, s, c); }
    // roll, pitch, yaw of one (stage) state.  The guards only fire for states the physics itself never produces
    // (a caller-injected attitude far outside the wrapped/clamped ranges, or a substep so long that a stage overshoots).
    // GUARDED = false: the caller has bounded the angles of all four stages for this substep (substep_angles_in_sfu_range).
    template <bool GUARDED = true>
    static __device__ __forceinline__ void attitude_trig(float phi, float theta, float psi, float* sphi, float* cphi, float* sth,
                                                         float* cth, float* spsi, float* cpsi) {
#if ACB_F32_MUFU_TRIG
        if (GUARDED && fmaxf(fmaxf(fabsf(phi), fabsf(theta)), fabsf(psi)) > 12.0f) {
            F32::sincos(phi, sphi, cphi);
            F32::sincos(theta, sth, cth);
            F32::sincos(psi, spsi, cpsi);
        } else {
            sincos_mufu(phi, sphi, cphi);
            sincos_mufu(theta, sth, cth);
            sincos_mufu(psi, spsi, cpsi);
        }
#else
        if (fabsf(theta) > 1.6f || fabsf(phi) > 4.7f) {
            F32::sincos(phi, sphi, cphi);
            F32::sincos(theta, sth, cth);
        } else {
            sincos_roll(phi, sphi, cphi);
            sincos_pitch(theta, sth, cth);
        }
        if (fabsf(psi) > 12.0f) F32::sincos(psi, spsi, cpsi);
        else sincos_yaw(psi, spsi, cpsi);
#endif
    }
    static __device__ __forceinline__ float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
    // asin with the common case (|x| <= 0.5, i.e. sideslip below 30 deg) as straight-line code
    static __device__ __forceinline__ float asin(float x) {
        const float ax = fabsf(x);
        float z = x * x, t = x;
        const bool big = ax > 0.5f;
        if (big) { z = fmaf(ax, -0.5f, 0.5f); t = sqrt(z); }
        float p = fmaf(z, 4.2163199048e-2f, 2.4181311049e-2f);
        p = fmaf(p, z, 4.5470025998e-2f);
        p = fmaf(p, z, 7.4953002686e-2f);
        p = fmaf(p, z, 1.6666752422e-1f);
        float r = fmaf(p * z, t, t);
        if (big) r = copysignf(fmaf(r, -2.0f, 1.57079637050628662109375f), x);
        return r;
    }
};

// What the K substeps of one control step share: the held controls folded into the coefficients
// they multiply (simplified_6dof.py:383-390, 402-404, 415-435).
struct HeldControlsF32 {
    float cl_c;        // cl_0 + cl_elevator * elevator_rad
    float cy;          // cy_rudder * rudder_rad
    float l_c, m_c, n_c;   // (cl_aileron aileron_rad, cm_elevator elevator_rad, cn_rudder rudder_rad) * 0.5 rho S (b|c) / I
    float thrust_max_m;  // max_thrust * throttle / mass
};

__device__ __forceinline__ HeldControlsF32 hold_controls(const DynParams<float>& P, const float (&c)[4]) {
    HeldControlsF32 H;
    H.cl_c = fmaf(P.cl_elev_rad, c[0], P.cl_0);
    H.m_c = P.qm_elev * c[0];
    H.l_c = P.ql_ail * c[1];
    H.cy = P.cy_rud_rad * c[2];
    H.n_c = P.qn_rud * c[2];
    H.thrust_max_m = (P.max_thrust * P.inv_mass) * c[3];
    return H;
}

// _dynamics for a stage state t[12] whose attitude sines/cosines are already known.
__device__ __forceinline__ void dynamics_f32(const DynParams<float>& P, const HeldControlsF32& H, const float (&t)[12],
                                             float sphi, float cphi, float sth, float cth, float spsi, float cpsi,
                                             float (&k)[12]) {
    const float u = t[3], v = t[4], w = t[5], theta = t[7], p = t[9], q = t[10], r = t[11];

    // airspeed, 1/V_safe (:363-364): V = V2 * rsqrt(V2); 1/max(V, Vmin) = min(1/V, 1/Vmin)
    const float V2 = fmaf(w, w, fmaf(v, v, u * u));
    const float rV = FastF32::rsqrt(fmaxf(V2, 1e-20f));
    const float V = V2 * rV;
    const float inv_Vs = fminf(rV, P.inv_min_airspeed);

    // u_safe and alpha (:367-373).  u_safe < 0 puts atan2 beyond +-90 deg, i.e. on the clip, with the sign of w.
    const float au = fabsf(u);
    const float us = (au > 1e-6f) ? copysignf(fmaxf(au, P.min_u_velocity), u) : P.min_u_velocity;
    float tcl = FastF32::clamp(w * FastF32::rcp(us), -P.tan_max_alpha, P.tan_max_alpha);
    tcl = (us < 0.0f) ? copysignf(P.tan_max_alpha, w) : tcl;
    const float alpha = F32::atan_unit(tcl);                 // odd polynomial, valid on [-1, 1]
    const float cos_alpha = FastF32::rsqrt(fmaf(tcl, tcl, 1.0f));
    const float sin_alpha = tcl * cos_alpha;

    // sideslip (:376) and dynamic pressure * S (:379, 392)
    const float beta = FastF32::asin(FastF32::clamp(v * inv_Vs, -1.0f, 1.0f));

    // aerodynamic + thrust + gravity accelerations (:383-411, 455-460); 1/m is folded into qS and thrust on the way in
    const float qS_m = P.half_rho_S_m * V2;
    const float cl = fmaf(P.cl_alpha, alpha, H.cl_c);
    const float cd = fmaf(P.cd_alpha2, alpha * alpha, P.cd_0);
    const float lift = qS_m * cl, drag = qS_m * cd;
    const float thrust = H.thrust_max_m * fmaxf(0.0f, fmaf(-V, P.inv_vt0, 1.0f));
    const float gcth = P.g * cth;
    const float ax = fmaf(-P.g, sth, fmaf(lift, sin_alpha, fmaf(-drag, cos_alpha, thrust)));
    const float ay = fmaf(gcth, sphi, qS_m * H.cy);
    const float az = fmaf(gcth, cphi, -fmaf(drag, sin_alpha, lift * cos_alpha));

    // moments / inertia (:415-435, 474-482); 0.5 rho S (b or c) / I is folded into every coefficient on the host
    const float l_a = V2 * fmaf(P.ql_beta, beta, fmaf(P.ql_damp * p, inv_Vs, H.l_c));
    const float m_a = V2 * fmaf(P.qm_damp * q, inv_Vs, fmaf(P.qm_alpha, alpha, H.m_c));
    const float n_a = V2 * fmaf(P.qn_beta, beta, fmaf(P.qn_damp * r, inv_Vs, H.n_c));

    // position kinematics (:440-452) as three plane rotations, R = Rz(psi) Ry(theta) Rx(phi)
    const float y1 = fmaf(cphi, v, -sphi * w), z1 = fmaf(sphi, v, cphi * w);
    const float x2 = fmaf(cth, u, sth * z1);
    k[0] = fmaf(cpsi, x2, -spsi * y1);
    k[1] = fmaf(spsi, x2, cpsi * y1);
    k[2] = fmaf(cth, z1, -sth * u);

    // translational dynamics (:455-460) with the acceleration clamp (:488-490)
    k[3] = FastF32::clamp(fmaf(r, v, fmaf(-q, w, ax)), -P.max_acceleration, P.max_acceleration);
    k[4] = FastF32::clamp(fmaf(p, w, fmaf(-r, u, ay)), -P.max_acceleration, P.max_acceleration);
    k[5] = FastF32::clamp(fmaf(q, u, fmaf(-p, v, az)), -P.max_acceleration, P.max_acceleration);

    // Euler-angle kinematics with the pitch clipped to +-max_pitch for tan/cos only (:463-471)
    const bool clipped = fabsf(theta) > P.max_pitch_rad;
    const float sth_s = clipped ? copysignf(P.sin_max_pitch, theta) : sth;
    const float cth_s = clipped ? P.cos_max_pitch : cth;
    const float rc = FastF32::rcp(cth_s);
    const float yaw_num = fmaf(sphi, q, cphi * r);
    k[6] = fmaf(sth_s * rc, yaw_num, p);
    k[7] = fmaf(cphi, q, -sphi * r);
    k[8] = yaw_num * rc;

    // Euler's equations (:474-487)
    k[9] = FastF32::clamp(fmaf(-P.gyro_x * q, r, l_a), -P.max_angular_acceleration, P.max_angular_acceleration);
    k[10] = FastF32::clamp(fmaf(-P.gyro_y * p, r, m_a), -P.max_angular_acceleration, P.max_angular_acceleration);
    k[11] = FastF32::clamp(fmaf(-P.gyro_z * p, q, n_a), -P.max_angular_acceleration, P.max_angular_acceleration);
}

// Cold path of the per-substep scrub (:258, 286-291): position nan->0 / +-inf->+-1e4, the rest -> 0.
static __device__ __noinline__ int repair_state_f32(float* s) {
    int repaired = 0;
    for (int i = 0; i < 12; ++i) {
        const float x = s[i];
        if (isfinite(x)) continue;
        s[i] = (i < 3 && x == x) ? (x > 0.0f ? 10000.0f : -10000.0f) : 0.0f;
        ++repaired;
    }
    return repaired;
}

// One test per substep instead of one per stage: a stage angle is angle + a k with a <= dt, |k[6..8]| <= (1 + 2 /
// cos(max_pitch)) max|rate| (:463-471 with the pitch clipped for tan / cos), and a stage rate is at most max|rate| + dt
// max_angular_acceleration (:485-487).  Within the bound no stage leaves the SFU's accurate range (|x| <= 12), so the
// per-stage test of attitude_trig -- a branch in front of every stage's MUFU block, which also fences the instruction
// scheduler -- can go; beyond it the substep runs the guarded code.  Both give the same bits.
__device__ __forceinline__ bool substep_angles_in_sfu_range(const DynParams<float>& P, const float (&s)[12], float dt) {
    const float ang = fmaxf(fmaxf(fabsf(s[6]), fabsf(s[7])), fabsf(s[8]));
    const float rate = fmaxf(fmaxf(fabsf(s[9]), fabsf(s[10])), fabsf(s[11]));
    return fmaf(dt * P.euler_rate_gain, rate + dt * P.max_angular_acceleration, ang) <= 11.5f;
}

// One RK4 substep (:248-254) and the post-step clamps (:258-291).
// MASKED (cold path only, backend_step_f32_nan_controls): derivative words whose bit is set in `mask` are zeroed after every
// stage -- what the reference's nan_to_num (:496-501) does to the words a NaN control surface poisons.
template <bool GUARDED, bool MASKED = false>
__device__ __forceinline__ int rk4_step_f32_impl(const DynParams<float>& P, const HeldControlsF32& H, float (&s)[12], float& time,
                                                 float dt, unsigned mask = 0u) {
    float k[12], acc[12], t[12];
    const float half_dt = 0.5f * dt;
    float sphi, cphi, sth, cth, spsi, cpsi;
    // stage 1 at the state itself
    FastF32::attitude_trig<GUARDED>(s[6], s[7], s[8], &sphi, &cphi, &sth, &cth, &spsi, &cpsi);
    dynamics_f32(P, H, s, sphi, cphi, sth, cth, spsi, cpsi, k);
    if (MASKED) {
#pragma unroll
        for (int i = 3; i < 12; ++i) k[i] = ((mask >> i) & 1u) ? 0.0f : k[i];
    }
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = k[i];
    // stages 2-4 at s + a k; weights 2, 2, 1.  Rolled: one more copy of the derivative code in the
    // instruction stream, not three.
#if ACB_F32_UNROLL_STAGES
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int stage = 1; stage < 4; ++stage) {
        const float a = (stage == 3) ? dt : half_dt;
        const float wgt = (stage == 3) ? 1.0f : 2.0f;
#pragma unroll
        for (int i = 0; i < 12; ++i) t[i] = fmaf(a, k[i], s[i]);
        FastF32::attitude_trig<GUARDED>(t[6], t[7], t[8], &sphi, &cphi, &sth, &cth, &spsi, &cpsi);
        dynamics_f32(P, H, t, sphi, cphi, sth, cth, spsi, cpsi, k);
        if (MASKED) {
#pragma unroll
            for (int i = 3; i < 12; ++i) k[i] = ((mask >> i) & 1u) ? 0.0f : k[i];
        }
#pragma unroll
        for (int i = 0; i < 12; ++i) acc[i] = fmaf(wgt, k[i], acc[i]);
    }
    const float dt6 = dt * (1.0f / 6.0f);
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = fmaf(dt6, acc[i], s[i]);
    time += dt;

    // :262-273
#pragma unroll
    for (int i = 3; i < 6; ++i) s[i] = FastF32::clamp(s[i], -P.max_velocity, P.max_velocity);
    s[6] = Math<float>::wrap_pi(s[6]);
    s[7] = FastF32::clamp(s[7], -P.max_pitch_rad, P.max_pitch_rad);
    s[8] = Math<float>::wrap_pi(s[8]);
#pragma unroll
    for (int i = 9; i < 12; ++i) s[i] = FastF32::clamp(s[i], -P.max_rate_rad, P.max_rate_rad);
    // :276-283  ground: altitude = -z < 0
    if (s[2] > 0.0f) {
        s[2] = 0.0f;
        s[5] = fmaxf(0.0f, s[5]);
    }
    // :258, 286-291 as one test: 0 * x is NaN for x = NaN or +-inf (fminf/fmaxf above turn a NaN into a bound,
    // so test the unclamped positions and attitude; velocities and rates are finite after their clamps unless NaN
    // went in, which the clamps map to a bound)
    // Only the positions, roll and yaw can still be non-finite here: the velocity, pitch and rate clamps above have mapped
    // NaN and +-inf to a bound.
    float chk = 0.0f;
#pragma unroll
    for (int i = 0; i < 12; ++i)
        if (i < 3 || i == 6 || i == 8) chk = fmaf(0.0f, s[i], chk);
    int repaired = 0;
    if (chk != 0.0f) {        // NaN compares unequal
        float tmp[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) tmp[i] = s[i];
        repaired = repair_state_f32(tmp);
#pragma unroll
        for (int i = 0; i < 12; ++i) s[i] = tmp[i];
    }
    return repaired;
}

// The substep with a test of every stage's angles (any state, any substep length).
__device__ __forceinline__ int rk4_step_f32(const DynParams<float>& P, const HeldControlsF32& H, float (&s)[12], float& time,
                                            float dt) {
    return rk4_step_f32_impl<true>(P, H, s, time, dt);
}
// Cold, out of line: the hot loop carries only the unguarded copy of the substep.
static __device__ __noinline__ int rk4_step_f32_guarded(const DynParams<float>& P, const HeldControlsF32& H, float* s, float* time,
                                                        float dt) {
    float x[12], t = *time;
#pragma unroll
    for (int i = 0; i < 12; ++i) x[i] = s[i];
    const int repaired = rk4_step_f32_impl<true>(P, H, x, t, dt);
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = x[i];
    *time = t;
    return repaired;
}

// A NaN held control (np.clip and set_controls let NaN through, simplified_6dof.py:221-226; +-inf are clipped to a bound)
// poisons, in the reference's _dynamics, exactly the derivative words that control enters, and nan_to_num zeroes them
// (:496-501): elevator -> CL -> Fx, Fz and the pitch moment (u', w', q'); aileron -> roll moment (p'); rudder -> side force and
// yaw moment (v', r'); throttle -> thrust (u').  Those words then stay frozen for the whole control step while the rest of
// the state integrates on.  fminf / fmaxf clamps would instead turn the NaN into a saturated acceleration, so a
// control step with a NaN control leaves the hot loop: this cold function evaluates the derivative with the control at 0 and
// zeroes the poisoned words after every stage.  Returns the number of zeroed derivative words per substep, summed (they are
// counted in the env's nonfinite_repairs statistic like the reference's logged "NaN/Inf detected in derivatives").
__device__ __forceinline__ unsigned nan_control_mask(const float (&c)[4]) {
    unsigned m = 0u;
    if (c[0] != c[0]) m |= (1u << 3) | (1u << 5) | (1u << 10);
    if (c[1] != c[1]) m |= (1u << 9);
    if (c[2] != c[2]) m |= (1u << 4) | (1u << 11);
    if (c[3] != c[3]) m |= (1u << 3);
    return m;
}
static __device__ __noinline__ int backend_step_f32_nan_controls(const DynParams<float>& P, float* s, float* time, const float* c,
                                                                 int K, float h) {
    float cc[4], x[12], t = *time;
#pragma unroll
    for (int i = 0; i < 4; ++i) cc[i] = c[i];
    const unsigned mask = nan_control_mask(cc);
#pragma unroll
    for (int i = 0; i < 4; ++i) cc[i] = (cc[i] != cc[i]) ? 0.0f : cc[i];
    const HeldControlsF32 H = hold_controls(P, cc);
#pragma unroll
    for (int i = 0; i < 12; ++i) x[i] = s[i];
    int repaired = 0;
#pragma unroll 1
    for (int k = 0; k < K; ++k) repaired += rk4_step_f32_impl<true, true>(P, H, x, t, h, mask) + __popc(mask);
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = x[i];
    *time = t;
    return repaired;
}

// SimulationAircraftBackend.step for the throughput build: K substeps of h under zero-order hold.
__device__ __forceinline__ int backend_step(const DynParams<float>& P, float (&s)[12], float& time, const float (&c)[4], int K,
                                            float h) {
    if (fmaf(0.0f, c[0], fmaf(0.0f, c[1], fmaf(0.0f, c[2], 0.0f * c[3]))) != 0.0f) {     // a NaN control: cold path
        float tmp[12], cc[4], tt = time;
#pragma unroll
        for (int i = 0; i < 12; ++i) tmp[i] = s[i];
#pragma unroll
        for (int i = 0; i < 4; ++i) cc[i] = c[i];
        const int r = backend_step_f32_nan_controls(P, tmp, &tt, cc, K, h);
#pragma unroll
        for (int i = 0; i < 12; ++i) s[i] = tmp[i];
        time = tt;
        return r;
    }
    const HeldControlsF32 H = hold_controls(P, c);
    int repaired = 0;
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
        if (substep_angles_in_sfu_range(P, s, h)) {
            repaired += rk4_step_f32_impl<false>(P, H, s, time, h);
        } else {
            float tmp[12], tt = time;          // only these copies live in local memory, not the loop-carried state / the held terms
            const HeldControlsF32 Hc = H;
#pragma unroll
            for (int i = 0; i < 12; ++i) tmp[i] = s[i];
            repaired += rk4_step_f32_guarded(P, Hc, tmp, &tt, h);
#pragma unroll
            for (int i = 0; i < 12; ++i) s[i] = tmp[i];
            time = tt;
        }
    }
    return repaired;
}

}  // namespace acb
