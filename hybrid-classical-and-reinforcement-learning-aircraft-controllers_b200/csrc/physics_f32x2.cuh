// physics_f32x2.cuh -- the throughput build's substep for TWO aircraft per thread on Blackwell's packed
// FP32 instructions (FFMA2 / FMUL2 / FADD2: one issue slot, two IEEE single-precision results).
//
// Why: env_step_kernel<float> is bound by the SM's instruction ISSUE rate (86 % of issue slots, FMA pipe 50 %,
// ALU 39 %, MUFU 36 %; profiles/r01c_ncu_env_step_f32_summary.csv), not by any one pipe.  A packed
// instruction carries the same FMA-pipe work in half the issue slots (tools/probes/ffma2_probe.cu: the
// substep's instruction mix runs 1.40x faster packed than scalar on B200), so the pair of aircraft a
// thread owns shares every FMA-pipe instruction; min/max, selects and MUFU calls stay per lane.
//
// Same model and THE SAME ARITHMETIC as physics_f32.cuh (Simplified6DOF._dynamics / step,
// simulation/simplified_6dof.py:228-293, 333-503): every lane operation below is the IEEE operation the
// scalar code performs, in the same order, so a lane's trajectory is bit-identical to the scalar
// kernel's (tests/test_gpu_pair_kernel.py).  Lane x = aircraft 2t, lane y = aircraft 2t+1.
#pragma once
#include "physics_f32.cuh"

#if !ACB_F32_MUFU_TRIG
#error "physics_f32x2.cuh implements the SFU trigonometry variant only"
#endif
#ifndef ACB_F32X2_UNROLL_STAGES
#define ACB_F32X2_UNROLL_STAGES 1
#endif

namespace acb {

using v2 = float2;

struct P2 {
    static __device__ __forceinline__ v2 bc(float c) { return make_float2(c, c); }
    static __device__ __forceinline__ v2 fma(v2 a, v2 b, v2 c) { return __ffma2_rn(a, b, c); }
    static __device__ __forceinline__ v2 mul(v2 a, v2 b) { return __fmul2_rn(a, b); }
    static __device__ __forceinline__ v2 add(v2 a, v2 b) { return __fadd2_rn(a, b); }
    static __device__ __forceinline__ v2 neg(v2 a) { return make_float2(-a.x, -a.y); }
    static __device__ __forceinline__ v2 vmin(v2 a, float b) { return make_float2(fminf(a.x, b), fminf(a.y, b)); }
    static __device__ __forceinline__ v2 vmax(v2 a, float b) { return make_float2(fmaxf(a.x, b), fmaxf(a.y, b)); }
    static __device__ __forceinline__ v2 clamp(v2 a, float lo, float hi) {
        return make_float2(FastF32::clamp(a.x, lo, hi), FastF32::clamp(a.y, lo, hi));
    }
    static __device__ __forceinline__ v2 rcp(v2 a) { return make_float2(FastF32::rcp(a.x), FastF32::rcp(a.y)); }
    static __device__ __forceinline__ v2 rsqrt(v2 a) { return make_float2(FastF32::rsqrt(a.x), FastF32::rsqrt(a.y)); }

    // F32::atan_unit on both lanes
    static __device__ __forceinline__ v2 atan_unit(v2 t) {
        const v2 z = mul(t, t);
        v2 p = fma(z, bc(0.00282363896258175373077393f), bc(-0.0159569028764963150024414f));
        p = fma(p, z, bc(0.0425049886107444763183594f));
        p = fma(p, z, bc(-0.0748900920152664184570312f));
        p = fma(p, z, bc(0.106347933411598205566406f));
        p = fma(p, z, bc(-0.142027363181114196777344f));
        p = fma(p, z, bc(0.199926957488059997558594f));
        p = fma(p, z, bc(-0.333331018686294555664062f));
        return fma(mul(p, z), t, t);
    }
    // FastF32::asin on both lanes: the lane-wise argument reduction stays scalar, the polynomial is shared
    static __device__ __forceinline__ v2 asin(v2 x) {
        const float ax0 = fabsf(x.x), ax1 = fabsf(x.y);
        v2 z = mul(x, x), t = x;
        const bool big0 = ax0 > 0.5f, big1 = ax1 > 0.5f;
        if (big0) { z.x = fmaf(ax0, -0.5f, 0.5f); t.x = FastF32::sqrt(z.x); }
        if (big1) { z.y = fmaf(ax1, -0.5f, 0.5f); t.y = FastF32::sqrt(z.y); }
        v2 p = fma(z, bc(4.2163199048e-2f), bc(2.4181311049e-2f));
        p = fma(p, z, bc(4.5470025998e-2f));
        p = fma(p, z, bc(7.4953002686e-2f));
        p = fma(p, z, bc(1.6666752422e-1f));
        v2 r = fma(mul(p, z), t, t);
        if (big0) r.x = copysignf(fmaf(r.x, -2.0f, 1.57079637050628662109375f), x.x);
        if (big1) r.y = copysignf(fmaf(r.y, -2.0f, 1.57079637050628662109375f), x.y);
        return r;
    }
    // The SFU branch of FastF32::attitude_trig on both lanes; the range guard is rk4_step_f32x2's, once per substep
    static __device__ __forceinline__ void attitude_trig(v2 phi, v2 theta, v2 psi, v2* sphi, v2* cphi, v2* sth, v2* cth, v2* spsi,
                                                         v2* cpsi) {
        FastF32::sincos_mufu(phi.x, &sphi->x, &cphi->x); FastF32::sincos_mufu(phi.y, &sphi->y, &cphi->y);
        FastF32::sincos_mufu(theta.x, &sth->x, &cth->x); FastF32::sincos_mufu(theta.y, &sth->y, &cth->y);
        FastF32::sincos_mufu(psi.x, &spsi->x, &cpsi->x); FastF32::sincos_mufu(psi.y, &spsi->y, &cpsi->y);
    }
    // Math<float>::wrap_pi on both lanes
    static __device__ __forceinline__ v2 wrap_pi(v2 a) {
        const v2 ks = mul(a, bc(0.159154943091895346f));
        const v2 k = make_float2(rintf(ks.x), rintf(ks.y));
        v2 r = fma(k, bc(-6.28318548202514648f), a);
        r = fma(k, bc(1.74845553146951715e-07f), r);
        return r;
    }
};

struct HeldControlsF32x2 {
    v2 cl_c, cy, l_c, m_c, n_c, thrust_max_m;
};

__device__ __forceinline__ HeldControlsF32x2 hold_controls2(const DynParams<float>& P, const float (&c0)[4], const float (&c1)[4]) {
    const HeldControlsF32 a = hold_controls(P, c0), b = hold_controls(P, c1);
    HeldControlsF32x2 H;
    H.cl_c = make_float2(a.cl_c, b.cl_c); H.cy = make_float2(a.cy, b.cy);
    H.l_c = make_float2(a.l_c, b.l_c); H.m_c = make_float2(a.m_c, b.m_c); H.n_c = make_float2(a.n_c, b.n_c);
    H.thrust_max_m = make_float2(a.thrust_max_m, b.thrust_max_m);
    return H;
}

// dynamics_f32 for the two lanes.  Line for line the scalar function; see it for the model references.
__device__ __forceinline__ void dynamics_f32x2(const DynParams<float>& P, const HeldControlsF32x2& H, const v2 (&t)[12], v2 sphi,
                                               v2 cphi, v2 sth, v2 cth, v2 spsi, v2 cpsi, v2 (&k)[12]) {
    using X = P2;
    const v2 u = t[3], v = t[4], w = t[5], theta = t[7], p = t[9], q = t[10], r = t[11];

    const v2 V2 = X::fma(w, w, X::fma(v, v, X::mul(u, u)));
    const v2 rV = X::rsqrt(X::vmax(V2, 1e-20f));
    const v2 V = X::mul(V2, rV);
    const v2 inv_Vs = X::vmin(rV, P.inv_min_airspeed);

    const float au0 = fabsf(u.x), au1 = fabsf(u.y);
    v2 us;
    us.x = (au0 > 1e-6f) ? copysignf(fmaxf(au0, P.min_u_velocity), u.x) : P.min_u_velocity;
    us.y = (au1 > 1e-6f) ? copysignf(fmaxf(au1, P.min_u_velocity), u.y) : P.min_u_velocity;
    v2 tcl = X::clamp(X::mul(w, X::rcp(us)), -P.tan_max_alpha, P.tan_max_alpha);
    tcl.x = (us.x < 0.0f) ? copysignf(P.tan_max_alpha, w.x) : tcl.x;
    tcl.y = (us.y < 0.0f) ? copysignf(P.tan_max_alpha, w.y) : tcl.y;
    const v2 alpha = X::atan_unit(tcl);
    const v2 cos_alpha = X::rsqrt(X::fma(tcl, tcl, X::bc(1.0f)));
    const v2 sin_alpha = X::mul(tcl, cos_alpha);

    const v2 beta = X::asin(X::clamp(X::mul(v, inv_Vs), -1.0f, 1.0f));

    const v2 qS_m = X::mul(X::bc(P.half_rho_S_m), V2);
    const v2 cl = X::fma(X::bc(P.cl_alpha), alpha, H.cl_c);
    const v2 cd = X::fma(X::bc(P.cd_alpha2), X::mul(alpha, alpha), X::bc(P.cd_0));
    const v2 lift = X::mul(qS_m, cl), drag = X::mul(qS_m, cd);
    const v2 thrust = X::mul(H.thrust_max_m, X::vmax(X::fma(V, X::bc(-P.inv_vt0), X::bc(1.0f)), 0.0f));
    const v2 gcth = X::mul(X::bc(P.g), cth);
    const v2 ax = X::fma(X::bc(-P.g), sth, X::fma(lift, sin_alpha, X::fma(X::neg(drag), cos_alpha, thrust)));
    const v2 ay = X::fma(gcth, sphi, X::mul(qS_m, H.cy));
    const v2 az = X::fma(gcth, cphi, X::neg(X::fma(drag, sin_alpha, X::mul(lift, cos_alpha))));

    const v2 l_a = X::mul(V2, X::fma(X::bc(P.ql_beta), beta, X::fma(X::mul(X::bc(P.ql_damp), p), inv_Vs, H.l_c)));
    const v2 m_a = X::mul(V2, X::fma(X::mul(X::bc(P.qm_damp), q), inv_Vs, X::fma(X::bc(P.qm_alpha), alpha, H.m_c)));
    const v2 n_a = X::mul(V2, X::fma(X::bc(P.qn_beta), beta, X::fma(X::mul(X::bc(P.qn_damp), r), inv_Vs, H.n_c)));

    const v2 y1 = X::fma(cphi, v, X::mul(X::neg(sphi), w)), z1 = X::fma(sphi, v, X::mul(cphi, w));
    const v2 x2 = X::fma(cth, u, X::mul(sth, z1));
    k[0] = X::fma(cpsi, x2, X::mul(X::neg(spsi), y1));
    k[1] = X::fma(spsi, x2, X::mul(cpsi, y1));
    k[2] = X::fma(cth, z1, X::mul(X::neg(sth), u));

    k[3] = X::clamp(X::fma(r, v, X::fma(X::neg(q), w, ax)), -P.max_acceleration, P.max_acceleration);
    k[4] = X::clamp(X::fma(p, w, X::fma(X::neg(r), u, ay)), -P.max_acceleration, P.max_acceleration);
    k[5] = X::clamp(X::fma(q, u, X::fma(X::neg(p), v, az)), -P.max_acceleration, P.max_acceleration);

    const bool clipped0 = fabsf(theta.x) > P.max_pitch_rad, clipped1 = fabsf(theta.y) > P.max_pitch_rad;
    v2 sth_s, cth_s;
    sth_s.x = clipped0 ? copysignf(P.sin_max_pitch, theta.x) : sth.x;
    sth_s.y = clipped1 ? copysignf(P.sin_max_pitch, theta.y) : sth.y;
    cth_s.x = clipped0 ? P.cos_max_pitch : cth.x;
    cth_s.y = clipped1 ? P.cos_max_pitch : cth.y;
    const v2 rc = X::rcp(cth_s);
    const v2 yaw_num = X::fma(sphi, q, X::mul(cphi, r));
    k[6] = X::fma(X::mul(sth_s, rc), yaw_num, p);
    k[7] = X::fma(cphi, q, X::mul(X::neg(sphi), r));
    k[8] = X::mul(yaw_num, rc);

    k[9] = X::clamp(X::fma(X::mul(X::bc(-P.gyro_x), q), r, l_a), -P.max_angular_acceleration, P.max_angular_acceleration);
    k[10] = X::clamp(X::fma(X::mul(X::bc(-P.gyro_y), p), r, m_a), -P.max_angular_acceleration, P.max_angular_acceleration);
    k[11] = X::clamp(X::fma(X::mul(X::bc(-P.gyro_z), p), q, n_a), -P.max_angular_acceleration, P.max_angular_acceleration);
}

// Cold path of the substep: a lane whose Euler angles could leave the SFU's accurate range during the stages (a caller-
// injected attitude far outside the wrapped / clamped ranges, or a very long substep) runs the one-aircraft code, which
// guards every stage (FastF32::attitude_trig).
static __device__ __noinline__ int rk4_step_lanes_f32(const DynParams<float>& P, const HeldControlsF32x2& H, v2* s, float dt) {
    int repaired = 0;
#pragma unroll 1
    for (int l = 0; l < 2; ++l) {
        HeldControlsF32 h;
        h.cl_c = l ? H.cl_c.y : H.cl_c.x; h.cy = l ? H.cy.y : H.cy.x; h.l_c = l ? H.l_c.y : H.l_c.x;
        h.m_c = l ? H.m_c.y : H.m_c.x; h.n_c = l ? H.n_c.y : H.n_c.x; h.thrust_max_m = l ? H.thrust_max_m.y : H.thrust_max_m.x;
        float x[12], time = 0.0f;
#pragma unroll
        for (int i = 0; i < 12; ++i) x[i] = l ? s[i].y : s[i].x;
        repaired += rk4_step_f32(P, h, x, time, dt);
#pragma unroll
        for (int i = 0; i < 12; ++i) { if (l) s[i].y = x[i]; else s[i].x = x[i]; }
    }
    return repaired;
}

// rk4_step_f32 for the two lanes; returns the number of repaired (non-finite) state words of both.
__device__ __forceinline__ int rk4_step_f32x2(const DynParams<float>& P, const HeldControlsF32x2& H, v2 (&s)[12], float dt) {
    using X = P2;
    // The one-aircraft code tests every stage's angles against the SFU range (|x| <= 12, FastF32::attitude_trig) -- a
    // branch in front of each stage's MUFU block, which also fences the instruction scheduler.  Here ONE test per substep
    // bounds all four stages: a stage angle is angle + a k with a <= dt, |k[6..8]| <= (1 + 2 / cos(max_pitch)) max|rate|
    // (:463-471 with the pitch clipped for tan / cos) and a stage rate is at most max|rate| + dt max_angular_acceleration
    // (:485-487).  Beyond the bound the substep runs the guarded one-aircraft code; within it both give the same bits.
    {
        const float ang0 = fmaxf(fmaxf(fabsf(s[6].x), fabsf(s[7].x)), fabsf(s[8].x));
        const float ang1 = fmaxf(fmaxf(fabsf(s[6].y), fabsf(s[7].y)), fabsf(s[8].y));
        const float rate0 = fmaxf(fmaxf(fabsf(s[9].x), fabsf(s[10].x)), fabsf(s[11].x));
        const float rate1 = fmaxf(fmaxf(fabsf(s[9].y), fabsf(s[10].y)), fabsf(s[11].y));
        const float gain = dt * P.euler_rate_gain, dacc = dt * P.max_angular_acceleration;
        const float b0 = fmaf(gain, rate0 + dacc, ang0), b1 = fmaf(gain, rate1 + dacc, ang1);
        if (!(fmaxf(b0, b1) <= 11.5f)) {
            v2 tmp[12];                       // only these copies live in local memory, not the loop-carried state / the held terms
            const HeldControlsF32x2 Hc = H;
#pragma unroll
            for (int i = 0; i < 12; ++i) tmp[i] = s[i];
            const int repaired = rk4_step_lanes_f32(P, Hc, tmp, dt);
#pragma unroll
            for (int i = 0; i < 12; ++i) s[i] = tmp[i];
            return repaired;
        }
    }
    v2 k[12], acc[12], t[12];
    const float half_dt = 0.5f * dt;
    v2 sphi, cphi, sth, cth, spsi, cpsi;
    X::attitude_trig(s[6], s[7], s[8], &sphi, &cphi, &sth, &cth, &spsi, &cpsi);
    dynamics_f32x2(P, H, s, sphi, cphi, sth, cth, spsi, cpsi, k);
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = k[i];
#if ACB_F32X2_UNROLL_STAGES
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int stage = 1; stage < 4; ++stage) {
        const float a = (stage == 3) ? dt : half_dt;
        // With the stages unrolled the compiler would know the weight, turn fma(1, k, acc) into k + acc and contract that
        // add with the multiply that produced k[8] (= yaw_num * rc) -- one rounding less than the one-aircraft kernel's
        // fma(wgt, k, acc) with its run-time wgt.  The weights come from the parameter block (constant bank), which
        // neither nvvm nor ptxas can fold, so the accumulation stays an FMA.
        const float wgt = (stage == 3) ? P.rk_one : P.rk_two;
#pragma unroll
        for (int i = 0; i < 12; ++i) t[i] = X::fma(X::bc(a), k[i], s[i]);
        X::attitude_trig(t[6], t[7], t[8], &sphi, &cphi, &sth, &cth, &spsi, &cpsi);
        dynamics_f32x2(P, H, t, sphi, cphi, sth, cth, spsi, cpsi, k);
#pragma unroll
        for (int i = 0; i < 12; ++i) acc[i] = X::fma(X::bc(wgt), k[i], acc[i]);
    }
    const float dt6 = dt * (1.0f / 6.0f);
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = X::fma(X::bc(dt6), acc[i], s[i]);

#pragma unroll
    for (int i = 3; i < 6; ++i) s[i] = X::clamp(s[i], -P.max_velocity, P.max_velocity);
    s[6] = X::wrap_pi(s[6]);
    s[7] = X::clamp(s[7], -P.max_pitch_rad, P.max_pitch_rad);
    s[8] = X::wrap_pi(s[8]);
#pragma unroll
    for (int i = 9; i < 12; ++i) s[i] = X::clamp(s[i], -P.max_rate_rad, P.max_rate_rad);
    if (s[2].x > 0.0f) { s[2].x = 0.0f; s[5].x = fmaxf(0.0f, s[5].x); }
    if (s[2].y > 0.0f) { s[2].y = 0.0f; s[5].y = fmaxf(0.0f, s[5].y); }
    v2 chk = X::bc(0.0f);
#pragma unroll
    for (int i = 0; i < 12; ++i)
        if (i < 3 || i == 6 || i == 8) chk = X::fma(X::bc(0.0f), s[i], chk);   // the clamped words are finite (physics_f32.cuh)
    int repaired = 0;
    if (chk.x != 0.0f || chk.y != 0.0f) {
        float tmp[12];
        if (chk.x != 0.0f) {
#pragma unroll
            for (int i = 0; i < 12; ++i) tmp[i] = s[i].x;
            repaired += repair_state_f32(tmp);
#pragma unroll
            for (int i = 0; i < 12; ++i) s[i].x = tmp[i];
        }
        if (chk.y != 0.0f) {
#pragma unroll
            for (int i = 0; i < 12; ++i) tmp[i] = s[i].y;
            repaired += repair_state_f32(tmp);
#pragma unroll
            for (int i = 0; i < 12; ++i) s[i].y = tmp[i];
        }
    }
    return repaired;
}

static __device__ __noinline__ int backend_step2_lanes(const DynParams<float>& P, v2* s, const float* c0, const float* c1, int K, float h) {
    int repaired = 0;
#pragma unroll 1
    for (int l = 0; l < 2; ++l) {
        float x[12], c[4], time = 0.0f;
#pragma unroll
        for (int i = 0; i < 12; ++i) x[i] = l ? s[i].y : s[i].x;
#pragma unroll
        for (int i = 0; i < 4; ++i) c[i] = l ? c1[i] : c0[i];
        repaired += backend_step(P, x, time, c, K, h);
#pragma unroll
        for (int i = 0; i < 12; ++i) { if (l) s[i].y = x[i]; else s[i].x = x[i]; }
    }
    return repaired;
}
__device__ __forceinline__ int backend_step2_cold(const DynParams<float>& P, v2 (&s)[12], const float (&c0)[4], const float (&c1)[4],
                                                  int K, float h) {
    v2 tmp[12];
    float a[4], b[4];
#pragma unroll
    for (int i = 0; i < 12; ++i) tmp[i] = s[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i] = c0[i]; b[i] = c1[i]; }
    const int r = backend_step2_lanes(P, tmp, a, b, K, h);
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = tmp[i];
    return r;
}

// K substeps of h under zero-order hold for the pair (simulation_backend.py:95-99).
// The simulation time (t += h per substep, :254) is advanced by the caller: nothing in the substep reads it.
__device__ __forceinline__ int backend_step2(const DynParams<float>& P, v2 (&s)[12], const float (&c0)[4],
                                             const float (&c1)[4], int K, float h) {
    // a NaN control in either lane: both lanes take the one-aircraft code for this control step (physics_f32.cuh,
    // backend_step_f32_nan_controls) -- the same bits as the one-aircraft kernel, nothing added to the hot loop
    {
        float chk = 0.0f;
#pragma unroll
        for (int i = 0; i < 4; ++i) chk = fmaf(0.0f, c0[i], fmaf(0.0f, c1[i], chk));
        if (chk != 0.0f) return backend_step2_cold(P, s, c0, c1, K, h);
    }
    const HeldControlsF32x2 H = hold_controls2(P, c0, c1);
    int repaired = 0;
#pragma unroll 1
    for (int k = 0; k < K; ++k) repaired += rk4_step_f32x2(P, H, s, h);
    return repaired;
}

}  // namespace acb
