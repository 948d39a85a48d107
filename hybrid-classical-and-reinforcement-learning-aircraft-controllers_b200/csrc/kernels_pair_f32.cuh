// kernels_pair_f32.cuh -- env_step for the throughput build with TWO aircraft per thread (aircraft 2t and 2t+1 of the
// range), the K substeps on packed FP32 instructions (physics_f32x2.cuh).  Everything outside the substep loop
// (controller, reward, termination, reset) is the scalar code of env.cuh run once per lane, so a lane computes
// exactly what env_step_kernel<float> computes for that aircraft -- bit for bit (tests/test_gpu_pair_kernel.py).
// Neighbouring aircraft are neighbouring words of every field-major array, so each field moves as ONE 8-byte
// access per thread (LDG.64 / STG.64, 256 contiguous bytes per warp): half the memory instructions as well.
// Included by kernels_f32.cu only.
#pragma once
#include "physics_f32x2.cuh"   // (included from kernels.cuh, after the one-aircraft kernels)

namespace acb {

// Resident blocks per SM the register allocation aims at: 3 blocks x 128 threads = 12 warps = 168 registers.
#ifndef ACB_PAIR_MIN_BLOCKS
#define ACB_PAIR_MIN_BLOCKS 3
#endif

__device__ __forceinline__ v2 ld2(const float* base, int f, int64_t n, int64_t i) {
    return *reinterpret_cast<const v2*>(base + (int64_t)f * n + i);
}
__device__ __forceinline__ void st2(float* base, int f, int64_t n, int64_t i, float a, float b) {
    *reinterpret_cast<v2*>(base + (int64_t)f * n + i) = make_float2(a, b);
}
__device__ __forceinline__ float lane(const v2& a, int l) { return l ? a.y : a.x; }

__device__ __forceinline__ void load_pid2(const float* pid, int64_t n, int64_t i, int loop, PidMem (&m0)[3], PidMem (&m1)[3]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int b = (loop * 3 + a) * 3;
        const v2 x = ld2(pid, b + 0, n, i), y = ld2(pid, b + 1, n, i), z = ld2(pid, b + 2, n, i);
        m0[a].integral = x.x; m0[a].error_prev = y.x; m0[a].derivative_filtered = z.x;
        m1[a].integral = x.y; m1[a].error_prev = y.y; m1[a].derivative_filtered = z.y;
    }
}
__device__ __forceinline__ void store_pid2(float* pid, int64_t n, int64_t i, int loop, const PidMem (&m0)[3], const PidMem (&m1)[3]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int b = (loop * 3 + a) * 3;
        st2(pid, b + 0, n, i, m0[a].integral, m1[a].integral);
        st2(pid, b + 1, n, i, m0[a].error_prev, m1[a].error_prev);
        st2(pid, b + 2, n, i, m0[a].derivative_filtered, m1[a].derivative_filtered);
    }
}

template <int BLOCK>
__device__ __forceinline__ void env_step_pair_body(const DynParams<float>& P, const CtrlParams<float>& C, const EnvConst<float>& E,
                                                   const EnvPtrs<float>& B, const float* __restrict__ action, int64_t first,
                                                   int64_t pairs) {
    // aircraft [first, first + 2 pairs) of the fleet; first and the field stride n are even (checked by the launcher)
    __shared__ __align__(16) float stage[2 * BLOCK * ACB200_OBS_DIM];
    __shared__ double block_stats[ACB200_STATS_FIELDS];
    const int64_t n = B.f.n;
    const int64_t pair_idx = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    const int64_t block_first = first + (int64_t)blockIdx.x * (2 * BLOCK);
    const int64_t i = first + 2 * pair_idx;
    const bool valid = pair_idx < pairs;
    if (threadIdx.x < ACB200_STATS_FIELDS) block_stats[threadIdx.x] = 0.0;

    float o[2][18];
    float st[ACB200_STATS_FIELDS];
#pragma unroll
    for (int k = 0; k < ACB200_STATS_FIELDS; ++k) st[k] = 0.0f;

    if (valid) {
        EnvRegs<float> e[2];
        v2 s2[12], time2;
        // (A) controller, once per lane
#pragma unroll
        for (int k = 0; k < 12; ++k) s2[k] = ld2(B.f.state, k, n, i);
        time2 = ld2(B.f.state, 12, n, i);
        float act[2][4], c[2][4], residual_bonus[2];
        {
            float act_in[2][4] = {{0.0f, 0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f, 0.0f}};
            if (action) {
                const float4 a0 = reinterpret_cast<const float4*>(action)[i], a1 = reinterpret_cast<const float4*>(action)[i + 1];
                act_in[0][0] = a0.x; act_in[0][1] = a0.y; act_in[0][2] = a0.z; act_in[0][3] = a0.w;
                act_in[1][0] = a1.x; act_in[1][1] = a1.y; act_in[1][2] = a1.z; act_in[1][3] = a1.w;
            }
            float att[2][4] = {{0.0f, 0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f, 0.0f}};
            float rcmd[2][3] = {{0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}};
            PidMem rate[2][3], ang[2][3];
            if (E.controller != ACB200_CTRL_EXTERNAL) load_pid2(B.f.pid, n, i, 0, rate[0], rate[1]);
            if (E.controller == ACB200_CTRL_ATTITUDE_PID) {
                load_pid2(B.f.pid, n, i, 1, ang[0], ang[1]);
#pragma unroll
                for (int k = 0; k < 4; ++k) { const v2 a = ld2(B.att_cmd, k, n, i); att[0][k] = a.x; att[1][k] = a.y; }
            } else if (E.controller != ACB200_CTRL_EXTERNAL) {
#pragma unroll
                for (int k = 0; k < 3; ++k) { const v2 a = ld2(B.rate_cmd, k, n, i); rcmd[0][k] = a.x; rcmd[1][k] = a.y; }
            }
#pragma unroll
            for (int l = 0; l < 2; ++l) {
                float sl[12];
#pragma unroll
                for (int k = 0; k < 12; ++k) sl[k] = lane(s2[k], l);
                env_action(C, E, sl, rate[l], ang[l], rcmd[l], att[l], act_in[l], act[l], c[l], residual_bonus[l]);
            }
            if (E.controller != ACB200_CTRL_EXTERNAL) store_pid2(B.f.pid, n, i, 0, rate[0], rate[1]);
            if (E.controller == ACB200_CTRL_ATTITUDE_PID) store_pid2(B.f.pid, n, i, 1, ang[0], ang[1]);
        }
        // (B) physics: rate_env.py:234-237, both lanes on packed instructions.  What the loop does not touch (the clipped
        // action, the residual bonus, the simulation time) waits in the thread's own rows of the observation staging
        // buffer, which is idle until the end of the step: 12 registers less across the loop.
        float* const stash = stage + threadIdx.x * (2 * ACB200_OBS_DIM);
#pragma unroll
        for (int k = 0; k < 4; ++k) *reinterpret_cast<v2*>(stash + 2 * k) = make_float2(act[0][k], act[1][k]);
        *reinterpret_cast<v2*>(stash + 8) = make_float2(residual_bonus[0], residual_bonus[1]);
        *reinterpret_cast<v2*>(stash + 10) = time2;
        const int repaired = backend_step2(P, s2, c[0], c[1], E.K, E.h);
#pragma unroll
        for (int k = 0; k < 4; ++k) { const v2 a = *reinterpret_cast<const v2*>(stash + 2 * k); act[0][k] = a.x; act[1][k] = a.y; }
        { const v2 a = *reinterpret_cast<const v2*>(stash + 8); residual_bonus[0] = a.x; residual_bonus[1] = a.y; }
        time2 = *reinterpret_cast<const v2*>(stash + 10);
        for (int k = 0; k < E.K; ++k) time2 = P2::add(time2, P2::bc(E.h));       // t += dt once per substep (:254)
        // (C) the env's own memory
        {
            v2 x;
#pragma unroll
            for (int k = 0; k < 3; ++k) { x = ld2(B.rate_cmd, k, n, i); e[0].rate_cmd[k] = x.x; e[1].rate_cmd[k] = x.y; }
#pragma unroll
            for (int k = 0; k < 4; ++k) { x = ld2(B.prev_action, k, n, i); e[0].prev_action[k] = x.x; e[1].prev_action[k] = x.y; }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                x = ld2(B.rew_mem, k, n, i); e[0].prev_err[k] = x.x; e[1].prev_err[k] = x.y;
                x = ld2(B.rew_mem, 3 + k, n, i); e[0].sign_changes[k] = x.x; e[1].sign_changes[k] = x.y;
            }
            x = ld2(B.rew_mem, 6, n, i); e[0].settle_timer = x.x; e[1].settle_timer = x.y;
#pragma unroll
            for (int k = 0; k < 7; ++k) { x = ld2(B.sched, k, n, i); e[0].sched[k] = x.x; e[1].sched[k] = x.y; }
            x = ld2(B.env_time, 0, n, i); e[0].env_time = x.x; e[1].env_time = x.y;
            x = ld2(B.difficulty, 0, n, i); e[0].difficulty = x.x; e[1].difficulty = x.y;
            x = ld2(B.ep_return, 0, n, i); e[0].ep_return = x.x; e[1].ep_return = x.y;
            const int2 sc = *reinterpret_cast<const int2*>(B.step_count + i);
            e[0].step_count = sc.x; e[1].step_count = sc.y;
            const uint2 ei = *reinterpret_cast<const uint2*>(B.episode_idx + i);
            e[0].episode_idx = ei.x; e[1].episode_idx = ei.y;
            const uchar2 is = *reinterpret_cast<const uchar2*>(B.is_settled + i);
            e[0].is_settled = is.x; e[1].is_settled = is.y;
            const uchar2 ct = *reinterpret_cast<const uchar2*>(B.command_type + i);
            e[0].command_type = ct.x; e[1].command_type = ct.y;
        }
        float reward[2], comp[2][8];
        bool terminated[2], truncated[2];
        st[14] = (float)repaired;
        st[15] = 2.0f;
#pragma unroll
        for (int l = 0; l < 2; ++l) {
#pragma unroll
            for (int k = 0; k < 12; ++k) e[l].s[k] = lane(s2[k], l);
            e[l].time = lane(time2, l);
            derived_outputs(e[l].s, e[l].d);
            env_after_physics(E, e[l], act[l], residual_bonus[l], comp[l], reward[l], terminated[l], truncated[l], i + l);
            env_observe(e[l], o[l]);
#pragma unroll
            for (int k = 0; k < 8; ++k) st[6 + k] += comp[l][k];
            if (terminated[l] || truncated[l]) {
                st[0] += 1.0f; st[1] += e[l].ep_return; st[2] += e[l].ep_return * e[l].ep_return; st[3] += (float)e[l].step_count;
                st[4] += (terminated[l] && !truncated[l]) ? 1.0f : 0.0f;
                st[5] += truncated[l] ? 1.0f : 0.0f;
                if (E.auto_reset) {
                    // SB3 VecEnv semantics: keep the terminal observation, reset, hand back the reset observation
                    if (B.terminal_obs) {
#pragma unroll
                        for (int k = 0; k < ACB200_OBS_DIM; ++k) B.terminal_obs[(i + l) * ACB200_OBS_DIM + k] = o[l][k];
                    }
                    env_reset_one(E, i + l, e[l]);
                    env_observe(e[l], o[l]);
#pragma unroll
                    for (int k = 0; k < ACB200_PID_FIELDS; ++k) stf(B.f.pid, k, n, i + l, 0.0f);   // agent.reset()
#pragma unroll
                    for (int k = 0; k < 4; ++k) c[l][k] = 0.0f;
                }
            }
        }
        // scatter, 8 bytes per field and thread
        *reinterpret_cast<v2*>(B.reward + i) = make_float2(reward[0], reward[1]);
        if (B.reward_f32) *reinterpret_cast<v2*>(B.reward_f32 + i) = make_float2(reward[0], reward[1]);
        *reinterpret_cast<uchar2*>(B.terminated + i) = make_uchar2((uint8_t)terminated[0], (uint8_t)terminated[1]);
        *reinterpret_cast<uchar2*>(B.truncated + i) = make_uchar2((uint8_t)truncated[0], (uint8_t)truncated[1]);
        if (B.components) {
#pragma unroll
            for (int k = 0; k < 8; ++k) st2(B.components, k, n, i, comp[0][k], comp[1][k]);
        }
#pragma unroll
        for (int k = 0; k < 12; ++k) st2(B.f.state, k, n, i, e[0].s[k], e[1].s[k]);
        st2(B.f.state, 12, n, i, e[0].time, e[1].time);
#pragma unroll
        for (int k = 0; k < 4; ++k) st2(B.f.derived, k, n, i, e[0].d[k], e[1].d[k]);
        if (B.f.controls) {
#pragma unroll
            for (int k = 0; k < 4; ++k) st2(B.f.controls, k, n, i, c[0][k], c[1][k]);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) st2(B.rate_cmd, k, n, i, e[0].rate_cmd[k], e[1].rate_cmd[k]);
#pragma unroll
        for (int k = 0; k < 4; ++k) st2(B.prev_action, k, n, i, e[0].prev_action[k], e[1].prev_action[k]);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            st2(B.rew_mem, k, n, i, e[0].prev_err[k], e[1].prev_err[k]);
            st2(B.rew_mem, 3 + k, n, i, e[0].sign_changes[k], e[1].sign_changes[k]);
        }
        st2(B.rew_mem, 6, n, i, e[0].settle_timer, e[1].settle_timer);
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            if (e[l].step_count == 0) {   // the schedule only changes at a reset
#pragma unroll
                for (int k = 0; k < 7; ++k) stf(B.sched, k, n, i + l, e[l].sched[k]);
            }
        }
        st2(B.env_time, 0, n, i, e[0].env_time, e[1].env_time);
        *reinterpret_cast<int2*>(B.step_count + i) = make_int2(e[0].step_count, e[1].step_count);
        *reinterpret_cast<uint2*>(B.episode_idx + i) = make_uint2(e[0].episode_idx, e[1].episode_idx);
        *reinterpret_cast<uchar2*>(B.is_settled + i) = make_uchar2(e[0].is_settled, e[1].is_settled);
        st2(B.ep_return, 0, n, i, e[0].ep_return, e[1].ep_return);
    }

    // observations: float[n][18] row-major; the thread's two rows are adjacent (36 floats = 9 x 16 bytes)
    if (valid) {
        float* row = stage + threadIdx.x * (2 * ACB200_OBS_DIM);
#pragma unroll
        for (int k = 0; k < ACB200_OBS_DIM; ++k) { row[k] = o[0][k]; row[ACB200_OBS_DIM + k] = o[1][k]; }
    }
    __syncthreads();
    {
        const int64_t end = first + 2 * pairs;
        const int64_t rows = (end - block_first < 2 * BLOCK) ? (end - block_first) : 2 * BLOCK;
        const int count = (int)rows * ACB200_OBS_DIM;        // rows is even: count is a multiple of 4
        float* dst = B.obs + block_first * ACB200_OBS_DIM;
        if (((uintptr_t)dst & 15) == 0) {
            for (int k = threadIdx.x; k < count / 4; k += BLOCK)
                reinterpret_cast<float4*>(dst)[k] = reinterpret_cast<const float4*>(stage)[k];
        } else {
            for (int k = threadIdx.x; k < count; k += BLOCK) dst[k] = stage[k];
        }
    }

    if (B.stats) {
#pragma unroll
        for (int k = 0; k < ACB200_STATS_FIELDS; ++k) {
            float v = st[k];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
            if ((threadIdx.x & 31) == 0) atomicAdd(&block_stats[k], (double)v);
        }
        __syncthreads();
        if (threadIdx.x < ACB200_STATS_FIELDS) atomicAdd(&B.stats[threadIdx.x], block_stats[threadIdx.x]);
    }
}

template <int BLOCK, int MIN_BLOCKS>
__global__ void __launch_bounds__(BLOCK, MIN_BLOCKS)
env_step_pair_kernel(const __grid_constant__ DynParams<float> P, const __grid_constant__ CtrlParams<float> C,
                     const __grid_constant__ EnvConst<float> E, EnvPtrs<float> B, const float* __restrict__ action, int64_t first,
                     int64_t pairs) {
    env_step_pair_body<BLOCK>(P, C, E, B, action, first, pairs);
}
// Block shape of the pair kernel: threads per block x resident blocks per SM the register allocation aims at.
// ACB200_PAIR_VARIANT picks one of the compiled shapes (measurements); the default is the fastest measured on B200.
// Measured on B200 (2^20 aircraft, profiles/r01g_pair_variants.md): at K = 20 substeps the 168-register shape without
// spills (12 warps per SM) is fastest; the 128-register shapes spill ~250 bytes and lose what the extra warps gain.  Below
// K = 4 the step is a memory-bound gather / scatter where the one-aircraft kernel (24 warps per SM) is as fast or
// faster, so the launcher keeps it there (pair_variant() < 0) unless ACB200_PAIR_VARIANT names a shape.
#define ACB_PAIR_VARIANTS(X) X(0, 128, 3) X(1, 128, 4) X(2, 64, 7)
static inline int pair_variant(int K) {
    const char* v = getenv("ACB200_PAIR_VARIANT");
    return (v && v[0] >= '0' && v[0] <= '2') ? v[0] - '0' : (K < 4 ? -1 : 0);
}
static inline cudaError_t launch_env_step_pair(const DynParams<float>& P, const CtrlParams<float>& C, const EnvConst<float>& E,
                                               const EnvPtrs<float>& e, const float* action, int64_t first, int64_t pairs,
                                               cudaStream_t st) {
    switch (pair_variant(E.K)) {
#define ACB_PAIR_CASE(id, block, minb)                                                                                        \
    case id:                                                                                                                  \
        env_step_pair_kernel<block, minb><<<(unsigned)((pairs + block - 1) / block), block, 0, st>>>(P, C, E, e, action, first, \
                                                                                                    pairs);                   \
        break;
        ACB_PAIR_VARIANTS(ACB_PAIR_CASE)
#undef ACB_PAIR_CASE
    }
    return cudaGetLastError();
}
static inline const void* env_step_pair_entry() {
    switch (pair_variant(20)) {
#define ACB_PAIR_CASE(id, block, minb) case id: return (const void*)env_step_pair_kernel<block, minb>;
        ACB_PAIR_VARIANTS(ACB_PAIR_CASE)
#undef ACB_PAIR_CASE
    }
    return nullptr;
}

// The pair kernel needs 8-byte access to every float field (first, n even; bases 8-byte aligned) and 2-byte access to the flags.
static inline bool pair_layout_ok(const EnvPtrs<float>& e, int64_t first) {
    if ((e.f.n & 1) || (first & 1)) return false;
    const void* w8[] = {e.f.state, e.f.controls, e.f.derived, e.f.pid, e.att_cmd, e.rate_cmd, e.prev_action, e.rew_mem, e.sched,
                        e.env_time, e.step_count, e.difficulty, e.episode_idx, e.ep_return, e.reward, e.components, e.reward_f32};
    for (const void* p : w8)
        if (p && ((uintptr_t)p & 7)) return false;
    const void* w2[] = {e.is_settled, e.command_type, e.terminated, e.truncated};
    for (const void* p : w2)
        if (p && ((uintptr_t)p & 1)) return false;
    return true;
}

}  // namespace acb
