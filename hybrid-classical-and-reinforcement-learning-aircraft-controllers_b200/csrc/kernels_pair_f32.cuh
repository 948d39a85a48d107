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
    const int64_t n = B.f.n;
    const int64_t pair_idx = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    const int64_t block_first = first + (int64_t)blockIdx.x * (2 * BLOCK);
    const int64_t i = first + 2 * pair_idx;
    const bool valid = pair_idx < pairs;
    float st[ACB200_STATS_FIELDS];
#pragma unroll
    for (int k = 0; k < ACB200_STATS_FIELDS; ++k) st[k] = 0.0f;

    if (valid) {
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
        if (B.f.controls) {      // the held surfaces leave now (a reset below zeroes its lane's words again)
#pragma unroll
            for (int k = 0; k < 4; ++k) st2(B.f.controls, k, n, i, c[0][k], c[1][k]);
        }
        float* const stash = stage + threadIdx.x * (2 * ACB200_OBS_DIM) + 24;
#pragma unroll
        for (int k = 0; k < 4; ++k) *reinterpret_cast<v2*>(stash + 2 * k) = make_float2(act[0][k], act[1][k]);
        *reinterpret_cast<v2*>(stash + 8) = make_float2(residual_bonus[0], residual_bonus[1]);
        *reinterpret_cast<v2*>(stash + 10) = time2;
        const int repaired = backend_step2(P, s2, c[0], c[1], E.K, E.h);
        time2 = *reinterpret_cast<const v2*>(stash + 10);
        // (C) reward, termination, reset and scatter, ONE LANE AT A TIME in a rolled loop: the two lanes' bookkeeping words
        // together (2 x ~45 registers + 2 observation rows) would be the register peak of the kernel, above the substep
        // loop's 139; per lane they fit under it, the code is half as long, and the 4-byte accesses of the two passes
        // meet in the same 32-byte sectors.  The lane's observation row goes straight into the staging buffer.
        st[14] = (float)repaired;
        st[15] = 2.0f;
        for (int k = 0; k < E.K; ++k) time2 = P2::add(time2, P2::bc(E.h));       // t += dt once per substep (:254)
#pragma unroll 1
        for (int l = 0; l < 2; ++l) {
            const int64_t il = i + l;
            EnvRegs<float> e;
            float act[4];
            {   // parked before the loop in floats 24..35 of the thread's staging rows (lane 1's row is written last)
                const float* stash = stage + threadIdx.x * (2 * ACB200_OBS_DIM) + 24;
#pragma unroll
                for (int k = 0; k < 4; ++k) act[k] = stash[2 * k + l];
            }
            const float residual_bonus = stage[threadIdx.x * (2 * ACB200_OBS_DIM) + 32 + l];
#pragma unroll
            for (int k = 0; k < 12; ++k) e.s[k] = lane(s2[k], l);
            e.time = lane(time2, l);
#pragma unroll
            for (int k = 0; k < 3; ++k) e.rate_cmd[k] = ldf(B.rate_cmd, k, n, il);
#pragma unroll
            for (int k = 0; k < 4; ++k) e.prev_action[k] = ldf(B.prev_action, k, n, il);
#pragma unroll
            for (int k = 0; k < 3; ++k) { e.prev_err[k] = ldf(B.rew_mem, k, n, il); e.sign_changes[k] = ldf(B.rew_mem, 3 + k, n, il); }
            e.settle_timer = ldf(B.rew_mem, 6, n, il);
#pragma unroll
            for (int k = 0; k < 7; ++k) e.sched[k] = ldf(B.sched, k, n, il);
            e.env_time = B.env_time[il];
            e.step_count = B.step_count[il];
            e.episode_idx = B.episode_idx[il];
            e.is_settled = B.is_settled[il];
            e.command_type = B.command_type[il];
            e.difficulty = B.difficulty[il];
            e.ep_return = B.ep_return[il];
            derived_outputs(e.s, e.d);

            float comp[8], reward, o[18];
            bool terminated, truncated;
            env_after_physics(E, e, act, residual_bonus, comp, reward, terminated, truncated, il);
            env_observe(e, o);
            B.reward[il] = reward;
            if (B.reward_f32) B.reward_f32[il] = reward;
            B.terminated[il] = (uint8_t)terminated;
            B.truncated[il] = (uint8_t)truncated;
            if (B.components) {
#pragma unroll
                for (int k = 0; k < 8; ++k) stf(B.components, k, n, il, comp[k]);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) st[6 + k] += comp[k];
            if (terminated || truncated) {
                st[0] += 1.0f; st[1] += e.ep_return; st[2] += e.ep_return * e.ep_return; st[3] += (float)e.step_count;
                st[4] += (terminated && !truncated) ? 1.0f : 0.0f;
                st[5] += truncated ? 1.0f : 0.0f;
                if (E.auto_reset) {
                    // SB3 VecEnv semantics: keep the terminal observation, reset, hand back the reset observation
                    if (B.terminal_obs) {
#pragma unroll
                        for (int k = 0; k < ACB200_OBS_DIM; ++k) B.terminal_obs[il * ACB200_OBS_DIM + k] = o[k];
                    }
                    env_reset_one(E, il, e);
                    env_observe(e, o);
#pragma unroll
                    for (int k = 0; k < ACB200_PID_FIELDS; ++k) stf(B.f.pid, k, n, il, 0.0f);   // agent.reset()
                    if (B.f.controls) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) stf(B.f.controls, k, n, il, 0.0f);
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < 12; ++k) stf(B.f.state, k, n, il, e.s[k]);
            stf(B.f.state, 12, n, il, e.time);
#pragma unroll
            for (int k = 0; k < 4; ++k) stf(B.f.derived, k, n, il, e.d[k]);
#pragma unroll
            for (int k = 0; k < 3; ++k) stf(B.rate_cmd, k, n, il, e.rate_cmd[k]);
#pragma unroll
            for (int k = 0; k < 4; ++k) stf(B.prev_action, k, n, il, e.prev_action[k]);
#pragma unroll
            for (int k = 0; k < 3; ++k) { stf(B.rew_mem, k, n, il, e.prev_err[k]); stf(B.rew_mem, 3 + k, n, il, e.sign_changes[k]); }
            stf(B.rew_mem, 6, n, il, e.settle_timer);
            if (e.step_count == 0) {   // the schedule only changes at a reset
#pragma unroll
                for (int k = 0; k < 7; ++k) stf(B.sched, k, n, il, e.sched[k]);
            }
            B.env_time[il] = e.env_time;
            B.step_count[il] = e.step_count;
            B.episode_idx[il] = e.episode_idx;
            B.is_settled[il] = e.is_settled;
            B.ep_return[il] = e.ep_return;
            float* row = stage + threadIdx.x * (2 * ACB200_OBS_DIM) + l * ACB200_OBS_DIM;
#pragma unroll
            for (int k = 0; k < ACB200_OBS_DIM; ++k) row[k] = o[k];
        }
    }

    // observations: float[n][18] row-major; the block's rows are one contiguous run of up to 18 KB, which ONE bulk copy of the
    // TMA engine sends (cp.async.bulk shared -> global; rows is even, so the run is a multiple of 16 bytes and starts on one)
    // instead of nine float4 load / store pairs per thread
    const int64_t end = first + 2 * pairs;
    const int64_t rows = (end - block_first < 2 * BLOCK) ? (end - block_first) : 2 * BLOCK;
    const int count = (int)rows * ACB200_OBS_DIM;            // a multiple of 4
    float* dst = B.obs + block_first * ACB200_OBS_DIM;
    const bool bulk = ((uintptr_t)dst & 15) == 0;            // block-uniform
    if (bulk) stagedetail::fence_async_smem();                // the rows written above, for the engine
    __syncthreads();
    if (bulk) {
        if (threadIdx.x == 0 && count > 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(stagedetail::smem_u32(stage)),
                         "r"((uint32_t)count * 4u) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    } else {
        for (int k = threadIdx.x; k < count; k += BLOCK) dst[k] = stage[k];
    }

    if (B.stats) block_accumulate_env_stats<float, BLOCK>(st, B.stats);
    if (bulk && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the stage must outlive the read
}

template <int BLOCK, int MIN_BLOCKS>
__global__ void __launch_bounds__(BLOCK, MIN_BLOCKS)
env_step_pair_kernel(const __grid_constant__ DynParams<float> P, const __grid_constant__ CtrlParams<float> C,
                     const __grid_constant__ EnvConst<float> E, EnvPtrs<float> B, const float* __restrict__ action, int64_t first,
                     int64_t pairs) {
    env_step_pair_body<BLOCK>(P, C, E, B, action, first, pairs);
}
// Block shape of the pair kernel: threads per block x resident blocks per SM the register allocation aims at.
// ACB200_PAIR_VARIANT picks the other compiled shape (measurements).  With the post-physics part run one lane at a time the
// kernel needs 148 registers uncapped and fits 128 without a spill, so the default is 4 blocks x 128 threads = 16 warps per
// SM: as fast as the 12-warp shape at K = 20 (0.4726 vs 0.4731 ms per 2^20 aircraft) and faster than the one-aircraft kernel
// at K = 1 (0.1263 vs 0.1283 ms); profiles/r01g_pair_variants.md.
#define ACB_PAIR_VARIANTS(X) X(0, 128, 3) X(1, 128, 4)
static inline int pair_variant(int) {
    const char* v = getenv("ACB200_PAIR_VARIANT");
    return (v && v[0] >= '0' && v[0] <= '1') ? v[0] - '0' : 1;
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

// ---- the fleet step (backend.set_controls + backend.step, with or without an agent in front) for aircraft pairs ----------
// LEVEL 0: in = clipped-on-entry surface commands (step_surface_kernel); ACB200_LEVEL_RATE / _ATTITUDE: in = the agent's
// command (step_cascade_kernel).  Same lane arithmetic as those kernels; the K substeps on packed FP32.
template <int LEVEL>
__global__ void __launch_bounds__(kBlock, 4)
step_fleet_pair_kernel(const __grid_constant__ DynParams<float> P, const __grid_constant__ CtrlParams<float> C, FleetPtrs<float> f,
                       const float* __restrict__ in, PidDt dt_ctrl, int K, float h, float* __restrict__ surfaces_out, int64_t pairs) {
    const int64_t n = f.n;
    const int64_t pair_idx = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (pair_idx >= pairs) return;
    const int64_t i = 2 * pair_idx;
    v2 s2[12], time2;
#pragma unroll
    for (int k = 0; k < 12; ++k) s2[k] = ld2(f.state, k, n, i);
    time2 = ld2(f.state, 12, n, i);
    float cm[2][4], surf[2][4], c[2][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const v2 a = ld2(in, k, n, i); cm[0][k] = a.x; cm[1][k] = a.y; }
    if (LEVEL == 0) {
#pragma unroll
        for (int l = 0; l < 2; ++l) clip_controls(cm[l], c[l]);
    } else {
        PidMem rate[2][3], ang[2][3];
        load_pid2(f.pid, n, i, 0, rate[0], rate[1]);
        if (LEVEL == ACB200_LEVEL_ATTITUDE) load_pid2(f.pid, n, i, 1, ang[0], ang[1]);
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            float sl[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) sl[k] = lane(s2[k], l);
            if (LEVEL == ACB200_LEVEL_ATTITUDE) attitude_agent(C, ang[l], rate[l], cm[l][0], cm[l][1], cm[l][2], cm[l][3], sl, dt_ctrl, surf[l]);
            else rate_agent(C, rate[l], cm[l][0], cm[l][1], cm[l][2], cm[l][3], sl[9], sl[10], sl[11], dt_ctrl, surf[l]);
            clip_controls(surf[l], c[l]);
        }
        if (LEVEL == ACB200_LEVEL_ATTITUDE) store_pid2(f.pid, n, i, 1, ang[0], ang[1]);
        store_pid2(f.pid, n, i, 0, rate[0], rate[1]);
        if (surfaces_out) {
#pragma unroll
            for (int k = 0; k < 4; ++k) st2(surfaces_out, k, n, i, surf[0][k], surf[1][k]);
        }
    }
    if (f.controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) st2(f.controls, k, n, i, c[0][k], c[1][k]);
    }
    backend_step2(P, s2, c[0], c[1], K, h);
    for (int k = 0; k < K; ++k) time2 = P2::add(time2, P2::bc(h));            // t += dt once per substep (:254)
#pragma unroll
    for (int k = 0; k < 12; ++k) *reinterpret_cast<v2*>(f.state + (int64_t)k * n + i) = s2[k];
    *reinterpret_cast<v2*>(f.state + 12 * n + i) = time2;
    if (f.derived) {
        float d[2][4];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            float sl[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) sl[k] = lane(s2[k], l);
            derived_outputs(sl, d[l]);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) st2(f.derived, k, n, i, d[0][k], d[1][k]);
    }
}

// ---- levels 2 and 1 (HSAAgent / MissionPlanner + WaypointAgent on top of the cascade) for aircraft pairs ------------------
// The guidance and the nine fp32 PIDs run per lane (the one-aircraft code of agents.cuh, one lane at a time), the K
// substeps packed.  A pair with an aircraft whose mission is complete or completes in this step (the one-aircraft kernel
// leaves that aircraft untouched) takes the one-aircraft code for both lanes, out of line.
template <int LEVEL>
static __device__ __noinline__ void guided_step_one_cold(const DynParams<float>& P, const CtrlParams<float>& C, const FleetPtrs<float>& f,
                                                         float* pid_outer, const float* cmd, const MissionPtrs<float>& M, PidDt dt_ctrl,
                                                         int K, float h, float* surfaces_out, float* hsa_cmd_out, int64_t i, float* ms_out) {
    float ms[kMissionStats];
#pragma unroll
    for (int k = 0; k < kMissionStats; ++k) ms[k] = 0.0f;
    guided_step_one<float, LEVEL>(P, C, f, pid_outer, cmd, M, dt_ctrl, K, h, surfaces_out, hsa_cmd_out, i, ms);
#pragma unroll
    for (int k = 0; k < kMissionStats; ++k) ms_out[k] += ms[k];
}

#ifndef ACB_GUIDED_PAIR_MIN_BLOCKS
#define ACB_GUIDED_PAIR_MIN_BLOCKS 3
#endif
template <int LEVEL>
__device__ __forceinline__ void step_guided_pair_body(const DynParams<float>& P, const CtrlParams<float>& C, const FleetPtrs<float>& f,
                                                      float* __restrict__ pid_outer, const float* __restrict__ cmd,
                                                      const MissionPtrs<float>& M, PidDt dt_ctrl, int K, float h,
                                                      float* __restrict__ surfaces_out, float* __restrict__ hsa_cmd_out, int64_t pairs,
                                                      float (&ms)[kMissionStats]) {
    const int64_t n = f.n;
    const int64_t pair_idx = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (pair_idx >= pairs) return;
    const int64_t i = 2 * pair_idx;
    v2 s2[12], time2;
#pragma unroll
    for (int k = 0; k < 12; ++k) s2[k] = ld2(f.state, k, n, i);
    time2 = ld2(f.state, 12, n, i);
    int2 idx2 = make_int2(0, 0);
    if (LEVEL == ACB200_LEVEL_WAYPOINT) {
        idx2 = *reinterpret_cast<const int2*>(M.wp_index + i);
        bool leaves = false;
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            const int idx = l ? idx2.y : idx2.x;
            if (idx >= M.n_waypoints) { leaves = true; continue; }
            const float dx = M.waypoints[idx * 4 + 0] - lane(s2[0], l), dy = M.waypoints[idx * 4 + 1] - lane(s2[1], l),
                        dz = M.waypoints[idx * 4 + 2] - lane(s2[2], l);
            if (Math<float>::norm3(dx, dy, dz) < M.acceptance_radius && idx + 1 >= M.n_waypoints && !M.loop) leaves = true;
        }
        if (leaves) {
            float tmp[kMissionStats];
#pragma unroll
            for (int k = 0; k < kMissionStats; ++k) tmp[k] = 0.0f;
            guided_step_one_cold<LEVEL>(P, C, f, pid_outer, cmd, M, dt_ctrl, K, h, surfaces_out, hsa_cmd_out, i, tmp);
            guided_step_one_cold<LEVEL>(P, C, f, pid_outer, cmd, M, dt_ctrl, K, h, surfaces_out, hsa_cmd_out, i + 1, tmp);
#pragma unroll
            for (int k = 0; k < kMissionStats; ++k) ms[k] = tmp[k];
            return;
        }
    }
    // All of the pair's controller memory is requested at once (8 bytes per field and thread: one memory round trip for both
    // lanes); the guidance and the nine PIDs then run one lane at a time on their halves.
    v2 c2[4], pidv[ACB200_PID_FIELDS], outv[9], surf2[4], hc2[3];
#pragma unroll
    for (int k = 0; k < ACB200_PID_FIELDS; ++k) pidv[k] = ld2(f.pid, k, n, i);
#pragma unroll
    for (int k = 0; k < 9; ++k) outv[k] = ld2(pid_outer, k, n, i);
    if (LEVEL != ACB200_LEVEL_WAYPOINT) {
#pragma unroll
        for (int k = 0; k < 3; ++k) hc2[k] = ld2(cmd, k, n, i);
    }
    // the derived outputs the agents read are loaded when the fleet keeps them (guided_front, kernels.cuh): the same words
    // the one-aircraft kernel reads, so the two kernels cannot drift apart over how an inlined copy of derived_outputs rounds
    v2 dv[4];
    if (f.derived) {
#pragma unroll
        for (int k = 0; k < 4; ++k) dv[k] = ld2(f.derived, k, n, i);
    }
#pragma unroll 1
    for (int l = 0; l < 2; ++l) {
        const int64_t il = i + l;
        float sl[12], d[4], hc[3], surf[4], c[4];
#pragma unroll
        for (int k = 0; k < 12; ++k) sl[k] = lane(s2[k], l);
        if (f.derived) {
#pragma unroll
            for (int k = 0; k < 4; ++k) d[k] = lane(dv[k], l);
        } else {
            derived_outputs(sl, d);
        }
        if (LEVEL == ACB200_LEVEL_WAYPOINT) {
            int32_t idx = l ? idx2.y : idx2.x;
            float wp[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) wp[k] = M.waypoints[idx * 4 + k];
            // MissionPlanner.update (mission_planner.py:128-171): inside the acceptance radius -> next waypoint
            bool advanced = false, lapped = false;
            if (Math<float>::norm3(wp[0] - sl[0], wp[1] - sl[1], wp[2] - sl[2]) < M.acceptance_radius) {
                idx += 1;
                advanced = true;
                if (idx >= M.n_waypoints && M.loop) { idx = 0; lapped = true; }
                M.wp_index[il] = idx;
#pragma unroll
                for (int k = 0; k < 4; ++k) wp[k] = M.waypoints[idx * 4 + k];
            }
            if (M.stats) mission_stats_one(M, idx, wp, sl, d, advanced, lapped, ms);
            waypoint_guidance(C, M.guidance, wp, sl, d, hc);
#pragma unroll
            for (int k = 0; k < 3; ++k) { if (l) hc2[k].y = hc[k]; else hc2[k].x = hc[k]; }
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) hc[k] = lane(hc2[k], l);
        }
        PidMem rate[3], ang[3], outer[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            rate[a] = PidMem{lane(pidv[a * 3], l), lane(pidv[a * 3 + 1], l), lane(pidv[a * 3 + 2], l)};
            ang[a] = PidMem{lane(pidv[9 + a * 3], l), lane(pidv[9 + a * 3 + 1], l), lane(pidv[9 + a * 3 + 2], l)};
            outer[a] = PidMem{lane(outv[a * 3], l), lane(outv[a * 3 + 1], l), lane(outv[a * 3 + 2], l)};
        }
        hsa_agent(C, outer, ang, rate, hc[0], hc[1], hc[2], sl, d, dt_ctrl, surf);
        clip_controls(surf, c);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float r3[3] = {rate[a].integral, rate[a].error_prev, rate[a].derivative_filtered};
            const float a3[3] = {ang[a].integral, ang[a].error_prev, ang[a].derivative_filtered};
            const float o3[3] = {outer[a].integral, outer[a].error_prev, outer[a].derivative_filtered};
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                if (l) { pidv[a * 3 + j].y = r3[j]; pidv[9 + a * 3 + j].y = a3[j]; outv[a * 3 + j].y = o3[j]; }
                else { pidv[a * 3 + j].x = r3[j]; pidv[9 + a * 3 + j].x = a3[j]; outv[a * 3 + j].x = o3[j]; }
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (l) { c2[k].y = c[k]; surf2[k].y = surf[k]; } else { c2[k].x = c[k]; surf2[k].x = surf[k]; }
        }
    }
#pragma unroll
    for (int k = 0; k < ACB200_PID_FIELDS; ++k) *reinterpret_cast<v2*>(f.pid + (int64_t)k * n + i) = pidv[k];
#pragma unroll
    for (int k = 0; k < 9; ++k) *reinterpret_cast<v2*>(pid_outer + (int64_t)k * n + i) = outv[k];
    if (LEVEL == ACB200_LEVEL_WAYPOINT && hsa_cmd_out) {
#pragma unroll
        for (int k = 0; k < 3; ++k) *reinterpret_cast<v2*>(hsa_cmd_out + (int64_t)k * n + i) = hc2[k];
    }
    if (surfaces_out) {
#pragma unroll
        for (int k = 0; k < 4; ++k) *reinterpret_cast<v2*>(surfaces_out + (int64_t)k * n + i) = surf2[k];
    }
    if (f.controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) *reinterpret_cast<v2*>(f.controls + (int64_t)k * n + i) = c2[k];
    }
    {
        const float c0[4] = {c2[0].x, c2[1].x, c2[2].x, c2[3].x}, c1[4] = {c2[0].y, c2[1].y, c2[2].y, c2[3].y};
        backend_step2(P, s2, c0, c1, K, h);
    }
    for (int k = 0; k < K; ++k) time2 = P2::add(time2, P2::bc(h));            // t += dt once per substep (:254)
#pragma unroll
    for (int k = 0; k < 12; ++k) *reinterpret_cast<v2*>(f.state + (int64_t)k * n + i) = s2[k];
    *reinterpret_cast<v2*>(f.state + 12 * n + i) = time2;
    if (f.derived) {
        v2 d2[4];
#pragma unroll 1
        for (int l = 0; l < 2; ++l) {
            float sl[12], d[4];
#pragma unroll
            for (int k = 0; k < 12; ++k) sl[k] = lane(s2[k], l);
            derived_outputs(sl, d);
#pragma unroll
            for (int k = 0; k < 4; ++k) { if (l) d2[k].y = d[k]; else d2[k].x = d[k]; }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) *reinterpret_cast<v2*>(f.derived + (int64_t)k * n + i) = d2[k];
    }
}

template <int LEVEL>
__global__ void __launch_bounds__(kBlock, ACB_GUIDED_PAIR_MIN_BLOCKS)
step_guided_pair_kernel(const __grid_constant__ DynParams<float> P, const __grid_constant__ CtrlParams<float> C, FleetPtrs<float> f,
                        float* __restrict__ pid_outer, const float* __restrict__ cmd, MissionPtrs<float> M, PidDt dt_ctrl, int K,
                        float h, float* __restrict__ surfaces_out, float* __restrict__ hsa_cmd_out, int64_t pairs) {
    float ms[kMissionStats];
#pragma unroll
    for (int k = 0; k < kMissionStats; ++k) ms[k] = 0.0f;
    step_guided_pair_body<LEVEL>(P, C, f, pid_outer, cmd, M, dt_ctrl, K, h, surfaces_out, hsa_cmd_out, pairs, ms);
    if (LEVEL == ACB200_LEVEL_WAYPOINT && M.stats) block_accumulate_mission(ms, M.stats);
}

// Whole fleets only (n even, 8-byte aligned fields); otherwise the one-aircraft kernels run.
static inline bool fleet_pair_ok(const FleetPtrs<float>& f, const void* in, const void* out, int K, const void* x0 = nullptr,
                                 const void* x1 = nullptr, const void* x2 = nullptr) {
    const char* pair_env = getenv("ACB200_ENV_PAIR");
    if (pair_env && pair_env[0] == '0') return false;
    if (pair_variant(K) < 0 || f.n < 2 || (f.n & 1)) return false;
    const void* w8[] = {f.state, f.controls, f.derived, f.pid, in, out, x0, x1, x2};
    for (const void* p : w8)
        if (p && ((uintptr_t)p & 7)) return false;
    return true;
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
