// kernels.cuh -- the __global__ kernels: one thread owns one aircraft; HBM is touched only at the
// control-step boundary through coalesced structure-of-arrays loads and stores; everything between
// (K RK4 substeps x 4 derivative evaluations, PID, reward, termination, reset) runs out of registers.
#pragma once
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "env.cuh"
#include "kernels.h"

namespace acb {

constexpr int kBlock = 128;

template <typename T> __device__ __forceinline__ T ldf(const T* base, int f, int64_t n, int64_t i) { return base[(int64_t)f * n + i]; }
template <typename T> __device__ __forceinline__ void stf(T* base, int f, int64_t n, int64_t i, T v) { base[(int64_t)f * n + i] = v; }

__device__ __forceinline__ void load_pid(const float* pid, int64_t n, int64_t i, int loop, PidMem (&m)[3]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int b = (loop * 3 + a) * 3;
        m[a].integral = ldf(pid, b + 0, n, i);
        m[a].error_prev = ldf(pid, b + 1, n, i);
        m[a].derivative_filtered = ldf(pid, b + 2, n, i);
    }
}
__device__ __forceinline__ void store_pid(float* pid, int64_t n, int64_t i, int loop, const PidMem (&m)[3]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int b = (loop * 3 + a) * 3;
        stf(pid, b + 0, n, i, m[a].integral);
        stf(pid, b + 1, n, i, m[a].error_prev);
        stf(pid, b + 2, n, i, m[a].derivative_filtered);
    }
}

// ---- Simplified6DOF.reset for a masked subset (simplified_6dof.py:189-212) ------------------------
// Resident blocks per SM the register allocation of the fp32 build aims at (measured on B200, r01d; the fp64 parity
// build keeps the compiler's own choice).  env_step: 6 blocks = 80 registers, 8 bytes of spill: +3.5 % at K=1, +0.3 % at
// K=20 over 96 registers (8 blocks = 64 registers spills 156 bytes and loses 2 %).  step_guided (K=1 missions, the most
// memory-bound kernel): 8 blocks = 64 registers: +14 % over 126 registers (6 blocks: -11 %, it spills 224 bytes there).
#ifndef ACB_GUIDED_MIN_BLOCKS
#define ACB_GUIDED_MIN_BLOCKS 8
#endif
#ifndef ACB_ENV_MIN_BLOCKS
#define ACB_ENV_MIN_BLOCKS 6
#endif
template <typename T> struct MinBlocks {
    static constexpr int env = sizeof(T) == 4 ? ACB_ENV_MIN_BLOCKS : 1;
    static constexpr int guided = sizeof(T) == 4 ? ACB_GUIDED_MIN_BLOCKS : 1;
    static constexpr int cascade = sizeof(T) == 4 ? 4 : 1;      // step_surface / step_cascade: 128 registers
};

template <typename T>
__global__ void __launch_bounds__(kBlock) fleet_reset_kernel(FleetPtrs<T> f, const uint8_t* __restrict__ mask,
                                                             const T* __restrict__ init_state) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= f.n || (mask && !mask[i])) return;
    T s[12], time = T(0), d[4];
    if (init_state) {
#pragma unroll
        for (int k = 0; k < 12; ++k) s[k] = ldf(init_state, k, f.n, i);
        time = ldf(init_state, 12, f.n, i);
    } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) s[k] = T(0);
        s[2] = T(-100); s[3] = T(20);
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) stf(f.state, k, f.n, i, s[k]);
    stf(f.state, 12, f.n, i, time);
    if (f.controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(f.controls, k, f.n, i, T(0));
    }
    if (f.derived) {
        derived_outputs(s, d);
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(f.derived, k, f.n, i, d[k]);
    }
    if (f.pid) {
#pragma unroll
        for (int k = 0; k < ACB200_PID_FIELDS; ++k) stf(f.pid, k, f.n, i, 0.0f);
    }
}

// ---- K-A: backend.set_controls + backend.step ------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kBlock, MinBlocks<T>::cascade) step_surface_kernel(const __grid_constant__ DynParams<T> P, FleetPtrs<T> f,
                                                              const T* __restrict__ ctrl_in, int K, T h) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= f.n) return;
    T s[12], time, cin[4], c[4], d[4];
#pragma unroll
    for (int k = 0; k < 12; ++k) s[k] = ldf(f.state, k, f.n, i);
    time = ldf(f.state, 12, f.n, i);
#pragma unroll
    for (int k = 0; k < 4; ++k) cin[k] = ldf(ctrl_in, k, f.n, i);
    clip_controls(cin, c);
    backend_step(P, s, time, c, K, h);
#pragma unroll
    for (int k = 0; k < 12; ++k) stf(f.state, k, f.n, i, s[k]);
    stf(f.state, 12, f.n, i, time);
    if (f.controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(f.controls, k, f.n, i, c[k]);
    }
    if (f.derived) {
        derived_outputs(s, d);
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(f.derived, k, f.n, i, d[k]);
    }
}

// ---- K-B: agent.compute_action + backend.set_controls + backend.step ------------------------------
template <typename T, int LEVEL>
__global__ void __launch_bounds__(kBlock, MinBlocks<T>::cascade) step_cascade_kernel(const __grid_constant__ DynParams<T> P,
                                                              const __grid_constant__ CtrlParams<T> C, FleetPtrs<T> f,
                                                              const T* __restrict__ cmd, PidDt dt_ctrl, int K, T h,
                                                              T* __restrict__ surfaces_out) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= f.n) return;
    T s[12], time, cm[4], surf[4], c[4], d[4];
    PidMem rate[3], ang[3];
#pragma unroll
    for (int k = 0; k < 12; ++k) s[k] = ldf(f.state, k, f.n, i);
    time = ldf(f.state, 12, f.n, i);
#pragma unroll
    for (int k = 0; k < 4; ++k) cm[k] = ldf(cmd, k, f.n, i);
    load_pid(f.pid, f.n, i, 0, rate);
    if (LEVEL == ACB200_LEVEL_ATTITUDE) {
        load_pid(f.pid, f.n, i, 1, ang);
        attitude_agent(C, ang, rate, cm[0], cm[1], cm[2], cm[3], s, dt_ctrl, surf);
        store_pid(f.pid, f.n, i, 1, ang);
    } else {
        rate_agent(C, rate, cm[0], cm[1], cm[2], cm[3], s[9], s[10], s[11], dt_ctrl, surf);
    }
    store_pid(f.pid, f.n, i, 0, rate);
    clip_controls(surf, c);
    backend_step(P, s, time, c, K, h);
#pragma unroll
    for (int k = 0; k < 12; ++k) stf(f.state, k, f.n, i, s[k]);
    stf(f.state, 12, f.n, i, time);
    if (surfaces_out) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(surfaces_out, k, f.n, i, surf[k]);
    }
    if (f.controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(f.controls, k, f.n, i, c[k]);
    }
    if (f.derived) {
        derived_outputs(s, d);
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(f.derived, k, f.n, i, d[k]);
    }
}

// ---- levels 2 and 1: HSAAgent / (MissionPlanner + WaypointAgent) + set_controls + step ---------------------
__device__ __forceinline__ void load_outer(const float* p, int64_t n, int64_t i, PidMem (&m)[3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        m[k].integral = ldf(p, k * 3 + 0, n, i); m[k].error_prev = ldf(p, k * 3 + 1, n, i); m[k].derivative_filtered = ldf(p, k * 3 + 2, n, i);
    }
}
__device__ __forceinline__ void store_outer(float* p, int64_t n, int64_t i, const PidMem (&m)[3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        stf(p, k * 3 + 0, n, i, m[k].integral); stf(p, k * 3 + 1, n, i, m[k].error_prev); stf(p, k * 3 + 2, n, i, m[k].derivative_filtered);
    }
}

template <typename T> struct MissionPtrs {
    const T* waypoints;
    int32_t n_waypoints, guidance;
    int32_t* wp_index;
    int32_t loop;
    T acceptance_radius;
    double* stats;      // double[ACB200_MISSION_STATS_FIELDS] accumulated over calls; nullable
};

// Mission statistics (BASELINE config 5: "episode-stat allreduce"): what one aircraft adds in one control step, from the state
// the guidance sees (before the physics step).  Field order = ACB200_MISSION_STATS_* in acb200.h.
constexpr int kMissionStats = 10;
template <typename T>
__device__ __forceinline__ void mission_stats_one(const MissionPtrs<T>& M, int32_t idx, const T (&wp)[4], const T (&s)[12], const T (&d)[4],
                                                  bool advanced, bool lapped, T (&ms)[kMissionStats]) {
    using Mt = Math<T>;
    ms[0] += T(1);                                   // aircraft-steps flown
    ms[1] += advanced ? T(1) : T(0);                 // waypoints reached (MissionPlanner.update returned True)
    ms[2] += lapped ? T(1) : T(0);                   // laps closed (looping missions)
    // cross-track error: horizontal distance from the leg (previous waypoint -> active waypoint)
    int32_t prev = idx - 1;
    if (prev < 0) prev = M.loop ? M.n_waypoints - 1 : -1;
    T ct = T(0);
    if (prev >= 0) {
        const T ax = M.waypoints[prev * 4 + 0], ay = M.waypoints[prev * 4 + 1];
        const T bx = wp[0] - ax, by = wp[1] - ay, px = s[0] - ax, py = s[1] - ay;
        const T len2 = Mt::mad(bx, bx, by * by), cross = Mt::mad(px, by, -(py * bx));
        if (len2 > T(1e-12)) ct = Mt::kParity ? Mt::abs(Mt::div(cross, Mt::sqrt(len2))) : Mt::abs(cross) * (T)rsqrtf((float)len2);
    }
    ms[4] += ct; ms[5] = Mt::mad(ct, ct, ms[5]);
    ms[6] += Mt::abs(d[1] - (-wp[2]));               // |altitude - waypoint altitude|
    ms[7] += Mt::norm3(wp[0] - s[0], wp[1] - s[1], wp[2] - s[2]);
    ms[8] += d[0];                                   // airspeed
    ms[9] += (d[1] < T(5)) ? T(1) : T(0);            // below the env's crash floor (rate_env.py:437-460)
}

// Sum of every thread's mission contributions into the global vector.  Every thread of the block must call it.
// At K = 1 the guided kernel is issue-bound, so the reduction is written for instruction count: the five 0/1 counts go
// through the integer warp reduction (one REDUX each), four of the five float sums through a TRANSPOSED warp reduction
// (lanes split the fields between them while they fold: 2 + 1 + 3 = 6 shuffles for four values instead of 20), the fifth
// through a plain tree; warp results meet in shared memory without atomics and ten threads add them to the global
// (double) accumulator.  First version (a five-step tree per field, shared double atomics): +0.017 ms on a 0.106 ms step.
template <typename T>
__device__ __forceinline__ void block_accumulate_mission(const T (&v)[kMissionStats], double* __restrict__ global) {
    __shared__ T part[kBlock / 32][kMissionStats];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // counts: fields 0, 1, 2, 3, 9
    const int c0 = __reduce_add_sync(full, (int)v[0]), c1 = __reduce_add_sync(full, (int)v[1]), c2 = __reduce_add_sync(full, (int)v[2]),
              c3 = __reduce_add_sync(full, (int)v[3]), c9 = __reduce_add_sync(full, (int)v[9]);
    // float sums 4..7, transposed: after the xor-16 step a lane keeps two fields, after xor-8 one
    const bool hi = (lane & 16) != 0, mid = (lane & 8) != 0;
    T x0 = hi ? v[6] : v[4], x1 = hi ? v[7] : v[5];
    const T y0 = hi ? v[4] : v[6], y1 = hi ? v[5] : v[7];
    x0 += __shfl_xor_sync(full, y0, 16); x1 += __shfl_xor_sync(full, y1, 16);
    T z = mid ? x1 : x0;
    const T w = mid ? x0 : x1;
    z += __shfl_xor_sync(full, w, 8);
    z += __shfl_xor_sync(full, z, 4); z += __shfl_xor_sync(full, z, 2); z += __shfl_xor_sync(full, z, 1);
    // lanes 0 / 8 / 16 / 24 now hold the warp totals of fields 4 / 5 / 6 / 7
    T a = v[8];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) a += __shfl_xor_sync(full, a, off);
    if ((lane & 7) == 0) part[warp][4 + (lane >> 3)] = z;
    if (lane == 0) {
        part[warp][0] = (T)c0; part[warp][1] = (T)c1; part[warp][2] = (T)c2; part[warp][3] = (T)c3; part[warp][9] = (T)c9;
        part[warp][8] = a;
    }
    __syncthreads();
    if (threadIdx.x < kMissionStats) {
        double t = 0.0;
#pragma unroll
        for (int wv = 0; wv < kBlock / 32; ++wv) t += (double)part[wv][threadIdx.x];
        if (t != 0.0) atomicAdd(&global[threadIdx.x], t);
    }
}

// One aircraft's guided step in two halves (step_guided_kernel runs the block's statistics reduction between them, so that the
// ten statistics words are dead before the register-hungry half starts; guided_step_one below is the two back to back).
//
// guided_front: state and derived outputs in, MissionPlanner.update + mission statistics + WaypointAgent guidance (level 1) or
// the caller's HSA command (level 2).  Returns false when the aircraft takes no step (mission complete).
// The derived outputs the agents read (airspeed, altitude, heading: what get_state() hands them) are LOADED from f.derived
// when the fleet has that buffer: every entry point that writes the state refreshes it (acb200.h), so it holds
// derived_outputs(state) bit for bit and the 140 instructions of recomputing it are saved; with derived == NULL they are
// recomputed here.
// Where a guided step finds the words it reads AND writes back (state, derived outputs, the 27 words of controller memory):
// the fleet's own arrays (stride n = fleet size, index = aircraft) or the block's shared-memory stage (stride 128, index =
// thread; step_guided_staged_kernel).  Everything else (waypoint index, commands, optional outputs, held controls) is
// always addressed in the fleet with (gn, gi).
template <typename T> struct GuidedMem {
    T* state;
    T* derived;      // nullable in the fleet view
    float* pid;
    float* outer;
    int64_t n, i;
};

template <typename T, int LEVEL>
__device__ __forceinline__ bool guided_front(const CtrlParams<T>& C, const GuidedMem<T>& m, int64_t gn, int64_t gi,
                                             const T* __restrict__ cmd, const MissionPtrs<T>& M, T* __restrict__ hsa_cmd_out,
                                             T (&s)[12], T& time, T (&d)[4], T (&hc)[3], T (&ms)[kMissionStats]) {
#pragma unroll
    for (int k = 0; k < 12; ++k) s[k] = ldf(m.state, k, m.n, m.i);
    time = ldf(m.state, 12, m.n, m.i);
    if (m.derived) {
#pragma unroll
        for (int k = 0; k < 4; ++k) d[k] = ldf(m.derived, k, m.n, m.i);
    } else {
        derived_outputs(s, d);
    }
    if (LEVEL == ACB200_LEVEL_WAYPOINT) {
        int32_t idx = M.wp_index[gi];
        if (idx >= M.n_waypoints) return false;           // mission complete: examples/03_waypoint_square_demo.py:196-203 stops stepping
        T wp[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) wp[k] = M.waypoints[idx * 4 + k];
        // MissionPlanner.update (mission_planner.py:128-171): 3-D distance below the acceptance radius -> next waypoint
        bool advanced = false, lapped = false;
        if (Math<T>::norm3(wp[0] - s[0], wp[1] - s[1], wp[2] - s[2]) < M.acceptance_radius) {
            idx += 1;
            advanced = true;
            if (idx >= M.n_waypoints && M.loop) { idx = 0; lapped = true; }
            M.wp_index[gi] = idx;
            if (idx >= M.n_waypoints) { ms[1] += T(1); ms[3] += T(1); return false; }   // reached the last waypoint: mission complete
#pragma unroll
            for (int k = 0; k < 4; ++k) wp[k] = M.waypoints[idx * 4 + k];
        }
        if (M.stats) mission_stats_one(M, idx, wp, s, d, advanced, lapped, ms);
        waypoint_guidance(C, M.guidance, wp, s, d, hc);
        if (hsa_cmd_out) {
#pragma unroll
            for (int k = 0; k < 3; ++k) stf(hsa_cmd_out, k, gn, gi, hc[k]);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 3; ++k) hc[k] = ldf(cmd, k, gn, gi);
    }
    return true;
}

// guided_back: HSAAgent (levels 2-4: nine PIDs) on the command, set_controls, K substeps, state / controls / derived out.
template <typename T>
__device__ __forceinline__ void guided_back(const DynParams<T>& P, const CtrlParams<T>& C, const GuidedMem<T>& m, int64_t gn,
                                            int64_t gi, T* __restrict__ controls, PidDt dt_ctrl, int K, T h,
                                            T* __restrict__ surfaces_out, T (&s)[12], T time, const T (&d_in)[4], const T (&hc)[3]) {
    T surf[4], c[4], d[4];
    PidMem rate[3], ang[3], outer[3];
    load_pid(m.pid, m.n, m.i, 0, rate);
    load_pid(m.pid, m.n, m.i, 1, ang);
    load_outer(m.outer, m.n, m.i, outer);
    hsa_agent(C, outer, ang, rate, hc[0], hc[1], hc[2], s, d_in, dt_ctrl, surf);
    store_pid(m.pid, m.n, m.i, 0, rate);
    store_pid(m.pid, m.n, m.i, 1, ang);
    store_outer(m.outer, m.n, m.i, outer);
    clip_controls(surf, c);
    backend_step(P, s, time, c, K, h);
#pragma unroll
    for (int k = 0; k < 12; ++k) stf(m.state, k, m.n, m.i, s[k]);
    stf(m.state, 12, m.n, m.i, time);
    if (surfaces_out) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(surfaces_out, k, gn, gi, surf[k]);
    }
    if (controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(controls, k, gn, gi, c[k]);
    }
    if (m.derived) {
        derived_outputs(s, d);
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(m.derived, k, m.n, m.i, d[k]);
    }
}

// One aircraft's whole guided step (the cold path of the pair kernel).
template <typename T, int LEVEL>
__device__ __forceinline__ void guided_step_one(const DynParams<T>& P, const CtrlParams<T>& C, const FleetPtrs<T>& f,
                                                float* __restrict__ pid_outer, const T* __restrict__ cmd, const MissionPtrs<T>& M,
                                                PidDt dt_ctrl, int K, T h, T* __restrict__ surfaces_out, T* __restrict__ hsa_cmd_out,
                                                int64_t i, T (&ms)[kMissionStats]) {
    T s[12], time, d[4], hc[3];
    const GuidedMem<T> m{f.state, f.derived, f.pid, pid_outer, f.n, i};
    if (guided_front<T, LEVEL>(C, m, f.n, i, cmd, M, hsa_cmd_out, s, time, d, hc, ms))
        guided_back<T>(P, C, m, f.n, i, f.controls, dt_ctrl, K, h, surfaces_out, s, time, d, hc);
}

template <typename T, int LEVEL>
__global__ void __launch_bounds__(kBlock, MinBlocks<T>::guided) step_guided_kernel(const __grid_constant__ DynParams<T> P,
                                                             const __grid_constant__ CtrlParams<T> C, FleetPtrs<T> f,
                                                             float* __restrict__ pid_outer, const T* __restrict__ cmd,
                                                             MissionPtrs<T> M, PidDt dt_ctrl, int K, T h,
                                                             T* __restrict__ surfaces_out, T* __restrict__ hsa_cmd_out) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const GuidedMem<T> m{f.state, f.derived, f.pid, pid_outer, f.n, i};
    T s[12], time = T(0), d[4], hc[3];
    bool active = false;
    {
        T ms[kMissionStats];
#pragma unroll
        for (int k = 0; k < kMissionStats; ++k) ms[k] = T(0);
        if (i < f.n) active = guided_front<T, LEVEL>(C, m, f.n, i, cmd, M, hsa_cmd_out, s, time, d, hc, ms);
        if (LEVEL == ACB200_LEVEL_WAYPOINT && M.stats) block_accumulate_mission(ms, M.stats);   // uniform branch: whole block or nobody
    }
    if (active) guided_back<T>(P, C, m, f.n, i, f.controls, dt_ctrl, K, h, surfaces_out, s, time, d, hc);
}

// ---- the same step with the block's words staged through shared memory by the TMA engine (fp32, K small) ----------------
// At K = 1 the guided step is bound by issue slots, and around every global access two of its instructions are the 64-bit
// address (field-major arrays: base + (field n + aircraft) 4, an IADD3 pair per access): 420 of the 2 088 warp instructions
// of 32 aircraft-steps.  Here the four arrays the step reads AND writes back -- state [13][n], derived [4][n], rate + angle
// loop memory [18][n], outer loop memory [9][n] -- are described by 2-D tensor maps (inner dimension = aircraft, outer =
// field), and ONE thread per block asks the TMA engine for the block's box of each: 128 aircraft x all fields, landing as
// stage[field][thread] (cp.async.bulk.tensor.2d, completion on an mbarrier).  The threads read and write their column with
// immediate offsets (LDS / STS, no address arithmetic), and four tensor stores send the boxes back.  The engine clips the
// ragged last block itself (loads zero-fill, stores skip what lies beyond n).  Held controls, waypoint index and the
// optional outputs keep their direct accesses.  An aircraft that takes no step leaves its column as loaded, so its words
// are rewritten with the values they had.
// Needs: fleet size a multiple of 4 (row pitch a multiple of 16 bytes) and 16-byte aligned arrays, derived != NULL.
constexpr int kGuidedStageRows = 13 + 4 + ACB200_PID_FIELDS + 9;

struct GuidedMaps {
    CUtensorMap state, derived, pid, outer;
};

namespace stagedetail {
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// bounded: a protocol bug must end as a trapped kernel, not as a hung GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    long long t0 = 0;
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 0xFFFu) != 0) continue;
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > 4000000000LL) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int c0, int c1, uint32_t src) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                 ::"l"(map), "r"(c0), "r"(c1), "r"(src) : "memory");
}
__device__ __forceinline__ void bulk_commit_and_wait_read() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
}  // namespace stagedetail

template <int LEVEL>
__global__ void __launch_bounds__(kBlock, ACB_GUIDED_MIN_BLOCKS)
step_guided_staged_kernel(const __grid_constant__ DynParams<float> P, const __grid_constant__ CtrlParams<float> C,
                          const __grid_constant__ GuidedMaps maps, FleetPtrs<float> f, const float* __restrict__ cmd,
                          MissionPtrs<float> M, PidDt dt_ctrl, int K, float h, float* __restrict__ surfaces_out,
                          float* __restrict__ hsa_cmd_out) {
    namespace sd = stagedetail;
    __shared__ __align__(128) float stage[kGuidedStageRows][kBlock];
    __shared__ __align__(8) uint64_t full_bar;
    constexpr int rDer = 13, rPid = 17, rOut = 17 + ACB200_PID_FIELDS;
    const int first = (int)blockIdx.x * kBlock;                       // tensor coordinates are 32-bit: the launcher checks n < 2^31
    const int64_t i = (int64_t)first + threadIdx.x;
    const uint32_t bar = sd::smem_u32(&full_bar);
    if (threadIdx.x == 0) {
        sd::mbar_init(bar, 1);
        sd::mbar_expect_tx(bar, (uint32_t)sizeof(stage));             // the engine reports whole boxes, clipped or not
        sd::tma_load_2d(sd::smem_u32(&stage[0][0]), &maps.state, first, 0, bar);
        sd::tma_load_2d(sd::smem_u32(&stage[rDer][0]), &maps.derived, first, 0, bar);
        sd::tma_load_2d(sd::smem_u32(&stage[rPid][0]), &maps.pid, first, 0, bar);
        sd::tma_load_2d(sd::smem_u32(&stage[rOut][0]), &maps.outer, first, 0, bar);
    }
    __syncthreads();                                                  // the barrier is initialised before anybody polls it
    sd::mbar_wait(bar, 0);

    const GuidedMem<float> m{&stage[0][0], &stage[rDer][0], &stage[rPid][0], &stage[rOut][0], kBlock, threadIdx.x};
    float s[12], time = 0.0f, d[4], hc[3];
    bool active = false;
    {
        float ms[kMissionStats];
#pragma unroll
        for (int k = 0; k < kMissionStats; ++k) ms[k] = 0.0f;
        if (i < f.n) active = guided_front<float, LEVEL>(C, m, f.n, i, cmd, M, hsa_cmd_out, s, time, d, hc, ms);
        if (LEVEL == ACB200_LEVEL_WAYPOINT && M.stats) block_accumulate_mission(ms, M.stats);
    }
    if (active) guided_back<float>(P, C, m, f.n, i, f.controls, dt_ctrl, K, h, surfaces_out, s, time, d, hc);
    sd::fence_async_smem();                   // the columns written above, made visible to the TMA engine
    __syncthreads();
    if (threadIdx.x == 0) {
        sd::tma_store_2d(&maps.state, first, 0, sd::smem_u32(&stage[0][0]));
        sd::tma_store_2d(&maps.derived, first, 0, sd::smem_u32(&stage[rDer][0]));
        sd::tma_store_2d(&maps.pid, first, 0, sd::smem_u32(&stage[rPid][0]));
        sd::tma_store_2d(&maps.outer, first, 0, sd::smem_u32(&stage[rOut][0]));
        sd::bulk_commit_and_wait_read();      // the stage must outlive the engine's reads
    }
}

// 2-D tensor map of a field-major float array [rows][n] with boxes of box_cols aircraft x all rows: encoded once per
// (array, size, shape) and kept (the encoding is host arithmetic on the pointer and the extents; no driver state is attached).
static inline bool cached_tensor_map(const void* base, int64_t n, uint32_t rows, uint32_t box_cols, CUtensorMap* out) {
    typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    struct Entry { const void* base; int64_t n; uint32_t rows, box; CUtensorMap map; };
    static std::mutex mu;
    static std::vector<Entry> cache;
    static void* fn = nullptr;
    static bool unavailable = false;
    std::lock_guard<std::mutex> lock(mu);
    if (unavailable) return false;
    for (const Entry& e : cache)
        if (e.base == base && e.n == n && e.rows == rows && e.box == box_cols) { *out = e.map; return true; }
    if (!fn) {
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) { unavailable = true; return false; }
    }
    Entry e{base, n, rows, box_cols, {}};
    const cuuint64_t dims[2] = {(cuuint64_t)n, rows};
    const cuuint64_t strides[1] = {(cuuint64_t)n * sizeof(float)};
    const cuuint32_t box[2] = {box_cols, rows};
    const cuuint32_t estr[2] = {1, 1};
    if (((EncodeTiled)fn)(&e.map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    if (cache.size() >= 256) cache.erase(cache.begin());
    cache.push_back(e);
    *out = e.map;
    return true;
}
// What every staged array must satisfy: 16-byte aligned base, row pitch a multiple of 16 bytes, 32-bit coordinates.
static inline bool tma_stageable(const void* const* arrays, int count, int64_t n) {
    if ((n & 3) != 0 || n >= (int64_t(1) << 31)) return false;
    for (int a = 0; a < count; ++a)
        if (!arrays[a] || ((uintptr_t)arrays[a] & 15)) return false;
    return true;
}
static inline bool guided_maps_for(const FleetPtrs<float>& f, float* pid_outer, GuidedMaps* out) {
    return cached_tensor_map(f.state, f.n, 13, kBlock, &out->state) && cached_tensor_map(f.derived, f.n, 4, kBlock, &out->derived) &&
           cached_tensor_map(f.pid, f.n, ACB200_PID_FIELDS, kBlock, &out->pid) && cached_tensor_map(pid_outer, f.n, 9, kBlock, &out->outer);
}

// ---- env memory <-> registers ----------------------------------------------------------------------
template <typename T> __device__ __forceinline__ void env_gather(const EnvPtrs<T>& B, int64_t i, EnvRegs<T>& e) {
    const int64_t n = B.f.n;
#pragma unroll
    for (int k = 0; k < 12; ++k) e.s[k] = ldf(B.f.state, k, n, i);
    e.time = ldf(B.f.state, 12, n, i);
#pragma unroll
    for (int k = 0; k < 4; ++k) e.d[k] = ldf(B.f.derived, k, n, i);
    load_pid(B.f.pid, n, i, 0, e.rate);
    load_pid(B.f.pid, n, i, 1, e.ang);
#pragma unroll
    for (int k = 0; k < 3; ++k) e.rate_cmd[k] = ldf(B.rate_cmd, k, n, i);
#pragma unroll
    for (int k = 0; k < 4; ++k) e.prev_action[k] = ldf(B.prev_action, k, n, i);
#pragma unroll
    for (int k = 0; k < 3; ++k) { e.prev_err[k] = ldf(B.rew_mem, k, n, i); e.sign_changes[k] = ldf(B.rew_mem, 3 + k, n, i); }
    e.settle_timer = ldf(B.rew_mem, 6, n, i);
#pragma unroll
    for (int k = 0; k < 7; ++k) e.sched[k] = ldf(B.sched, k, n, i);
    e.env_time = B.env_time[i];
    e.step_count = B.step_count[i];
    e.episode_idx = B.episode_idx[i];
    e.is_settled = B.is_settled[i];
    e.command_type = B.command_type[i];
    e.difficulty = B.difficulty[i];
    e.ep_return = B.ep_return[i];
}

template <typename T> __device__ __forceinline__ void env_scatter(const EnvPtrs<T>& B, int64_t i, const EnvRegs<T>& e) {
    const int64_t n = B.f.n;
#pragma unroll
    for (int k = 0; k < 12; ++k) stf(B.f.state, k, n, i, e.s[k]);
    stf(B.f.state, 12, n, i, e.time);
#pragma unroll
    for (int k = 0; k < 4; ++k) stf(B.f.derived, k, n, i, e.d[k]);
    store_pid(B.f.pid, n, i, 0, e.rate);
    store_pid(B.f.pid, n, i, 1, e.ang);
#pragma unroll
    for (int k = 0; k < 3; ++k) stf(B.rate_cmd, k, n, i, e.rate_cmd[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) stf(B.prev_action, k, n, i, e.prev_action[k]);
#pragma unroll
    for (int k = 0; k < 3; ++k) { stf(B.rew_mem, k, n, i, e.prev_err[k]); stf(B.rew_mem, 3 + k, n, i, e.sign_changes[k]); }
    stf(B.rew_mem, 6, n, i, e.settle_timer);
#pragma unroll
    for (int k = 0; k < 7; ++k) stf(B.sched, k, n, i, e.sched[k]);
    B.env_time[i] = e.env_time;
    B.step_count[i] = e.step_count;
    B.episode_idx[i] = e.episode_idx;
    B.is_settled[i] = e.is_settled;
    B.ep_return[i] = e.ep_return;
}

// Observations are float[n][18] row-major (the gym / policy-GEMM layout): stage the block's rows in
// shared memory and write them out as one contiguous, fully coalesced run.
__device__ __forceinline__ void store_obs_block(float* __restrict__ obs, int64_t first, int64_t end, const float (&o)[18],
                                                bool valid, float* stage) {
    if (valid) {
#pragma unroll
        for (int k = 0; k < ACB200_OBS_DIM; ++k) stage[threadIdx.x * ACB200_OBS_DIM + k] = o[k];
    }
    __syncthreads();
    const int64_t rows = (end - first < kBlock) ? (end - first) : kBlock;
    const int count = (int)rows * ACB200_OBS_DIM;
    float* dst = obs + first * ACB200_OBS_DIM;   // first*18*4 B = blockIdx*9216 B: 16-byte aligned when obs is
    if ((count & 3) == 0 && ((uintptr_t)dst & 15) == 0) {
        for (int k = threadIdx.x; k < count / 4; k += kBlock)
            reinterpret_cast<float4*>(dst)[k] = reinterpret_cast<const float4*>(stage)[k];
    } else {
        for (int k = threadIdx.x; k < count; k += kBlock) dst[k] = stage[k];
    }
    __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(kBlock) env_reset_kernel(EnvPtrs<T> B, const __grid_constant__ EnvConst<T> E,
                                                           const uint8_t* __restrict__ mask) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= B.f.n || (mask && !mask[i])) return;
    EnvRegs<T> e;
    env_gather(B, i, e);
    env_reset_one(E, i, e);
    env_scatter(B, i, e);
    if (B.f.controls) {
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(B.f.controls, k, B.f.n, i, T(0));
    }
    float o[18];
    env_observe(e, o);
#pragma unroll
    for (int k = 0; k < ACB200_OBS_DIM; ++k) B.obs[i * ACB200_OBS_DIM + k] = o[k];   // masked resets are sparse: direct rows
}

// Sum of every thread's 16 statistics contributions (STATS order of acb200.h) into the global vector; every thread of the
// block must call it.  Written for instruction count (it is ~6 % of a K = 1 step as a five-step tree per field): the six
// integer-valued fields -- episodes, length, crashes, truncations, repairs, aircraft-steps -- go through the integer warp
// reduction (one REDUX each); the eight reward components through a TRANSPOSED warp reduction (lanes split the fields
// between them while they fold: 4 + 2 + 1 + 2 = 9 shuffles for eight values); return and return^2 as a transposed pair
// (5 shuffles).  14 shuffles instead of 80; warp results meet in shared memory without atomics.
template <typename T, int BLOCK>
__device__ __forceinline__ void block_accumulate_env_stats(const T (&v)[ACB200_STATS_FIELDS], double* __restrict__ global) {
    __shared__ T part[BLOCK / 32][ACB200_STATS_FIELDS];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = __reduce_add_sync(full, (int)v[0]), c3 = __reduce_add_sync(full, (int)v[3]), c4 = __reduce_add_sync(full, (int)v[4]),
              c5 = __reduce_add_sync(full, (int)v[5]), c14 = __reduce_add_sync(full, (int)v[14]), c15 = __reduce_add_sync(full, (int)v[15]);
    const bool b16 = (lane & 16) != 0, b8 = (lane & 8) != 0, b4 = (lane & 4) != 0;
    // components 6..13
    T x[4], z[2];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        x[q] = b16 ? v[10 + q] : v[6 + q];
        const T y = b16 ? v[6 + q] : v[10 + q];
        x[q] += __shfl_xor_sync(full, y, 16);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        z[r] = b8 ? x[2 + r] : x[r];
        const T w = b8 ? x[r] : x[2 + r];
        z[r] += __shfl_xor_sync(full, w, 8);
    }
    T a = b4 ? z[1] : z[0];
    const T b = b4 ? z[0] : z[1];
    a += __shfl_xor_sync(full, b, 4);
    a += __shfl_xor_sync(full, a, 2); a += __shfl_xor_sync(full, a, 1);
    // lanes 0, 4, ..., 28 hold the warp totals of fields 6 .. 13
    // return (1) and return^2 (2)
    T r1 = b16 ? v[2] : v[1];
    const T r2 = b16 ? v[1] : v[2];
    r1 += __shfl_xor_sync(full, r2, 16);
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) r1 += __shfl_xor_sync(full, r1, off);
    // lane 0: field 1, lane 16: field 2
    if ((lane & 3) == 0) part[warp][6 + (lane >> 2)] = a;
    if ((lane & 15) == 0) part[warp][1 + (lane >> 4)] = r1;
    if (lane == 0) {
        part[warp][0] = (T)c0; part[warp][3] = (T)c3; part[warp][4] = (T)c4; part[warp][5] = (T)c5; part[warp][14] = (T)c14;
        part[warp][15] = (T)c15;
    }
    __syncthreads();
    if (threadIdx.x < ACB200_STATS_FIELDS) {
        double t = 0.0;
#pragma unroll
        for (int wv = 0; wv < BLOCK / 32; ++wv) t += (double)part[wv][threadIdx.x];
        if (t != 0.0) atomicAdd(&global[threadIdx.x], t);
    }
}

// ---- K-C: controller + RateControlEnv.step + reward + termination + in-kernel auto-reset -----------
// Loads are staged so that only what the substep loop needs is live across it: (A) state + PID memory
// -> controller -> PID memory stored back; (B) K RK4 substeps on 12 states + held controls; (C) the
// env's bookkeeping words are loaded only now -> reward / termination / (cold) Philox reset -> scatter.
template <typename T>
__device__ __forceinline__ void env_step_body(const DynParams<T>& P, const CtrlParams<T>& C, const EnvConst<T>& E, const EnvPtrs<T>& B,
                                              const float* __restrict__ action, int64_t first, int64_t count) {
    // aircraft [first, first + count) of the fleet; n stays the field stride of the SoA buffers
    __shared__ __align__(16) float stage[kBlock * ACB200_OBS_DIM];
    const int64_t n = B.f.n;
    const int64_t block_first = first + (int64_t)blockIdx.x * kBlock;
    const int64_t i = block_first + threadIdx.x;
    const bool valid = i < first + count;

    float o[18];
    T st[ACB200_STATS_FIELDS];
#pragma unroll
    for (int k = 0; k < ACB200_STATS_FIELDS; ++k) st[k] = T(0);

    if (valid) {
        EnvRegs<T> e;
        // (A) controller
#pragma unroll
        for (int k = 0; k < 12; ++k) e.s[k] = ldf(B.f.state, k, n, i);
        e.time = ldf(B.f.state, 12, n, i);
        float act_in[4] = {0.0f, 0.0f, 0.0f, 0.0f}, act[4];
        if (action) {
            const float4 a4 = reinterpret_cast<const float4*>(action)[i];
            act_in[0] = a4.x; act_in[1] = a4.y; act_in[2] = a4.z; act_in[3] = a4.w;
        }
        T c[4], residual_bonus;
        {
            T att[4] = {T(0), T(0), T(0), T(0)}, rcmd[3] = {T(0), T(0), T(0)};
            PidMem rate[3], ang[3];
            if (E.controller != ACB200_CTRL_EXTERNAL) load_pid(B.f.pid, n, i, 0, rate);
            if (E.controller == ACB200_CTRL_ATTITUDE_PID) {
                load_pid(B.f.pid, n, i, 1, ang);
#pragma unroll
                for (int k = 0; k < 4; ++k) att[k] = ldf(B.att_cmd, k, n, i);
            } else if (E.controller != ACB200_CTRL_EXTERNAL) {
#pragma unroll
                for (int k = 0; k < 3; ++k) rcmd[k] = ldf(B.rate_cmd, k, n, i);
            }
            env_action(C, E, e.s, rate, ang, rcmd, att, act_in, act, c, residual_bonus);
            if (E.controller != ACB200_CTRL_EXTERNAL) store_pid(B.f.pid, n, i, 0, rate);
            if (E.controller == ACB200_CTRL_ATTITUDE_PID) store_pid(B.f.pid, n, i, 1, ang);
        }
        // (B) physics: rate_env.py:234-237
        const int repaired = backend_step(P, e.s, e.time, c, E.K, E.h);
        derived_outputs(e.s, e.d);
        // (C) the env's own memory
#pragma unroll
        for (int k = 0; k < 3; ++k) e.rate_cmd[k] = ldf(B.rate_cmd, k, n, i);
#pragma unroll
        for (int k = 0; k < 4; ++k) e.prev_action[k] = ldf(B.prev_action, k, n, i);
#pragma unroll
        for (int k = 0; k < 3; ++k) { e.prev_err[k] = ldf(B.rew_mem, k, n, i); e.sign_changes[k] = ldf(B.rew_mem, 3 + k, n, i); }
        e.settle_timer = ldf(B.rew_mem, 6, n, i);
#pragma unroll
        for (int k = 0; k < 7; ++k) e.sched[k] = ldf(B.sched, k, n, i);
        e.env_time = B.env_time[i];
        e.step_count = B.step_count[i];
        e.episode_idx = B.episode_idx[i];
        e.is_settled = B.is_settled[i];
        e.command_type = B.command_type[i];
        e.difficulty = B.difficulty[i];
        e.ep_return = B.ep_return[i];

        T comp[8], reward;
        bool terminated, truncated;
        env_after_physics(E, e, act, residual_bonus, comp, reward, terminated, truncated, i);
        env_observe(e, o);

        B.reward[i] = reward;
        if (B.reward_f32) B.reward_f32[i] = (float)reward;
        B.terminated[i] = (uint8_t)terminated;
        B.truncated[i] = (uint8_t)truncated;
        if (B.components) {
#pragma unroll
            for (int k = 0; k < 8; ++k) stf(B.components, k, n, i, comp[k]);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) st[6 + k] = comp[k];
        st[14] = (T)repaired;
        st[15] = T(1);
        if (terminated || truncated) {
            st[0] = T(1); st[1] = e.ep_return; st[2] = e.ep_return * e.ep_return; st[3] = (T)e.step_count;
            st[4] = (terminated && !truncated) ? T(1) : T(0);
            st[5] = truncated ? T(1) : T(0);
            if (E.auto_reset) {
                // SB3 VecEnv semantics: keep the terminal observation, reset, hand back the reset observation
                if (B.terminal_obs) {
#pragma unroll
                    for (int k = 0; k < ACB200_OBS_DIM; ++k) B.terminal_obs[i * ACB200_OBS_DIM + k] = o[k];
                }
                env_reset_one(E, i, e);
                env_observe(e, o);
#pragma unroll
                for (int k = 0; k < ACB200_PID_FIELDS; ++k) stf(B.f.pid, k, n, i, 0.0f);   // agent.reset()
#pragma unroll
                for (int k = 0; k < 4; ++k) c[k] = T(0);
            }
        }
        // scatter
#pragma unroll
        for (int k = 0; k < 12; ++k) stf(B.f.state, k, n, i, e.s[k]);
        stf(B.f.state, 12, n, i, e.time);
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(B.f.derived, k, n, i, e.d[k]);
        if (B.f.controls) {
#pragma unroll
            for (int k = 0; k < 4; ++k) stf(B.f.controls, k, n, i, c[k]);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) stf(B.rate_cmd, k, n, i, e.rate_cmd[k]);
#pragma unroll
        for (int k = 0; k < 4; ++k) stf(B.prev_action, k, n, i, e.prev_action[k]);
#pragma unroll
        for (int k = 0; k < 3; ++k) { stf(B.rew_mem, k, n, i, e.prev_err[k]); stf(B.rew_mem, 3 + k, n, i, e.sign_changes[k]); }
        stf(B.rew_mem, 6, n, i, e.settle_timer);
        if (e.step_count == 0) {   // the schedule only changes at a reset
#pragma unroll
            for (int k = 0; k < 7; ++k) stf(B.sched, k, n, i, e.sched[k]);
        }
        B.env_time[i] = e.env_time;
        B.step_count[i] = e.step_count;
        B.episode_idx[i] = e.episode_idx;
        B.is_settled[i] = e.is_settled;
        B.ep_return[i] = e.ep_return;
    }
    store_obs_block(B.obs, block_first, first + count, o, valid, stage);

    if (B.stats) block_accumulate_env_stats<T, kBlock>(st, B.stats);      // uniform branch: whole block or nobody
}

// Default: ONE parameter set per launch, in the constant bank (__grid_constant__).
template <typename T>
__global__ void __launch_bounds__(kBlock, MinBlocks<T>::env) env_step_kernel(const __grid_constant__ DynParams<T> P,
                                                          const __grid_constant__ CtrlParams<T> C,
                                                          const __grid_constant__ EnvConst<T> E, EnvPtrs<T> B,
                                                          const float* __restrict__ action, int64_t first, int64_t count) {
    env_step_body<T>(P, C, E, B, action, first, count);
}

// Per-aircraft parameters (second half): the same step with every aircraft's OWN DynParams block,
// folded on the host by make_dyn_params exactly as the constant-bank block is and stored field-major, table[word][n] --
// a warp reads 128 contiguous bytes per word.  The block is loaded once per env step into registers (the compiler drops
// the words the build does not read) and held across the K substeps; with ~45 live parameters on top of the state the
// register allocation is wider than the default kernel's, hence the separate launch bound.  Throughput: the loads are
// ~200 B per aircraft against 553 B of state traffic, and the kernel stays compute-bound at K = 20.
constexpr int kDynParamWords = (int)(sizeof(DynParams<float>) / sizeof(float));
static_assert(sizeof(DynParams<float>) == kDynParamWords * sizeof(float) && sizeof(DynParams<double>) == kDynParamWords * sizeof(double),
              "DynParams must be a plain block of T");
template <typename T>
__global__ void __launch_bounds__(kBlock, sizeof(T) == 4 ? 3 : 1) env_step_table_kernel(const T* __restrict__ table,
                                                          const __grid_constant__ CtrlParams<T> C,
                                                          const __grid_constant__ EnvConst<T> E, EnvPtrs<T> B,
                                                          const float* __restrict__ action, int64_t first, int64_t count) {
    int64_t i = first + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= first + count) i = first + count - 1;          // tail threads read a valid row; env_step_body masks them out
    DynParams<T> P;
    T* w = reinterpret_cast<T*>(&P);
#pragma unroll
    for (int k = 0; k < kDynParamWords; ++k) w[k] = table[(int64_t)k * B.f.n + i];
    env_step_body<T>(P, C, E, B, action, first, count);
}

}  // namespace acb
#ifdef ACB_PAIR_KERNEL
#include "kernels_pair_f32.cuh"   // throughput build: two aircraft per thread on packed FP32 instructions
#endif
namespace acb {

// ---- launchers -------------------------------------------------------------------------------------
static inline unsigned grid_for(int64_t n) { return (unsigned)((n + kBlock - 1) / kBlock); }

// Smallest k such that k successive fp64 additions of dt reach 0.2 (SettlingTimeBonus, rewards.py:207-211):
// 11 for dt = 0.02, because ten additions give 0.19999999999999998.
static inline int settle_steps_for(double dt) {
    double t = 0.0;
    for (int k = 1; k < 1000000; ++k) {
        t += dt;
        if (t >= 0.2) return k;
    }
    return 1000000;
}

template <typename T> static inline EnvConst<T> make_env_const(const acb200_env_config& cfg, int K, double h) {
    EnvConst<T> E;
    E.dt = (T)cfg.dt; E.h = (T)h; E.pid_throttle = (T)cfg.pid_throttle;
    E.dt_f32 = (float)cfg.dt; E.residual_scale = (float)cfg.residual_scale;
    E.pid_dt = make_pid_dt(E.dt_f32);
    E.K = K; E.max_steps = cfg.max_steps; E.controller = cfg.controller; E.auto_reset = cfg.auto_reset;
    E.settle_steps = settle_steps_for(cfg.dt);
    E.seed = cfg.seed; E.id_offset = cfg.id_offset;
    return E;
}

template <typename T>
cudaError_t Launch<T>::fleet_reset(const FleetPtrs<T>& f, const uint8_t* mask, const T* init_state, cudaStream_t st) {
    if (f.n <= 0) return cudaSuccess;
    fleet_reset_kernel<T><<<grid_for(f.n), kBlock, 0, st>>>(f, mask, init_state);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::step_surface(const acb200_aircraft_params& ap, const FleetPtrs<T>& f, const T* ctrl_in, int K, double h,
                                    cudaStream_t st) {
    if (f.n <= 0) return cudaSuccess;
#ifdef ACB_PAIR_KERNEL
    if constexpr (sizeof(T) == 4) {
        if (fleet_pair_ok(f, ctrl_in, nullptr, K)) {
            step_fleet_pair_kernel<0><<<grid_for(f.n / 2), kBlock, 0, st>>>(make_dyn_params<float>(ap), CtrlParams<float>{}, f, ctrl_in,
                                                                           make_pid_dt(0.0f), K, (float)h, nullptr, f.n / 2);
            return cudaGetLastError();
        }
    }
#endif
    step_surface_kernel<T><<<grid_for(f.n), kBlock, 0, st>>>(make_dyn_params<T>(ap), f, ctrl_in, K, (T)h);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::step_cascade(const acb200_aircraft_params& ap, const acb200_controller_params& cp, const FleetPtrs<T>& f,
                                    int level, const T* cmd, double dt_ctrl, int K, double h, T* surfaces_out, cudaStream_t st) {
    if (f.n <= 0) return cudaSuccess;
    const DynParams<T> P = make_dyn_params<T>(ap);
    const CtrlParams<T> C = make_ctrl_params<T>(cp);
#ifdef ACB_PAIR_KERNEL
    if constexpr (sizeof(T) == 4) {
        if (fleet_pair_ok(f, cmd, surfaces_out, K)) {
            if (level == ACB200_LEVEL_ATTITUDE)
                step_fleet_pair_kernel<ACB200_LEVEL_ATTITUDE><<<grid_for(f.n / 2), kBlock, 0, st>>>(P, C, f, cmd, make_pid_dt((float)dt_ctrl), K, (float)h,
                                                                                                   surfaces_out, f.n / 2);
            else
                step_fleet_pair_kernel<ACB200_LEVEL_RATE><<<grid_for(f.n / 2), kBlock, 0, st>>>(P, C, f, cmd, make_pid_dt((float)dt_ctrl), K, (float)h,
                                                                                               surfaces_out, f.n / 2);
            return cudaGetLastError();
        }
    }
#endif
    if (level == ACB200_LEVEL_ATTITUDE)
        step_cascade_kernel<T, ACB200_LEVEL_ATTITUDE><<<grid_for(f.n), kBlock, 0, st>>>(P, C, f, cmd, make_pid_dt((float)dt_ctrl), K, (T)h, surfaces_out);
    else
        step_cascade_kernel<T, ACB200_LEVEL_RATE><<<grid_for(f.n), kBlock, 0, st>>>(P, C, f, cmd, make_pid_dt((float)dt_ctrl), K, (T)h, surfaces_out);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::step_guided(const acb200_aircraft_params& ap, const acb200_controller_params& cp, const FleetPtrs<T>& f,
                                   float* pid_outer, const T* cmd, const acb200_mission* m, double dt_ctrl, int K, double h,
                                   T* surfaces_out, T* hsa_cmd_out, cudaStream_t st) {
    if (f.n <= 0) return cudaSuccess;
    const DynParams<T> P = make_dyn_params<T>(ap);
    const CtrlParams<T> C = make_ctrl_params<T>(cp);
    MissionPtrs<T> M{nullptr, 0, 0, nullptr, 0, T(0), nullptr};
    if (m) M = MissionPtrs<T>{(const T*)m->waypoints, m->n_waypoints, m->guidance, m->wp_index, m->loop, (T)m->acceptance_radius, m->stats};
#ifdef ACB_PAIR_KERNEL
    if constexpr (sizeof(T) == 4) {
        // measured on B200 (profiles/r01g_fleet_probe.json): with the nine PIDs and the guidance run per lane the pair kernel wins
        // from a few substeps on (HSA level, K = 10: 0.297 -> 0.250 ms per 2^20 aircraft) and loses at K = 1 (0.074 -> 0.088 ms)
        if (K >= 4 && fleet_pair_ok(f, cmd, surfaces_out, K, pid_outer, hsa_cmd_out, m ? m->wp_index : nullptr)) {
            if (m)
                step_guided_pair_kernel<ACB200_LEVEL_WAYPOINT><<<grid_for(f.n / 2), kBlock, 0, st>>>(P, C, f, pid_outer, cmd, M, make_pid_dt((float)dt_ctrl), K,
                                                                                                    (float)h, surfaces_out, hsa_cmd_out, f.n / 2);
            else
                step_guided_pair_kernel<ACB200_LEVEL_HSA><<<grid_for(f.n / 2), kBlock, 0, st>>>(P, C, f, pid_outer, cmd, M, make_pid_dt((float)dt_ctrl), K,
                                                                                               (float)h, surfaces_out, hsa_cmd_out, f.n / 2);
            return cudaGetLastError();
        }
    }
#endif
    if constexpr (sizeof(T) == 4) {
        const char* staged_env = getenv("ACB200_GUIDED_STAGED");                // "0": the direct kernel (A/B runs, tests)
        const bool staged_off = staged_env && staged_env[0] == '0';
        const void* a16[] = {f.state, f.derived, f.pid, pid_outer};
        const bool ok = !staged_off && tma_stageable(a16, 4, f.n);
        GuidedMaps maps;
        if (ok && guided_maps_for(f, pid_outer, &maps)) {
            if (m)
                step_guided_staged_kernel<ACB200_LEVEL_WAYPOINT><<<grid_for(f.n), kBlock, 0, st>>>(P, C, maps, f, cmd, M, make_pid_dt((float)dt_ctrl), K,
                                                                                                  (float)h, surfaces_out, hsa_cmd_out);
            else
                step_guided_staged_kernel<ACB200_LEVEL_HSA><<<grid_for(f.n), kBlock, 0, st>>>(P, C, maps, f, cmd, M, make_pid_dt((float)dt_ctrl), K,
                                                                                             (float)h, surfaces_out, hsa_cmd_out);
            return cudaGetLastError();
        }
    }
    if (m) {
        step_guided_kernel<T, ACB200_LEVEL_WAYPOINT><<<grid_for(f.n), kBlock, 0, st>>>(P, C, f, pid_outer, cmd, M, make_pid_dt((float)dt_ctrl), K, (T)h,
                                                                                       surfaces_out, hsa_cmd_out);
    } else {
        step_guided_kernel<T, ACB200_LEVEL_HSA><<<grid_for(f.n), kBlock, 0, st>>>(P, C, f, pid_outer, cmd, M, make_pid_dt((float)dt_ctrl), K, (T)h,
                                                                                  surfaces_out, hsa_cmd_out);
    }
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::env_reset(const EnvPtrs<T>& e, const acb200_env_config& cfg, const uint8_t* mask, cudaStream_t st) {
    if (e.f.n <= 0) return cudaSuccess;
    env_reset_kernel<T><<<grid_for(e.f.n), kBlock, 0, st>>>(e, make_env_const<T>(cfg, 1, cfg.dt), mask);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::env_step_table(const void* table, const acb200_controller_params& cp, const EnvPtrs<T>& e,
                                      const acb200_env_config& cfg, int K, double h, const float* action, int64_t first, int64_t count,
                                      cudaStream_t st) {
    if (count <= 0) return cudaSuccess;
    env_step_table_kernel<T><<<grid_for(count), kBlock, 0, st>>>((const T*)table, make_ctrl_params<T>(cp), make_env_const<T>(cfg, K, h), e,
                                                                 action, first, count);
    return cudaGetLastError();
}

template <typename T> int Launch<T>::dyn_param_words() { return kDynParamWords; }

template <typename T> void Launch<T>::fold_params(const acb200_aircraft_params& ap, T* out, int64_t stride) {
    const DynParams<T> d = make_dyn_params<T>(ap);
    const T* w = reinterpret_cast<const T*>(&d);
    for (int k = 0; k < kDynParamWords; ++k) out[(int64_t)k * stride] = w[k];
}

template <typename T>
cudaError_t Launch<T>::env_step(const acb200_aircraft_params& ap, const acb200_controller_params& cp, const EnvPtrs<T>& e,
                                const acb200_env_config& cfg, int K, double h, const float* action, int64_t first, int64_t count,
                                cudaStream_t st) {
    if (count <= 0) return cudaSuccess;
#ifdef ACB_PAIR_KERNEL
    if constexpr (sizeof(T) == 4) {
        // Default path of the throughput build: aircraft pairs on packed FP32 instructions; a last odd aircraft (and any
        // layout without 8-byte field alignment) goes through the one-aircraft-per-thread kernel, which computes the
        // same bits.  ACB200_ENV_PAIR=0 forces the scalar kernel (A/B measurements, the bit-identity test).
        const char* pair_env = getenv("ACB200_ENV_PAIR");      // read per launch: the bit-identity test flips it in-process
        const bool pair_enabled = !(pair_env && pair_env[0] == '0');
        if (pair_enabled && count >= 2 && pair_variant(K) >= 0 && pair_layout_ok(e, first)) {
            const int64_t pairs = count / 2;
            cudaError_t err = launch_env_step_pair(make_dyn_params<float>(ap), make_ctrl_params<float>(cp),
                                                   make_env_const<float>(cfg, K, h), e, action, first, pairs, st);
            if (err != cudaSuccess || !(count & 1)) return err;
            first += 2 * pairs;
            count = 1;
        }
    }
#endif
    env_step_kernel<T><<<grid_for(count), kBlock, 0, st>>>(make_dyn_params<T>(ap), make_ctrl_params<T>(cp),
                                                           make_env_const<T>(cfg, K, h), e, action, first, count);
    return cudaGetLastError();
}

template <typename T> cudaError_t Launch<T>::kernel_info(const char* name, KernelInfo* out) {
    const void* fn = nullptr;
    if (!strcmp(name, "step_surface")) fn = (const void*)step_surface_kernel<T>;
    else if (!strcmp(name, "step_cascade_rate")) fn = (const void*)step_cascade_kernel<T, ACB200_LEVEL_RATE>;
    else if (!strcmp(name, "step_cascade_attitude")) fn = (const void*)step_cascade_kernel<T, ACB200_LEVEL_ATTITUDE>;
    else if (!strcmp(name, "env_step")) fn = (const void*)env_step_kernel<T>;
    else if (!strcmp(name, "env_step_table")) fn = (const void*)env_step_table_kernel<T>;
#ifdef ACB_PAIR_KERNEL
    else if (!strcmp(name, "env_step_pair") && sizeof(T) == 4) fn = env_step_pair_entry();
    else if (!strcmp(name, "step_waypoint_pair") && sizeof(T) == 4) fn = (const void*)step_guided_pair_kernel<ACB200_LEVEL_WAYPOINT>;
    else if (!strcmp(name, "step_surface_pair") && sizeof(T) == 4) fn = (const void*)step_fleet_pair_kernel<0>;
    else if (!strcmp(name, "step_cascade_attitude_pair") && sizeof(T) == 4) fn = (const void*)step_fleet_pair_kernel<ACB200_LEVEL_ATTITUDE>;
#endif
    else if (!strcmp(name, "step_waypoint_staged") && sizeof(T) == 4) fn = (const void*)step_guided_staged_kernel<ACB200_LEVEL_WAYPOINT>;
    else if (!strcmp(name, "step_hsa_staged") && sizeof(T) == 4) fn = (const void*)step_guided_staged_kernel<ACB200_LEVEL_HSA>;
    else if (!strcmp(name, "step_hsa")) fn = (const void*)step_guided_kernel<T, ACB200_LEVEL_HSA>;
    else if (!strcmp(name, "step_waypoint")) fn = (const void*)step_guided_kernel<T, ACB200_LEVEL_WAYPOINT>;
    else if (!strcmp(name, "env_reset")) fn = (const void*)env_reset_kernel<T>;
    else if (!strcmp(name, "fleet_reset")) fn = (const void*)fleet_reset_kernel<T>;
    if (!fn) return cudaErrorInvalidValue;
    cudaFuncAttributes a;
    const cudaError_t err = cudaFuncGetAttributes(&a, fn);
    if (err != cudaSuccess) return err;
    out->regs = a.numRegs; out->local_bytes = (int)a.localSizeBytes; out->max_threads = a.maxThreadsPerBlock;
    return cudaSuccess;
}

}  // namespace acb
