// kernels.h -- host-side launch interface between the C ABI (capi.cu) and the two kernel
// translation units (kernels_f64.cu: parity build, -fmad=false; kernels_f32.cu: throughput build).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/acb200.h"

namespace acb {

template <typename T> struct FleetPtrs {
    int64_t n;
    T* state;
    T* controls;
    T* derived;
    float* pid;
};

template <typename T> struct EnvPtrs {
    FleetPtrs<T> f;
    T* att_cmd;
    T* rate_cmd;
    float* prev_action;
    T* rew_mem;
    T* sched;
    T* env_time;
    int32_t* step_count;
    uint8_t* is_settled;
    uint8_t* command_type;
    float* difficulty;
    uint32_t* episode_idx;
    T* ep_return;
    float* obs;
    T* reward;
    uint8_t* terminated;
    uint8_t* truncated;
    float* terminal_obs;
    T* components;
    double* stats;
    float* reward_f32;   // optional float copy of the reward for the host-facing step
};

struct KernelInfo {
    int regs, local_bytes, max_threads;
};

template <typename T> struct Launch {
    static cudaError_t fleet_reset(const FleetPtrs<T>& f, const uint8_t* mask, const T* init_state, cudaStream_t st);
    static cudaError_t step_surface(const acb200_aircraft_params& ap, const FleetPtrs<T>& f, const T* ctrl_in, int K, double h,
                                    cudaStream_t st);
    static cudaError_t step_cascade(const acb200_aircraft_params& ap, const acb200_controller_params& cp, const FleetPtrs<T>& f,
                                    int level, const T* cmd, double dt_ctrl, int K, double h, T* surfaces_out, cudaStream_t st);
    static cudaError_t step_guided(const acb200_aircraft_params& ap, const acb200_controller_params& cp, const FleetPtrs<T>& f,
                                   float* pid_outer, const T* cmd, const acb200_mission* m /* NULL: level 2 */, double dt_ctrl, int K,
                                   double h, T* surfaces_out, T* hsa_cmd_out, cudaStream_t st);
    static cudaError_t env_reset(const EnvPtrs<T>& e, const acb200_env_config& cfg, const uint8_t* mask, cudaStream_t st);
    static cudaError_t env_step(const acb200_aircraft_params& ap, const acb200_controller_params& cp, const EnvPtrs<T>& e,
                                const acb200_env_config& cfg, int K, double h, const float* action, int64_t first, int64_t count,
                                cudaStream_t st);
    // per-aircraft parameter table: table[dyn_param_words()][e.f.n] of T on the device, rows folded by fold_params
    static cudaError_t env_step_table(const void* table, const acb200_controller_params& cp, const EnvPtrs<T>& e,
                                      const acb200_env_config& cfg, int K, double h, const float* action, int64_t first, int64_t count,
                                      cudaStream_t st);
    static int dyn_param_words();
    static void fold_params(const acb200_aircraft_params& ap, T* out /* host, word k at out[k * stride] */, int64_t stride);
    static cudaError_t kernel_info(const char* name, KernelInfo* out);
};

extern template struct Launch<double>;
extern template struct Launch<float>;

}  // namespace acb
