// ctx.h -- the library context shared by the translation units behind the C ABI (capi.cu, policy.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/acb200.h"

struct acb200_ctx {
    int device;
    cudaStream_t stream;
    acb200_aircraft_params aircraft;
    acb200_controller_params controller;
    int64_t launches;
    std::string error;
    // env_step_host pipeline: worker streams and their events, created on first use
    static const int kPipe = 3;
    cudaStream_t pipe[kPipe] = {nullptr, nullptr, nullptr};
    cudaEvent_t pipe_done[kPipe] = {nullptr, nullptr, nullptr};
    cudaEvent_t fork = nullptr;
    int host_chunks = 0;   // 0 = automatic
    // parameter groups (acb200_set_param_groups): group g owns aircraft [group_bounds[g], group_bounds[g+1]) of the fleet
    std::vector<acb200_aircraft_params> groups;
    std::vector<int64_t> group_bounds;
    // per-aircraft parameter table (acb200_set_param_table): caller-owned device memory, table[words][table_n] of `real`
    const void* param_table = nullptr;
    int64_t table_n = 0;
    int32_t table_precision = 0;
    // NCCL communicator of the statistics all-reduce (comm.cu); opaque here, created by acb200_comm_init
    void* comm = nullptr;
    int comm_world = 0, comm_rank = 0;
    // device scratch owned by the handle (acb200_env_gather_done: one int per 1024 aircraft)
    int32_t* scratch_i32 = nullptr;
    int scratch_i32_len = 0;
};

// internal (not part of the C ABI): the worker streams of the host-step / rollout pipelines, created on first use (capi.cu)
cudaError_t acb200_ctx_pipe_streams(acb200_ctx* ctx);
// acb200_env_step_range on a given stream instead of the context's (capi.cu)
int acb200_env_step_range_on(acb200_ctx* ctx, const acb200_env* env, const acb200_env_config* cfg, const float* action, int64_t first,
                             int64_t count, cudaStream_t stream);
