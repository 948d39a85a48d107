// rollout.cu -- generalized advantage estimation over a device-resident rollout ("rollout buffer
// on device, GAE computed on device").
//
// Replaces stable-baselines3's RolloutBuffer.compute_returns_and_advantage, which the reference's trainers run on the
// host after every rollout (PPO / RecurrentPPO built at learned_controllers/utils/training_utils.py:112-160 and
// train_overnight.py:151-167).  Third-party arithmetic, restated from its published form:
//     for t = T-1 .. 0:  nnt   = 1 - (t == T-1 ? dones : episode_starts[t+1])
//                        nv    = (t == T-1 ? last_values : values[t+1])
//                        delta = rewards[t] + gamma * nv * nnt - values[t]
//                        gae   = delta + gamma * lambda * nnt * gae
//                        advantages[t] = gae;   returns[t] = gae + values[t]
// numpy evaluates this in float32 (float32 arrays, Python-float scalars are weak), left to right; the kernel keeps that
// order with explicitly rounded operations, so the result is bit-identical to numpy's.
// One thread per environment walks its column backwards; [T][n] row-major storage makes every access coalesced.
#include <string>

#include "ctx.h"

namespace {

__global__ void __launch_bounds__(256) gae_kernel(int T, int64_t n, const float* __restrict__ rewards, const float* __restrict__ values,
                                                  const uint8_t* __restrict__ episode_starts, const float* __restrict__ last_values,
                                                  const uint8_t* __restrict__ dones, float gamma, float gamma_lambda,
                                                  float* __restrict__ advantages, float* __restrict__ returns) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    float next_value = last_values[e];
    float nnt = 1.0f - (dones[e] ? 1.0f : 0.0f);
    float gae = 0.0f;
#pragma unroll 4
    for (int t = T - 1; t >= 0; --t) {
        const int64_t i = (int64_t)t * n + e;
        const float v = values[i], r = rewards[i];
        const float delta = __fsub_rn(__fadd_rn(r, __fmul_rn(__fmul_rn(gamma, next_value), nnt)), v);
        gae = __fadd_rn(delta, __fmul_rn(__fmul_rn(gamma_lambda, nnt), gae));
        advantages[i] = gae;
        returns[i] = __fadd_rn(gae, v);
        next_value = v;
        nnt = 1.0f - (episode_starts[i] ? 1.0f : 0.0f);
    }
}

}  // namespace

extern "C" int acb200_gae(acb200_ctx* ctx, int32_t T, int64_t n, const float* rewards, const float* values,
                          const uint8_t* episode_starts, const float* last_values, const uint8_t* dones, double gamma,
                          double gae_lambda, float* advantages, float* returns) {
    if (!ctx || !rewards || !values || !episode_starts || !last_values || !dones || !advantages || !returns) {
        if (ctx) ctx->error = "gae: NULL argument";
        return ACB200_ENULL;
    }
    if (T < 0 || n < 0 || !(gamma == gamma) || !(gae_lambda == gae_lambda)) {
        ctx->error = "gae: invalid size or coefficient";
        return ACB200_EVALUE;
    }
    if (T == 0 || n == 0) return ACB200_OK;
    cudaSetDevice(ctx->device);
    // numpy: gamma * next_values -> float32(gamma) * float32; gamma * gae_lambda is a Python (double) product first
    const float g = (float)gamma, gl = (float)(gamma * gae_lambda);
    gae_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(T, n, rewards, values, episode_starts, last_values, dones, g, gl,
                                                                       advantages, returns);
    ctx->launches += 1;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        ctx->error = std::string("gae_kernel: ") + cudaGetErrorString(e);
        return ACB200_ECUDA;
    }
    return ACB200_OK;
}
