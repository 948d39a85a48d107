// sensor.cu -- NoisySensorInterface.update for the whole fleet (interfaces/sensor.py:137-263 in the reference):
// measured = true + sigma * N(0,1) per channel, gyro bias added to the rates, then both bias vectors take one
// random-walk step.  One thread per aircraft, field-major loads/stores (coalesced), no generator state in HBM:
// the 20 normals of (seed, global id, tick) are a pure function (Philox4x32-10 + Box-Muller), so the noise does not
// depend on how the fleet is sharded.  Compiled without FMA contraction so the fp64 build follows the CPU restatement.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

#include "../../include/acb200.h"
#include "ctx.h"
#include "philox.cuh"

namespace {

constexpr int kBlock = 256;
constexpr uint32_t kSensorStream = 0x53454E00u;   // 'SEN\0': keeps these draws apart from the reset sampler's blocks

struct SensorSigma {
    double pos, vel, att, gyro, airspeed, altitude, gyro_walk, accel_walk;
};

template <typename T>
struct Normals;

// two standard normals from two 32-bit words: u1 = (w0 + 0.5) / 2^32 in (0,1), angle = 2 pi (w1 + 0.5) / 2^32
template <>
struct Normals<double> {
    static __device__ __forceinline__ void pair(uint32_t w0, uint32_t w1, double& z0, double& z1) {
        const double u1 = ((double)w0 + 0.5) * (1.0 / 4294967296.0);
        const double u2 = ((double)w1 + 0.5) * (1.0 / 4294967296.0);
        const double r = sqrt(-2.0 * log(u1));
        double s, c;
        sincospi(2.0 * u2, &s, &c);
        z0 = r * c;
        z1 = r * s;
    }
};
template <>
struct Normals<float> {
    static __device__ __forceinline__ void pair(uint32_t w0, uint32_t w1, float& z0, float& z1) {
        const float u1 = __fmaf_rn((float)w0, 1.0f / 4294967296.0f, 0.5f / 4294967296.0f);   // (0, 1]
        const float u2 = __fmaf_rn((float)w1, 1.0f / 4294967296.0f, 0.5f / 4294967296.0f);
        const float r = sqrtf(-2.0f * logf(u1));
        float s, c;
        sincospif(2.0f * u2, &s, &c);
        z0 = r * c;
        z1 = r * s;
    }
};

template <typename T>
__global__ void __launch_bounds__(kBlock) sensor_update_kernel(int64_t n, const T* __restrict__ state, const T* __restrict__ derived,
                                                               T* __restrict__ bias, T* __restrict__ m_state,
                                                               T* __restrict__ m_derived, SensorSigma sg, int enabled,
                                                               uint64_t seed, int64_t id_offset, uint32_t tick) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    T s[13], d[4];
#pragma unroll
    for (int f = 0; f < 13; ++f) s[f] = state[(int64_t)f * n + i];
#pragma unroll
    for (int f = 0; f < 4; ++f) d[f] = derived[(int64_t)f * n + i];
    if (enabled) {
        T b[6], z[20];
#pragma unroll
        for (int f = 0; f < 6; ++f) b[f] = bias[(int64_t)f * n + i];
        const acb::Philox g(seed, (uint64_t)(id_offset + i), tick);
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            uint32_t w[4];
            g.block(kSensorStream + k, w);
            Normals<T>::pair(w[0], w[1], z[4 * k + 0], z[4 * k + 1]);
            Normals<T>::pair(w[2], w[3], z[4 * k + 2], z[4 * k + 3]);
        }
        // draw order of NoisySensorInterface.update: position, velocity, attitude, gyro, airspeed, altitude, then the walks
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            s[a] = s[a] + (T)sg.pos * z[a];
            s[3 + a] = s[3 + a] + (T)sg.vel * z[3 + a];
            s[6 + a] = s[6 + a] + (T)sg.att * z[6 + a];
            s[9 + a] = (s[9 + a] + (T)sg.gyro * z[9 + a]) + b[a];      // the bias BEFORE this update's walk step
            b[a] = b[a] + (T)sg.gyro_walk * z[14 + a];
            b[3 + a] = b[3 + a] + (T)sg.accel_walk * z[17 + a];
        }
        d[0] = d[0] + (T)sg.airspeed * z[12];
        d[1] = d[1] + (T)sg.altitude * z[13];
#pragma unroll
        for (int f = 0; f < 6; ++f) bias[(int64_t)f * n + i] = b[f];
    }
#pragma unroll
    for (int f = 0; f < 13; ++f) m_state[(int64_t)f * n + i] = s[f];
#pragma unroll
    for (int f = 0; f < 4; ++f) m_derived[(int64_t)f * n + i] = d[f];
}

template <typename T>
__global__ void sensor_reset_kernel(int64_t n, T* __restrict__ bias, const uint8_t* __restrict__ mask) {
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n || (mask && !mask[i])) return;
#pragma unroll
    for (int f = 0; f < 6; ++f) bias[(int64_t)f * n + i] = (T)0;
}

int launch_check(acb200_ctx* ctx, const char* what) {
    ctx->launches += 1;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        ctx->error = std::string(what) + ": " + cudaGetErrorString(e);
        return ACB200_ECUDA;
    }
    return ACB200_OK;
}

}  // namespace

extern "C" void acb200_default_sensor_config(acb200_sensor_config* c) {
    if (!c) return;
    c->enabled = 1;
    c->_pad = 0;
    c->imu_gyro_stddev = 0.01;          // sensor.py:176-182
    c->imu_accel_stddev = 0.1;
    c->gps_position_stddev = 1.0;
    c->gps_velocity_stddev = 0.1;
    c->airspeed_stddev = 0.5;
    c->altitude_stddev = 0.5;
    c->attitude_stddev = 0.01;
    c->gyro_bias_walk_stddev = 0.0001;  // sensor.py:230-231 (literals there)
    c->accel_bias_walk_stddev = 0.001;
    c->seed = 0;
    c->id_offset = 0;
}

extern "C" int acb200_sensor_update(acb200_ctx* ctx, const acb200_fleet* f, const acb200_sensor_config* cfg, void* bias,
                                    uint32_t tick, void* measured_state, void* measured_derived) {
    if (!ctx) return ACB200_ENULL;
    if (!f || !cfg || !measured_state || !measured_derived || !f->state || !f->derived || (cfg->enabled && !bias)) {
        ctx->error = "sensor_update: NULL argument";
        return ACB200_ENULL;
    }
    if (f->n < 0 || (f->precision != 64 && f->precision != 32)) {
        ctx->error = "sensor_update: invalid fleet size or precision";
        return ACB200_EVALUE;
    }
    const double sd[9] = {cfg->imu_gyro_stddev, cfg->imu_accel_stddev, cfg->gps_position_stddev, cfg->gps_velocity_stddev,
                          cfg->airspeed_stddev, cfg->altitude_stddev, cfg->attitude_stddev, cfg->gyro_bias_walk_stddev,
                          cfg->accel_bias_walk_stddev};
    for (double v : sd)
        if (!(v >= 0.0) || v > 1.79e308) {          // Generator.normal raises ValueError for scale < 0 (and NaN)
            ctx->error = "sensor_update: standard deviations must be finite and >= 0";
            return ACB200_EVALUE;
        }
    if (f->n == 0) return ACB200_OK;
    cudaSetDevice(ctx->device);
    const SensorSigma sg = {cfg->gps_position_stddev, cfg->gps_velocity_stddev, cfg->attitude_stddev, cfg->imu_gyro_stddev,
                            cfg->airspeed_stddev,     cfg->altitude_stddev,     cfg->gyro_bias_walk_stddev,
                            cfg->accel_bias_walk_stddev};
    const unsigned grid = (unsigned)((f->n + kBlock - 1) / kBlock);
    if (f->precision == 64)
        sensor_update_kernel<double><<<grid, kBlock, 0, ctx->stream>>>(f->n, (const double*)f->state, (const double*)f->derived,
                                                                       (double*)bias, (double*)measured_state,
                                                                       (double*)measured_derived, sg, cfg->enabled, cfg->seed,
                                                                       cfg->id_offset, tick);
    else
        sensor_update_kernel<float><<<grid, kBlock, 0, ctx->stream>>>(f->n, (const float*)f->state, (const float*)f->derived,
                                                                      (float*)bias, (float*)measured_state, (float*)measured_derived,
                                                                      sg, cfg->enabled, cfg->seed, cfg->id_offset, tick);
    return launch_check(ctx, "sensor_update_kernel");
}

extern "C" int acb200_sensor_reset(acb200_ctx* ctx, int64_t n, int32_t precision, void* bias, const uint8_t* mask) {
    if (!ctx) return ACB200_ENULL;
    if (!bias) {
        ctx->error = "sensor_reset: NULL bias";
        return ACB200_ENULL;
    }
    if (n < 0 || (precision != 64 && precision != 32)) {
        ctx->error = "sensor_reset: invalid size or precision";
        return ACB200_EVALUE;
    }
    if (n == 0) return ACB200_OK;
    cudaSetDevice(ctx->device);
    const unsigned grid = (unsigned)((n + kBlock - 1) / kBlock);
    if (precision == 64)
        sensor_reset_kernel<double><<<grid, kBlock, 0, ctx->stream>>>(n, (double*)bias, mask);
    else
        sensor_reset_kernel<float><<<grid, kBlock, 0, ctx->stream>>>(n, (float*)bias, mask);
    return launch_check(ctx, "sensor_reset_kernel");
}
