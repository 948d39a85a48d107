// telemetry.cu -- sampling of a subset of the fleet for the reference's analysis tools (visualization/logger.py:72-147 keeps
// time, position, velocity, attitude, angular rate, airspeed, altitude and the four surfaces per logged aircraft).
// One launch gathers the k sampled aircraft's 21 fields (state 13 | derived 4 | controls 4) into one slot of a
// caller-owned device ring [capacity][21][k]; the ring leaves the device once, when the log is written.
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/acb200.h"
#include "ctx.h"

namespace {

constexpr int kBlock = 128;

template <typename T>
__global__ void telemetry_gather_kernel(int64_t n, int64_t k, const int64_t* __restrict__ ids, const T* __restrict__ state,
                                        const T* __restrict__ derived, const T* __restrict__ controls, T* __restrict__ out) {
    const int64_t j = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (j >= k) return;
    const int64_t i = ids[j];
    const bool ok = i >= 0 && i < n;                       // an id outside the fleet logs NaN, it does not fault
    const T nan = (T)__int_as_float(0x7fc00000);
#pragma unroll
    for (int f = 0; f < ACB200_STATE_FIELDS; ++f) out[(int64_t)f * k + j] = ok ? state[(int64_t)f * n + i] : nan;
#pragma unroll
    for (int f = 0; f < ACB200_DERIVED_FIELDS; ++f)
        out[(int64_t)(ACB200_STATE_FIELDS + f) * k + j] = ok ? derived[(int64_t)f * n + i] : nan;
#pragma unroll
    for (int f = 0; f < ACB200_CTRL_FIELDS; ++f)
        out[(int64_t)(ACB200_STATE_FIELDS + ACB200_DERIVED_FIELDS + f) * k + j] = (ok && controls) ? controls[(int64_t)f * n + i] : nan;
}

}  // namespace

extern "C" int acb200_telemetry_record(acb200_ctx* ctx, const acb200_fleet* f, const int64_t* ids, int64_t k, void* out) {
    if (!ctx) return ACB200_ENULL;
    if (!f || !ids || !out || !f->state || !f->derived) {
        ctx->error = "telemetry_record: NULL argument";
        return ACB200_ENULL;
    }
    if (k < 0 || f->n < 0 || (f->precision != 64 && f->precision != 32)) {
        ctx->error = "telemetry_record: invalid size or precision";
        return ACB200_EVALUE;
    }
    if (k == 0) return ACB200_OK;
    cudaSetDevice(ctx->device);
    const unsigned grid = (unsigned)((k + kBlock - 1) / kBlock);
    if (f->precision == 64)
        telemetry_gather_kernel<double><<<grid, kBlock, 0, ctx->stream>>>(f->n, k, ids, (const double*)f->state, (const double*)f->derived,
                                                                         (const double*)f->controls, (double*)out);
    else
        telemetry_gather_kernel<float><<<grid, kBlock, 0, ctx->stream>>>(f->n, k, ids, (const float*)f->state, (const float*)f->derived,
                                                                        (const float*)f->controls, (float*)out);
    ctx->launches += 1;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        ctx->error = std::string("telemetry_gather_kernel: ") + cudaGetErrorString(e);
        return ACB200_ECUDA;
    }
    return ACB200_OK;
}
