// envtools.cu -- small device-side helpers around the env step that keep host traffic proportional to what happened:
// acb200_env_gather_done compacts the aircraft that finished an episode in the last step (about 0.2 % of the fleet per
// step at 500-step episodes) into an id-ordered list with their terminal and reset observations -- what SB3's VecEnv
// puts into infos[i]["terminal_observation"] -- so that a trainer whose controller runs in
// the kernel never has to pull the (n, 18) observation matrix over PCIe.
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "ctx.h"

namespace {

constexpr int kTile = 1024;   // aircraft per block: one count per tile, 1024 tiles at 2^20 aircraft

// pass 1: finished aircraft per tile
__global__ void __launch_bounds__(256) done_count_kernel(const uint8_t* __restrict__ terminated, const uint8_t* __restrict__ truncated,
                                                         int64_t n, int32_t* __restrict__ tile_counts) {
    __shared__ int warp_sums[8];
    const int64_t base = (int64_t)blockIdx.x * kTile;
    int c = 0;
#pragma unroll
    for (int r = 0; r < kTile / 256; ++r) {
        const int64_t i = base + r * 256 + threadIdx.x;
        if (i < n) c += (terminated[i] | truncated[i]) ? 1 : 0;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) c += __shfl_down_sync(0xffffffffu, c, off);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += warp_sums[w];
        tile_counts[blockIdx.x] = t;
    }
}

// pass 2: exclusive scan of the tile counts by ONE block (tiles <= a few thousand), total -> count_out
__global__ void __launch_bounds__(1024) done_scan_kernel(int32_t* __restrict__ tile_counts, int tiles, int64_t* __restrict__ count_out) {
    __shared__ int64_t carry;
    __shared__ int64_t warp_tot[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int start = 0; start < tiles; start += 1024) {
        const int t = start + threadIdx.x;
        const int64_t v = t < tiles ? tile_counts[t] : 0;
        int64_t x = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int64_t y = __shfl_up_sync(0xffffffffu, x, off);
            if ((threadIdx.x & 31) >= off) x += y;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            int64_t w = warp_tot[threadIdx.x];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int64_t y = __shfl_up_sync(0xffffffffu, w, off);
                if (threadIdx.x >= off) w += y;
            }
            warp_tot[threadIdx.x] = w;
        }
        __syncthreads();
        const int64_t before = carry + (threadIdx.x >= 32 ? warp_tot[(threadIdx.x >> 5) - 1] : 0) + (x - v);
        if (t < tiles) tile_counts[t] = (int32_t)before;      // exclusive prefix (fits: n < 2^31 finished aircraft)
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *count_out = carry;
}

// pass 3: every tile writes its finished aircraft, in id order, behind its prefix
__global__ void __launch_bounds__(256) done_write_kernel(const uint8_t* __restrict__ terminated, const uint8_t* __restrict__ truncated,
                                                         const float* __restrict__ terminal_obs, const float* __restrict__ obs,
                                                         int64_t n, const int32_t* __restrict__ tile_prefix, int64_t capacity,
                                                         int64_t* __restrict__ ids_out, float* __restrict__ terminal_out,
                                                         float* __restrict__ reset_out) {
    __shared__ int warp_sums[8];
    __shared__ int running;
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kTile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int r = 0; r < kTile / 256; ++r) {
        const int64_t i = base + r * 256 + threadIdx.x;
        const bool d = i < n && ((terminated[i] | truncated[i]) != 0);
        const unsigned m = __ballot_sync(0xffffffffu, d);
        if (lane == 0) warp_sums[warp] = __popc(m);
        __syncthreads();
        int before = running;
        for (int w = 0; w < warp; ++w) before += warp_sums[w];
        if (d) {
            const int64_t slot = (int64_t)tile_prefix[blockIdx.x] + before + __popc(m & ((1u << lane) - 1u));
            if (slot < capacity) {
                ids_out[slot] = i;
                if (terminal_out && terminal_obs) {
#pragma unroll
                    for (int k = 0; k < ACB200_OBS_DIM; ++k) terminal_out[slot * ACB200_OBS_DIM + k] = terminal_obs[i * ACB200_OBS_DIM + k];
                }
                if (reset_out) {
#pragma unroll
                    for (int k = 0; k < ACB200_OBS_DIM; ++k) reset_out[slot * ACB200_OBS_DIM + k] = obs[i * ACB200_OBS_DIM + k];
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int w = 0; w < 8; ++w) t += warp_sums[w];
            running += t;
        }
        __syncthreads();
    }
}

}  // namespace

extern "C" int acb200_env_gather_done(acb200_ctx* ctx, const acb200_env* e, int64_t capacity, int64_t* ids_out,
                                      float* terminal_obs_out, float* reset_obs_out, int64_t* count_out) {
    if (!ctx) return ACB200_ENULL;
    auto fail = [&](int code, const char* msg) { ctx->error = msg; return code; };
    if (!e || !e->terminated || !e->truncated || !e->obs) return fail(ACB200_ENULL, "env / flags / obs is NULL");
    if (!ids_out || !count_out) return fail(ACB200_ENULL, "ids_out / count_out is NULL");
    if (capacity < 0) return fail(ACB200_EVALUE, "capacity < 0");
    if (terminal_obs_out && !e->terminal_obs) return fail(ACB200_ENULL, "env.terminal_obs is NULL");
    const int64_t n = e->fleet.n;
    cudaSetDevice(ctx->device);
    if (n <= 0) {
        const cudaError_t z = cudaMemsetAsync(count_out, 0, sizeof(int64_t), ctx->stream);
        return z == cudaSuccess ? ACB200_OK : fail(ACB200_ECUDA, cudaGetErrorString(z));
    }
    const int tiles = (int)((n + kTile - 1) / kTile);
    if (ctx->scratch_i32_len < tiles) {          // the handle's own scratch: 4 bytes per 1024 aircraft
        if (ctx->scratch_i32) cudaFree(ctx->scratch_i32);
        ctx->scratch_i32 = nullptr; ctx->scratch_i32_len = 0;
        const cudaError_t a = cudaMalloc(&ctx->scratch_i32, (size_t)tiles * sizeof(int32_t));
        if (a != cudaSuccess) return fail(ACB200_ECUDA, cudaGetErrorString(a));
        ctx->scratch_i32_len = tiles;
    }
    done_count_kernel<<<tiles, 256, 0, ctx->stream>>>(e->terminated, e->truncated, n, ctx->scratch_i32);
    done_scan_kernel<<<1, 1024, 0, ctx->stream>>>(ctx->scratch_i32, tiles, count_out);
    done_write_kernel<<<tiles, 256, 0, ctx->stream>>>(e->terminated, e->truncated, e->terminal_obs, e->obs, n, ctx->scratch_i32,
                                                      capacity, ids_out, terminal_obs_out, reset_obs_out);
    ctx->launches += 3;
    const cudaError_t err = cudaGetLastError();
    return err == cudaSuccess ? ACB200_OK : fail(ACB200_ECUDA, cudaGetErrorString(err));
}

// =====================================================================================================================
// acb200_env_gather_done_noalloc -- the same compaction for callers that may not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The per-tile counts live in a file-scope device array instead of the handle's cudaMalloc'ed scratch:
//   * size: kNoAllocMaxTiles ints = 64 KiB, held by every GPU that loads the library until the process exits;
//   * limit: fleets of at most ACB200_NOALLOC_MAX_AIRCRAFT = kNoAllocMaxTiles * 1024 = 16 777 216 aircraft; larger n is
//     refused with ACB200_EVALUE (use acb200_env_gather_done);
//   * the array is SHARED by every call on that device and keeps the last call's contents.  Nothing in it has to start at
//     zero: pass 1 overwrites every tile count [0, tiles) the later passes read, so no clearing is needed; *count_out is
//     written by pass 2.  Two calls that overlap on the GPU (different streams, or two contexts on one device) race on
//     it: keep the calls of one device on one stream, or order them with events.
// In host code `g_noalloc_tile_counts` is only a host-side stand-in for the symbol, not a device address; the address
// comes from cudaGetSymbolAddress, once per device, with that device current.
// =====================================================================================================================
#include <mutex>

namespace {

constexpr int kNoAllocMaxTiles = 16384;
static __device__ __align__(256) int32_t g_noalloc_tile_counts[kNoAllocMaxTiles];

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
int32_t* g_noalloc_ptr[kMaxDevices] = {};      // device address of the array on device d, resolved on first use

// Call with `device` current.  Returns cudaSuccess and the device-side address of the scratch array.
cudaError_t noalloc_scratch(int device, int32_t** out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        int32_t* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_tile_counts);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" int64_t acb200_noalloc_max_aircraft(void) { return (int64_t)kNoAllocMaxTiles * kTile; }

extern "C" int acb200_env_gather_done_noalloc(acb200_ctx* ctx, const acb200_env* e, int64_t capacity, int64_t* ids_out,
                                              float* terminal_obs_out, float* reset_obs_out, int64_t* count_out) {
    if (!ctx) return ACB200_ENULL;
    auto fail = [&](int code, const std::string& msg) { ctx->error = msg; return code; };
    if (!e || !e->terminated || !e->truncated || !e->obs) return fail(ACB200_ENULL, "env / flags / obs is NULL");
    if (!ids_out || !count_out) return fail(ACB200_ENULL, "ids_out / count_out is NULL");
    if (capacity < 0) return fail(ACB200_EVALUE, "capacity < 0");
    if (terminal_obs_out && !e->terminal_obs) return fail(ACB200_ENULL, "env.terminal_obs is NULL");
    const int64_t n = e->fleet.n;
    if (n < 0) return fail(ACB200_EVALUE, "fleet.n < 0");
    if (n > (int64_t)kNoAllocMaxTiles * kTile)
        return fail(ACB200_EVALUE, "env_gather_done_noalloc: fleet of " + std::to_string(n) + " aircraft exceeds the static scratch (max " +
                                       std::to_string((int64_t)kNoAllocMaxTiles * kTile) + "); use acb200_env_gather_done");
    cudaError_t err = cudaSetDevice(ctx->device);
    if (err != cudaSuccess) return fail(ACB200_ECUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(err));
    if (n == 0) {
        err = cudaMemsetAsync(count_out, 0, sizeof(int64_t), ctx->stream);
        return err == cudaSuccess ? ACB200_OK : fail(ACB200_ECUDA, cudaGetErrorString(err));
    }
    int32_t* scratch = nullptr;
    if ((err = noalloc_scratch(ctx->device, &scratch)) != cudaSuccess)
        return fail(ACB200_ECUDA, std::string("cudaGetSymbolAddress(g_noalloc_tile_counts): ") + cudaGetErrorString(err));
    const int tiles = (int)((n + kTile - 1) / kTile);
    done_count_kernel<<<tiles, 256, 0, ctx->stream>>>(e->terminated, e->truncated, n, scratch);
    done_scan_kernel<<<1, 1024, 0, ctx->stream>>>(scratch, tiles, count_out);
    done_write_kernel<<<tiles, 256, 0, ctx->stream>>>(e->terminated, e->truncated, e->terminal_obs, e->obs, n, scratch, capacity, ids_out,
                                                      terminal_obs_out, reset_obs_out);
    ctx->launches += 3;
    err = cudaGetLastError();
    return err == cudaSuccess ? ACB200_OK : fail(ACB200_ECUDA, cudaGetErrorString(err));
}
