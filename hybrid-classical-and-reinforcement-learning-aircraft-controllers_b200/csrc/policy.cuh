// policy.cuh -- device-side building blocks of the fused policy-forward kernel (row a20): raw sm_100a PTX for
// mbarriers, bulk (TMA) copies, tensor-memory management and tcgen05.mma, plus the layer table the kernel walks.
//
// Operand layout (no tensor map needed): every A/B operand sits in shared memory as the UMMA "K-major, no swizzle"
// canonical layout -- 8-row x 16-byte core matrices, each 128 contiguous bytes:
//     element (row, k)  ->  ((k / 8) * ROWS + row) * 16 bytes + (k % 8) * 2 bytes          (bf16)
// so the stride between the two K-halves of one MMA (leading-dimension byte offset) is ROWS*16 and the stride between
// 8-row groups (stride-dimension byte offset) is 128.  Activations are written in this layout by the epilogue
// threads (one thread per aircraft row, one 16-byte store per 8 outputs), weights are packed into it on the host
// once, so a weight stage is ONE contiguous cp.async.bulk from L2.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace acbp {

constexpr int kTile = 128;               // aircraft per CTA == UMMA M == tensor-memory lanes
constexpr int kMaxLayers = 8;
#ifndef ACB_POLICY_STAGES
#define ACB_POLICY_STAGES 3
#endif
constexpr int kStagesDefault = ACB_POLICY_STAGES;   // weight ring depth of the standalone forward
constexpr int kStagesMax = 4;
#ifndef ACB_POLICY_STAGE_BYTES
#define ACB_POLICY_STAGE_BYTES 16384
#endif
constexpr uint32_t kStageBytes = ACB_POLICY_STAGE_BYTES;
constexpr uint32_t kXBytes = 98304;      // activation buffer X: up to K = 384 (features 128 | h 256)
constexpr uint32_t kYBytes = 65536;      // activation buffer Y: up to K = 256
constexpr int kMaxBias = 3584;           // every layer's bias, fp32, resident for the whole chain (3 536 with full-width stateless chunks)
constexpr uint32_t kBiasOff = kXBytes + kYBytes;
constexpr uint32_t kStageOff = kBiasOff + kMaxBias * 4;
// The ring depth is a template parameter of the kernel: 3 slots for the forward on its own; 2 slots (16 KB less shared
// memory: 204 KB) where a block of the env step kernel has to fit on the same SM beside the persistent policy block
// (acb200_policy_env_rollout).  Measured, 262 144 aircraft: 0.526 ms with 3 x 16 KB, 0.556 with 2 x 16 KB, 0.599 with 4 x 8 KB.
__host__ __device__ constexpr uint32_t bar_off(int stages) { return kStageOff + (uint32_t)stages * kStageBytes; }
__host__ __device__ constexpr uint32_t smem_bytes(int stages) { return bar_off(stages) + 256; }
constexpr uint32_t kKChunkBytes = kTile * 16;   // one 8-wide K slab of an activation buffer (2048 B)
constexpr int kHidden = 256;
constexpr uint32_t kHTileBytes = kHidden * kTile * 2;   // one tile of h, bf16, already in operand layout
constexpr uint32_t kHOff = 16 * kKChunkBytes;           // h sits behind the 128 feature columns in X
#ifndef ACB_EPI_WARPS
#define ACB_EPI_WARPS 8
#endif
constexpr int kEpiWarps = ACB_EPI_WARPS;  // kParts warps per tensor-memory lane quarter: each takes 1/kParts of a chunk's columns
constexpr int kParts = kEpiWarps / 4;
static_assert(kEpiWarps == 8 || kEpiWarps == 16, "epilogue warps: 8 or 16");
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = kEpiThreads + 64;   // + weight producer warp, MMA issuer warp
constexpr uint32_t kTmemCols = 512;      // two accumulator buffers of 256 fp32 columns

enum Epilogue : int { EPI_RELU = 0, EPI_LSTM0 = 1, EPI_LSTM = 2, EPI_HEAD = 3 };

struct LayerDesc {
    int epi;
    int K;          // operand depth as executed (multiple of 16)
    int K_s;        // depth of one weight stage
    int N_c;        // accumulator columns per chunk (multiple of 16, <= 256)
    int n_chunks;
    uint32_t in_off, out_off;   // byte offsets of the A operand / of output column 0 inside the activation area
    int bias_off;   // float offset into the bias array (permuted like the packed rows)
    int box_rows;   // CTA-pair mode: 128-byte rows of one block's half of a stage (the tensor map's box height)
};

struct PolicyParams {
    CUtensorMap wmap[kMaxLayers];   // CTA-pair mode: layer l's packed weights as rows of 128 bytes, box = one block's half stage
    int n_layers;
    int recurrent_layer;      // index of the EPI_LSTM layer, -1 if none
    LayerDesc L[kMaxLayers];
    const uint8_t* weights;   // packed stages in consumption order
    const float* bias;        // [n_bias] in execution order; sigmoid gates hold 0.5 * b
    int n_bias;
    const float* obs;         // [n][18]
    const uint8_t* episode_start;     // either flag set: fresh episode
    const uint8_t* episode_start_b;
    uint8_t* h_state;         // [tiles][32][128][8] bf16
    float* c_state;           // [tiles][256][128]
    float* action;            // [n][4]
    int64_t n;
    int* error_flag;
    long long* trace;         // debug & 4: clock64 stamps of block 0 ([0..1023] MMA thread, [1024..2047] epilogue thread 0)
    int debug;                // timing experiments only (ACB200_POLICY_DEBUG): 1 = no weight traffic, 2 = no gate math, 8 = no weight ring at all
};

// ---- shared-memory addresses, mbarriers -------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must end as a trapped kernel with a code in *err, not as a hung GPU.
// CLUSTER: the barrier collects arrivals from both CTAs of a pair (acquire at cluster scope).
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity);
template <bool CLUSTER = false>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err, int code) {
    auto ready = [&]() { return CLUSTER ? mbar_try_wait_cluster(bar, parity) : mbar_try_wait(bar, parity); };
    if (ready()) return;
    long long t0 = 0;
    uint32_t spins = 0;
    while (!ready()) {
        if ((++spins & 0xFFFu) != 0) continue;
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > 4000000000LL) {   // about two seconds
            if (err) *err = code;
            __threadfence_system();
            __trap();
        }
    }
}

// ---- thread-block cluster (CTA pair) ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on a barrier that may live in the other CTA of the pair (address from mapa)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
// same without ordering of this thread's earlier memory operations: enough where the arrival only says "my tensor-memory
// reads are done" (tcgen05.wait::ld + tcgen05.fence cover those); a cluster-scope release costs hundreds of cycles
__device__ __forceinline__ void mbar_arrive_cluster_relaxed(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {   // own barrier, arrivals from the pair
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Tiled (tensor-map) load of one box into THIS block's shared memory, completing on a barrier that may be the pair
// leader's: only the tensor form of the bulk copy takes .cta_group::2, which is what lets both halves of a weight
// stage report to the one barrier the MMA warp waits on.
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t cluster_bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(cluster_bar)
        : "memory");
}

// ---- bulk asynchronous copies (TMA engine, 1-D) ---------------------------------------------------------------------
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory -> visible to the async proxy (UMMA operand reads, bulk stores)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tensor memory + tcgen05 ---------------------------------------------------------------------------------------
template <int CTA>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {   // whole warp (the same warp in both CTAs of a pair)
    if (CTA == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
}
template <int CTA>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {    // whole warp, the allocating one
    if (CTA == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor: K-major, SWIZZLE_NONE, descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor: D = f32, A = B = bf16, both K-major, M = 128 per CTA (256 over a CTA pair)
__device__ __forceinline__ uint32_t instr_desc_bf16(int n_cols, int m_rows) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_cols >> 3) << 17) | ((uint32_t)(m_rows >> 4) << 24);
}
// CTA == 2: one instruction drives both SMs of the pair -- A rows 0..127 from the leader's shared memory, 128..255 from
// the peer's (same offsets), B columns split in halves between the two, D rows into each CTA's own tensor memory.
template <int CTA>
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    if (CTA == 2)
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
            : "memory");
    else
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
            : "memory");
}
// mbarrier arrive once every tcgen05 operation this thread issued so far has completed; for a CTA pair the arrive is
// multicast to the barrier at the same offset in both CTAs
template <int CTA>
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    if (CTA == 2) {
        const uint16_t mask = 3;
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                     "h"(mask)
                     : "memory");
    } else {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    }
}

// 16 consecutive fp32 accumulator columns of this thread's lane (32x32b: lane = 32 * (warp % 4) + laneid)
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&v)[16]) {
#ifdef ACB_DBG_NO_LDTM        // timing experiment only: no tensor-memory read, values the compiler cannot fold
#pragma unroll
    for (int q = 0; q < 16; ++q) asm volatile("mov.b32 %0, %1;" : "=r"(v[q]) : "r"(taddr + q));
    return;
#endif
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
// wait for the loads; the registers are operands so that no use can be scheduled above the wait
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                   "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
                 :
                 : "memory");
}

// ---- gate arithmetic (fp32; one MUFU per transcendental) ---------------------------------------------------------------
__device__ __forceinline__ float tanh_fast(float x) {
#ifdef ACB_DBG_NO_MUFU        // timing experiment only: the transcendental replaced by one FMA
    return fmaf(x, 0.25f, 0.125f);
#else
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#endif
}
// tanh on [-1, 1] on the FMA pipe: x * P(x^2), P of degree 3, max abs error 3.3e-5 (tanh.approx itself: ~5e-4).  Used where
// the argument is a product of a sigmoid and a tanh, so that one of the four transcendentals per stateless LSTM unit
// leaves the MUFU pipe, which is what bounds the epilogue (tools/fit_tanh_unit.py refits the coefficients).
__device__ __forceinline__ float tanh_unit(float x) {
    const float t = x * x;
    float p = fmaf(t, -0.024656429886817932f, 0.11541730910539627f);
    p = fmaf(t, p, -0.3288934528827667f);
    p = fmaf(t, p, 0.9996939301490784f);
    return x * p;
}
// sigmoid(z) with z = 2 * (0.5 * acc + 0.5 * b): the bias array holds 0.5 * b for these gates, so one FFMA feeds the MUFU
__device__ __forceinline__ float sigmoid_half(float acc, float half_b) { return fmaf(0.5f, tanh_fast(fmaf(acc, 0.5f, half_b)), 0.5f); }

// NaN-propagating max / min: torch.relu and np.clip hand a NaN through, fmaxf / fminf would swallow it
__device__ __forceinline__ float max_nan(float a, float b) {
    float y;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(y) : "f"(a), "f"(b));
    return y;
}
__device__ __forceinline__ float min_nan(float a, float b) {
    float y;
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(y) : "f"(a), "f"(b));
    return y;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
// 16 consecutive output columns of one row -> two 16-byte stores in operand layout
__device__ __forceinline__ void store_row16(uint8_t* act, uint32_t out_off, int k0, int row, const float (&h)[16]) {
    uint4 a = make_uint4(pack_bf16(h[0], h[1]), pack_bf16(h[2], h[3]), pack_bf16(h[4], h[5]), pack_bf16(h[6], h[7]));
    uint4 b = make_uint4(pack_bf16(h[8], h[9]), pack_bf16(h[10], h[11]), pack_bf16(h[12], h[13]), pack_bf16(h[14], h[15]));
    uint8_t* p = act + out_off + (uint32_t)(k0 >> 3) * kKChunkBytes + (uint32_t)row * 16;
    *reinterpret_cast<uint4*>(p) = a;
    *reinterpret_cast<uint4*>(p + kKChunkBytes) = b;
}
__device__ __forceinline__ void load_bias16(const float* b, float (&out)[16]) {   // shared memory, warp-uniform address
    const float4* p = reinterpret_cast<const float4*>(b);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float4 t = p[q];
        out[4 * q] = t.x; out[4 * q + 1] = t.y; out[4 * q + 2] = t.z; out[4 * q + 3] = t.w;
    }
}

}  // namespace acbp
