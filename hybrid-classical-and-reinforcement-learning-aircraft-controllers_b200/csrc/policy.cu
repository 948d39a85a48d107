// policy.cu -- learned rate policy, actor forward as ONE fused tcgen05 kernel per call.
//
// Reference path: model.predict(obs, state, deterministic=True) (controllers/learned_rate_agent.py:220-227) through
// LSTMFeaturesExtractor (learned_controllers/networks/lstm_policy.py:13-97), sb3-contrib's stateful actor LSTM, the
// pi-MLP of LSTMPolicy.get_policy_kwargs (lstm_policy.py:107-136) and the action head; SimpleMLPPolicy
// (lstm_policy.py:145-164) runs through the same kernel with a shorter layer table.
//
// One CTA owns 128 aircraft for the whole chain.  Activations never leave the SM: each layer's A operand is a bf16
// buffer in shared memory (X / Y ping-pong), its accumulator is one of two 256-column tensor-memory buffers, and the
// epilogue threads (one per aircraft row) turn accumulator columns into the next layer's operand -- bias, ReLU or
// LSTM gates in fp32 -- while the MMA thread already runs the next chunk into the other buffer.  Weights stream from
// L2 as prepacked stages through a three-slot bulk-copy ring; every bias sits in shared memory for the whole chain.  The recurrent state h travels as a ready-made operand
// tile (one bulk load, one bulk store per tile); c is fp32, read and written by the row's own thread.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "ctx.h"
#include "policy.cuh"

using namespace acbp;

namespace {

// barrier slots inside the 128-byte barrier block
enum { B_FULL = 0, B_EMPTY = 4, B_ACC_FULL = 8, B_ACC_EMPTY = 10, B_ACT = 12, B_HFULL = 13, B_BIAS = 14, B_HGO = 15, B_YFREE = 16, B_COUNT = 17 };
// B_HGO: h' may be stored (peer block of a pair); B_YFREE: the bulk store of h' has finished reading buffer Y

// CTA == 1: one thread block per 128-aircraft tile.  CTA == 2: a cluster of two blocks (an SM pair) works as one
// 256-row tile: the leader's MMA warp issues cta_group::2 instructions for both, each block stages only HALF of every
// weight stage (a 16 KB ring slot then carries twice the K depth: half the L2 traffic and half the ring round trips
// per FLOP), epilogues and recurrent state stay per block.
template <int CTA, int kStages>
__device__ __forceinline__ void policy_forward_body(const PolicyParams& P) {
    constexpr uint32_t kBarOff = bar_off(kStages);
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* act = smem;
    const float* sbias = reinterpret_cast<const float*>(smem + kBiasOff);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t bar0 = smem_base + kBarOff;
    auto bar = [&](int i) { return bar0 + 8u * (uint32_t)i; };
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kBarOff + 8 * B_COUNT);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler
    constexpr bool kPair = CTA == 2;
    const uint32_t rank = kPair ? cluster_ctarank() : 0u;      // 0: leader of the pair
    // Persistent: block (pair) b walks the tile groups b, b + G, b + 2G, ...; barriers, tensor memory and the biases are
    // set up once, the weight ring and the accumulator ring keep turning across tiles (the producer is already loading
    // the next tile's first stages while this tile's last layers drain).  Every role below runs the same sequence.
    const int64_t n_groups = ((P.n + kTile - 1) / kTile + CTA - 1) / CTA;     // tiles (single) or tile pairs
    const int64_t group0 = blockIdx.x / CTA, group_stride = gridDim.x / CTA;
    auto tile_of = [&](int64_t g) { return g * CTA + (int64_t)rank; };
    // Barriers the leader's MMA warp waits on collect arrivals from both blocks.  The leader's own warps arrive locally;
    // the peer's arrive remotely and RELAXED: what the arrival announces is either "my tensor-memory reads are done"
    // (tcgen05.wait::ld + tcgen05.fence) or "my rows of the next operand are written" -- rows that only the peer's own
    // tensor core will read, from the peer's own shared memory, and that every lane has already pushed through
    // fence.proxy.async before the warp's __syncwarp.  Nothing the leader's threads read depends on it, so the
    // cluster-scope release (a ~1000-cycle fence per layer hand-off, measured) buys nothing here.
    auto lead = [&](int i) { return kPair ? mapa(bar(i), 0) : bar(i); };
    auto arrive_lead = [&](int i) { if (kPair && rank != 0) mbar_arrive_cluster_relaxed(lead(i)); else mbar_arrive(bar(i)); };
    // "the operand of layer k is complete" (one lane per epilogue warp).  The peer block's MMA warp is not in the issue
    // chain, so it gets a single-use barrier for the one moment it cares about (h' ready to store; no phase it could miss).
    auto signal_operand = [&](int k) {
        arrive_lead(B_ACT);
        if (kPair && P.recurrent_layer >= 0 && k == P.recurrent_layer + 1) mbar_arrive(bar(B_HGO));
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(bar(B_FULL + s), 1); mbar_init(bar(B_EMPTY + s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar(B_ACC_FULL + a), 1); mbar_init(bar(B_ACC_EMPTY + a), kEpiWarps * CTA); }   // one arrival per epilogue warp
        mbar_init(bar(B_ACT), kEpiWarps * CTA);
        mbar_init(bar(B_HGO), kEpiWarps);
        mbar_init(bar(B_YFREE), 1);
        mbar_init(bar(B_HFULL), 1);
        mbar_init(bar(B_BIAS), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kEpiWarps + 1) tmem_alloc<CTA>(smem_u32(tmem_slot), kTmemCols);
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync();          // the leader's barriers exist before the peer arrives on them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == kEpiWarps) {
        // ================= producer: biases once, then the packed weight stages in consumption order =================
        if (lane == 0) {
            mbar_expect_tx(bar(B_BIAS), (uint32_t)P.n_bias * 4u);
            bulk_g2s(smem_base + kBiasOff, P.bias, (uint32_t)P.n_bias * 4u, bar(B_BIAS));
            uint32_t it = 0;
            for (int64_t g = group0; g < n_groups; g += group_stride) {
            const uint8_t* src = P.weights;
            for (int l = 0; l < P.n_layers; ++l) {
                const LayerDesc& L = P.L[l];
                const uint32_t bytes = (uint32_t)(L.N_c / CTA) * (uint32_t)L.K_s * 2u;   // this block's share of a stage
                const int n_stage = L.n_chunks * (L.K / L.K_s);
                for (int i = 0; i < n_stage; ++i, ++it) {
                    const uint32_t s = it % kStages, use = it / kStages;
                    if (P.debug & 8) continue;          // experiment: no ring at all (the MMA warp does not wait either)
                    if (use > 0) mbar_wait(bar(B_EMPTY + s), (use - 1) & 1, P.error_flag, 100 + l);
                    if (P.debug & 1) { if (rank == 0) mbar_arrive(bar(B_FULL + s)); continue; }
                    if (kPair) {
                        // both halves complete on the LEADER's barrier (tensor-map copy, .cta_group::2): it expects both
                        if (rank == 0) mbar_expect_tx(bar(B_FULL + s), 2u * bytes);
                        tma_load_2d_pair(smem_base + kStageOff + s * kStageBytes, &P.wmap[l], 0, (2 * i + (int)rank) * L.box_rows,
                                         lead(B_FULL + s));
                    } else {
                        mbar_expect_tx(bar(B_FULL + s), bytes);
                        bulk_g2s(smem_base + kStageOff + s * kStageBytes, src, bytes, bar(B_FULL + s));
                        src += bytes;
                    }
                }
            }
            }
        }
    } else if (warp == kEpiWarps + 1) {
        // ================= MMA issuer: the whole warp walks the chain (so that addresses and descriptors stay in
        // uniform registers); one elected lane issues the tensor-core instructions =================
        const bool leader = lane == 0;
        uint32_t slot = 0, slot_phase = 0, chunk_it = 0;
        const bool tr = (P.debug & 4) && blockIdx.x == 0 && leader;
        int ti = 0;
        if (tr) P.trace[ti++] = clock64();
        // h' -> HBM: one bulk store out of buffer Y; once its reads of Y are done the next tile's first layer may
        // overwrite Y (B_YFREE), the write itself only has to land before the kernel ends
        auto store_h = [&](int64_t tile) {
            if (leader) {
                bulk_s2g(P.h_state + tile * (int64_t)kHTileBytes, smem_base + kXBytes, kHTileBytes);
                bulk_commit();
            }
        };
        auto release_y = [&]() {       // not in the issue path: called once the tile's last MMAs are on their way
            if (leader) {
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                mbar_arrive(bar(B_YFREE));
            }
            __syncwarp();
        };
        uint32_t tile_it = 0;
        for (int64_t g = group0; g < n_groups; g += group_stride, ++tile_it) {
        const int64_t tile = tile_of(g);
        if (rank != 0) {
            // peer block of a pair: nothing to issue, only its own tile's recurrent state to move
            if (P.recurrent_layer >= 0) {
                mbar_wait(bar(B_HGO), tile_it & 1, P.error_flag, 251);
                store_h(tile);
                release_y();
            }
            continue;
        }
        const uint32_t act_base = tile_it * (uint32_t)(P.n_layers + 1);     // B_ACT completes n_layers + 1 phases per tile
        for (int l = 0; l < P.n_layers; ++l) {
            const LayerDesc& L = P.L[l];
            mbar_wait<kPair>(bar(B_ACT), (act_base + l) & 1, P.error_flag, 200 + l);   // operand of layer l is complete (in both blocks of a pair)
            if (tr && tile_it == (uint32_t)(P.debug >> 4)) P.trace[ti++] = clock64();
            if (P.recurrent_layer >= 0 && l == P.recurrent_layer + 1) store_h(tile);   // Y holds h' as a finished operand tile
            tc_fence_after();
            const uint32_t idesc = instr_desc_bf16(L.N_c, kTile * CTA);
            const uint32_t lbo_b = (uint32_t)(L.N_c / CTA) * 16u;
            const int stages_per_chunk = L.K / L.K_s, mma_per_stage = L.K_s / 16;
            // descriptors advance in their low word only (addresses stay below 2^18): +2 K slabs per MMA
            const uint64_t a_desc0 = smem_desc(smem_base + L.in_off, kKChunkBytes, 128);
            const uint64_t b_desc0 = smem_desc(smem_base + kStageOff, lbo_b, 128);
            const uint32_t a_step = (2u * kKChunkBytes) >> 4, b_step = (2u * lbo_b) >> 4;
            for (int c = 0; c < L.n_chunks; ++c, ++chunk_it) {
                const uint32_t a = chunk_it & 1, use = chunk_it >> 1;
                if (use > 0) {
                    mbar_wait<kPair>(bar(B_ACC_EMPTY + a), (use - 1) & 1, P.error_flag, 300 + l);
                    tc_fence_after();
                }
                if (tr && tile_it == (uint32_t)(P.debug >> 4)) P.trace[ti++] = clock64();
                const uint32_t d_tmem = tmem_base + a * 256u;
                uint64_t a_desc = a_desc0;
                uint32_t accumulate = 0;
                for (int ks = 0; ks < stages_per_chunk; ++ks) {
                    if (!(P.debug & 8)) mbar_wait<kPair>(bar(B_FULL + slot), slot_phase, P.error_flag, 400 + l);
                    tc_fence_after();
                    if (leader) {
                        uint64_t b_desc = b_desc0 + (uint64_t)(slot * (kStageBytes >> 4));
                        if (mma_per_stage == 2) {
#pragma unroll
                            for (int k = 0; k < 2; ++k) { umma_bf16<CTA>(d_tmem, a_desc + k * a_step, b_desc + k * b_step, idesc, accumulate | k); }
                        } else if (mma_per_stage == 4) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) { umma_bf16<CTA>(d_tmem, a_desc + k * a_step, b_desc + k * b_step, idesc, accumulate | k); }
                        } else {
                            for (int k = 0; k < mma_per_stage; ++k) { umma_bf16<CTA>(d_tmem, a_desc + k * a_step, b_desc + k * b_step, idesc, accumulate | k); }
                        }
                        if (!(P.debug & 8)) umma_commit<CTA>(bar(B_EMPTY + slot));   // stage free once these MMAs have read it
                    }
                    __syncwarp();
                    a_desc += (uint64_t)(mma_per_stage * a_step);
                    accumulate = 1;
                    if (++slot == kStages) { slot = 0; slot_phase ^= 1; }
                }
                if (leader) umma_commit<CTA>(bar(B_ACC_FULL + a));        // accumulator chunk complete
                if (tr && tile_it == (uint32_t)(P.debug >> 4)) P.trace[ti++] = clock64();
            }
        }
        // the last phase of B_ACT of this tile ("last layer's epilogue done") has no MMA behind it: step over it, so that
        // the wait for the next tile's first operand cannot be satisfied by it
        if (P.recurrent_layer >= 0) release_y();
        mbar_wait<kPair>(bar(B_ACT), (act_base + P.n_layers) & 1, P.error_flag, 299);
        }
        if (P.recurrent_layer >= 0 && leader) bulk_wait_all();
    } else {
        // ================= epilogue: warp w reads tensor-memory lanes 32*(w%4).., i.e. aircraft row
        // 32*(w%4)+lane, and takes part w/4 (of kParts) of every chunk's columns =================
        const int part = warp >> 2;
        const int row = (warp & 3) * 32 + lane;
        const uint32_t lane_base = ((uint32_t)((warp & 3) * 32)) << 16;
        uint32_t chunk_it = 0, tile_it = 0;
        const bool tr0 = (P.debug & 4) && blockIdx.x == 0 && threadIdx.x == 0;
        int ti = 1024;
        for (int64_t g = group0; g < n_groups; g += group_stride, ++tile_it) {
        const int64_t tile = tile_of(g);
        const bool tr = tr0 && tile_it == (uint32_t)(P.debug >> 4);
        const int64_t e = tile * kTile + row;
        const bool valid = e < P.n;
        const bool fresh = valid && ((P.episode_start != nullptr && P.episode_start[e] != 0) ||
                                     (P.episode_start_b != nullptr && P.episode_start_b[e] != 0));

        {
            // observation -> X as [hi | lo | hi | 0] bf16 columns (64 wide) against [W_hi | W_hi | W_lo | 0]:
            // x*W ~ (hi+lo)*W_hi + hi*W_lo keeps 16 mantissa bits of both (altitude ~100 m enters unnormalised)
            float o[ACB200_POLICY_OBS];
            if (valid) {
                const float2* src = reinterpret_cast<const float2*>(P.obs + e * ACB200_POLICY_OBS);
#pragma unroll
                for (int q = 0; q < ACB200_POLICY_OBS / 2; ++q) { const float2 t = __ldg(src + q); o[2 * q] = t.x; o[2 * q + 1] = t.y; }
            } else {
#pragma unroll
                for (int q = 0; q < ACB200_POLICY_OBS; ++q) o[q] = 0.f;
            }
            float col[64];
#pragma unroll
            for (int q = 0; q < ACB200_POLICY_OBS; ++q) {
                const float hi = __bfloat162float(__float2bfloat16_rn(o[q]));
                col[q] = hi;
                col[ACB200_POLICY_OBS + q] = o[q] - hi;
                col[2 * ACB200_POLICY_OBS + q] = hi;
            }
#pragma unroll
            for (int q = 3 * ACB200_POLICY_OBS; q < 64; ++q) col[q] = 0.f;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                if (g % kParts != part) continue;          // the four 16-column groups are shared out over the row's warps
                float t[16];
#pragma unroll
                for (int q = 0; q < 16; ++q) t[q] = col[16 * g + q];
                store_row16(act, P.L[0].in_off, 16 * g, row, t);
            }
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) signal_operand(0);
        if (tile_it == 0) mbar_wait(bar(B_BIAS), 0, P.error_flag, 700);
        // the first layer writes buffer Y: the previous tile's h' must have left it
        if (tile_it > 0 && P.recurrent_layer >= 0) mbar_wait(bar(B_YFREE), (tile_it - 1) & 1, P.error_flag, 701);

        if (tr) P.trace[ti++] = clock64();
        for (int l = 0; l < P.n_layers; ++l) {
            const LayerDesc& L = P.L[l];
            for (int c = 0; c < L.n_chunks; ++c, ++chunk_it) {
                const uint32_t a = chunk_it & 1;
                const uint32_t t0 = tmem_base + lane_base + a * 256u;
                if (tr) P.trace[ti++] = clock64();
                const float* bias = sbias + L.bias_off + c * L.N_c;
                if (P.debug & 2) {
                    mbar_wait(bar(B_ACC_FULL + a), (chunk_it >> 1) & 1, P.error_flag, 500 + l);
                    __syncwarp();
                    tc_fence_after();
                } else if (L.epi == EPI_RELU) {
                    const int w = L.N_c / kParts, j0 = part * w;      // w < 16 (the head's layer before it): part 0 takes all
                    mbar_wait(bar(B_ACC_FULL + a), (chunk_it >> 1) & 1, P.error_flag, 500 + l);
                    __syncwarp();
                    tc_fence_after();
                    for (int j = j0; j < j0 + w; j += 16) {
                        uint32_t v[16]; float b[16], h[16];
                        tmem_ld16_issue(t0 + j, v);
                        load_bias16(bias + j, b);
                        tmem_ld_wait(v);
#pragma unroll
                        for (int q = 0; q < 16; ++q) h[q] = max_nan(__uint_as_float(v[q]) + b[q], 0.f);
                        store_row16(act, L.out_off, c * L.N_c + j, row, h);
                    }
                } else if (L.epi == EPI_LSTM0) {
                    // h0 = c0 = 0 on every call (lstm_policy.py:86): c = sig(i) tanh(g), h = sig(o) tanh(c); f unused
                    // chunk = [i | g | o] of U units (N_c = 3U) or [i | (f: zero, unused) | g | o] (N_c = 4U = 256, full-width MMAs)
                    const bool four = L.N_c == 256;
                    const int U = four ? 64 : L.N_c / 3, w = U / kParts, j0 = part * w;
                    const int og = four ? 2 * U : U, oo = four ? 3 * U : 2 * U;
                    mbar_wait(bar(B_ACC_FULL + a), (chunk_it >> 1) & 1, P.error_flag, 500 + l);
                    __syncwarp();
                    tc_fence_after();
                    if (l == P.recurrent_layer - 2 && c == L.n_chunks - 1 && threadIdx.x == 0) {
                        // every MMA of this layer has completed: nothing reads X's h region until the actor LSTM, so
                        // h_prev can travel now, two layers ahead of its use (finish_policy checks the layer table)
                        mbar_expect_tx(bar(B_HFULL), kHTileBytes);
                        bulk_g2s(smem_base + kHOff, P.h_state + tile * (int64_t)kHTileBytes, kHTileBytes, bar(B_HFULL));
                    }
                    for (int j = j0; j < j0 + w; j += 16) {
                        uint32_t vi[16], vg[16], vo[16]; float bi[16], bg[16], bo[16], h[16];
                        tmem_ld16_issue(t0 + j, vi);
                        tmem_ld16_issue(t0 + og + j, vg);
                        tmem_ld16_issue(t0 + oo + j, vo);
                        tmem_ld_wait(vi); tmem_ld_wait(vg); tmem_ld_wait(vo);
                        load_bias16(bias + j, bi); load_bias16(bias + og + j, bg); load_bias16(bias + oo + j, bo);
#pragma unroll
                        for (int q = 0; q < 16; ++q) {
                            const float cc = sigmoid_half(__uint_as_float(vi[q]), bi[q]) * tanh_fast(__uint_as_float(vg[q]) + bg[q]);
                            h[q] = sigmoid_half(__uint_as_float(vo[q]), bo[q]) * tanh_unit(cc);     // |cc| < 1
                        }
                        store_row16(act, L.out_off, c * U + j, row, h);
                    }
                } else if (L.epi == EPI_LSTM) {
                    constexpr int kW = 64 / kParts;                  // U == 64 (finish_policy checks): kW / 16 groups of 16 units
                    const int U = L.N_c / 4, j0 = part * kW;
                    float* cst = P.c_state + (tile * kHidden + (int64_t)c * U + j0) * kTile + row;
                    float cp[kW];
                    const bool carry = valid && !fresh;
#pragma unroll
                    for (int q = 0; q < kW; ++q) cp[q] = carry ? cst[(int64_t)q * kTile] : 0.f;   // before the wait
                    mbar_wait(bar(B_ACC_FULL + a), (chunk_it >> 1) & 1, P.error_flag, 500 + l);
                    __syncwarp();
                    tc_fence_after();
#pragma unroll
                    for (int g = 0; g < kW / 16; ++g) {
                        const int j = j0 + g * 16;
                        uint32_t vi[16], vf[16], vg[16], vo[16]; float bi[16], bf[16], bg[16], bo[16], h[16];
                        tmem_ld16_issue(t0 + j, vi);
                        tmem_ld16_issue(t0 + U + j, vf);
                        tmem_ld16_issue(t0 + 2 * U + j, vg);
                        tmem_ld16_issue(t0 + 3 * U + j, vo);
                        tmem_ld_wait(vi); tmem_ld_wait(vf); tmem_ld_wait(vg); tmem_ld_wait(vo);
                        load_bias16(bias + j, bi); load_bias16(bias + U + j, bf);
                        load_bias16(bias + 2 * U + j, bg); load_bias16(bias + 3 * U + j, bo);
#pragma unroll
                        for (int q = 0; q < 16; ++q) {
                            const float cn = fmaf(sigmoid_half(__uint_as_float(vf[q]), bf[q]), cp[g * 16 + q],
                                                  sigmoid_half(__uint_as_float(vi[q]), bi[q]) * tanh_fast(__uint_as_float(vg[q]) + bg[q]));
                            h[q] = sigmoid_half(__uint_as_float(vo[q]), bo[q]) * tanh_fast(cn);
                            if (valid) cst[(int64_t)(g * 16 + q) * kTile] = cn;
                        }
                        store_row16(act, L.out_off, c * U + j, row, h);
                    }
                } else {   // EPI_HEAD: mean action, clipped to the action space (rate_env.py:128-138)
                    mbar_wait(bar(B_ACC_FULL + a), (chunk_it >> 1) & 1, P.error_flag, 500 + l);
                    __syncwarp();
                    tc_fence_after();
                    if (part == 0) {
                        uint32_t v[16];
                        tmem_ld16_issue(t0, v);
                        tmem_ld_wait(v);
                        if (valid) {
                            float4 o;
                            o.x = min_nan(max_nan(__uint_as_float(v[0]) + bias[0], -1.f), 1.f);
                            o.y = min_nan(max_nan(__uint_as_float(v[1]) + bias[1], -1.f), 1.f);
                            o.z = min_nan(max_nan(__uint_as_float(v[2]) + bias[2], -1.f), 1.f);
                            o.w = min_nan(max_nan(__uint_as_float(v[3]) + bias[3], 0.f), 1.f);
                            reinterpret_cast<float4*>(P.action)[e] = o;
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) arrive_lead(B_ACC_EMPTY + a);
                if (tr) P.trace[ti++] = clock64();
            }
            if (l == P.recurrent_layer - 1) {
                // h_prev has landed behind the features; a fresh episode starts from zeros ((1 - episode_start) * h)
                mbar_wait(bar(B_HFULL), tile_it & 1, P.error_flag, 600);
                if (fresh || !valid) {
                    const uint4 z = make_uint4(0, 0, 0, 0);
                    for (int kc = part * (kHidden / 8 / kParts); kc < (part + 1) * (kHidden / 8 / kParts); ++kc)
                        *reinterpret_cast<uint4*>(act + kHOff + (uint32_t)kc * kKChunkBytes + (uint32_t)row * 16) = z;
                }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) signal_operand(l + 1);
        }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync();          // nobody leaves while the other block may still touch its barriers or tensor memory
    if (warp == kEpiWarps + 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc<CTA>(tmem_base, kTmemCols);
    }
}

template <int CTA, int kStages>
__global__ void __launch_bounds__(kThreads, 1) policy_forward_kernel(const __grid_constant__ PolicyParams P) {
    policy_forward_body<CTA, kStages>(P);
}
// The variant that shares its SMs with a block of the env step kernel (acb200_policy_env_rollout): 2 ring slots (204 KB of
// shared memory) and at most 128 registers, so that 320 x 128 + 128 x 128 registers and 204 + 19 KB fit one SM together.
#ifndef ACB_POLICY_LEAN_REGS
#define ACB_POLICY_LEAN_REGS 128
#endif
__global__ void __maxnreg__(ACB_POLICY_LEAN_REGS) policy_forward_lean_kernel(const __grid_constant__ PolicyParams P) {
    policy_forward_body<2, 2>(P);
}

// ---- recurrent state <-> (n, 256) f32 ---------------------------------------------------------------------------
__global__ void state_export_kernel(int64_t n, const __nv_bfloat16* h_state, const float* c_state, float* h_out, float* c_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // over n * 256, unit fastest
    if (i >= n * kHidden) return;
    const int64_t e = i / kHidden; const int u = (int)(i % kHidden);
    const int64_t tile = e / kTile; const int r = (int)(e % kTile);
    if (h_out) h_out[i] = __bfloat162float(h_state[((tile * (kHidden / 8) + u / 8) * kTile + r) * 8 + u % 8]);
    if (c_out) c_out[i] = c_state[(tile * kHidden + u) * kTile + r];
}
__global__ void state_import_kernel(int64_t n, const float* h_in, const float* c_in, __nv_bfloat16* h_state, float* c_state) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * kHidden) return;
    const int64_t e = i / kHidden; const int u = (int)(i % kHidden);
    const int64_t tile = e / kTile; const int r = (int)(e % kTile);
    if (h_in) h_state[((tile * (kHidden / 8) + u / 8) * kTile + r) * 8 + u % 8] = __float2bfloat16_rn(h_in[i]);
    if (c_in) c_state[(tile * kHidden + u) * kTile + r] = c_in[i];
}

}  // namespace

// =====================================================================================================================
// Host side: layer tables, weight packing, the C ABI
// =====================================================================================================================
struct acb200_policy {
    int kind = 0;
    PolicyParams params{};          // layer table + device pointers of the packed weights / biases
    uint8_t* d_weights = nullptr;
    float* d_bias = nullptr;
    int* d_error = nullptr;            // mapped host memory: survives a trapped kernel
    int* h_error = nullptr;
    size_t weight_bytes = 0;
    double flops_per_aircraft = 0.0;   // algorithmic: 2 * out * in over the layers, h0 = 0 products excluded
    int device = 0;
    int resident = 0;                  // blocks (single) / clusters (pair) that fit on the device at once: the persistent grid
    int cta = 2;                       // thread blocks per MMA: 2 = SM pair (cta_group::2, default), 1 = single block
};

namespace {

// The bounded waits of the kernel write a code here before they trap; the flag lives in mapped host memory so that it
// can still be read (and is reported at exit) after the context is gone.
int* g_last_error_flag = nullptr;
void report_trap() {
    if (g_last_error_flag && *g_last_error_flag)
        fprintf(stderr, "acb200 policy_forward_kernel: bounded wait expired, code %d\n", *g_last_error_flag);
}
cudaError_t alloc_error_flag(int** dev, int** host) {
    cudaError_t e = cudaHostAlloc(host, sizeof(int), cudaHostAllocMapped);
    if (e != cudaSuccess) return e;
    **host = 0;
    if ((e = cudaHostGetDevicePointer(dev, *host, 0)) != cudaSuccess) return e;
    static bool registered = false;
    if (!registered) { atexit(report_trap); registered = true; }
    g_last_error_flag = *host;
    return cudaSuccess;
}

int pfail(acb200_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->error = msg;
    return code;
}

uint16_t bf16_bits(float x) {   // round to nearest even, like __float2bfloat16_rn
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

// A layer as the packer sees it: a dense [n_chunks * N_c][K] matrix in EXECUTION order (rows permuted so that one
// chunk holds every gate of its units, columns padded / duplicated), and its bias in the same row order.
struct HostLayer {
    LayerDesc d;
    std::vector<float> w, b;
    double real_flops;
};

HostLayer make_layer(int epi, int K, int K_s, int N_c, int n_chunks, uint32_t in_off, uint32_t out_off) {
    HostLayer h;
    h.d = LayerDesc{epi, K, K_s, N_c, n_chunks, in_off, out_off, 0};
    h.w.assign((size_t)N_c * n_chunks * K, 0.f);
    h.b.assign((size_t)N_c * n_chunks, 0.f);
    h.real_flops = 0.0;
    return h;
}

// Linear(in -> out): rows as they are; obs_split lays the columns out for the split observation operand
HostLayer dense_layer(int epi, const float* W, const float* B, int out, int in, int N_c, int K, int K_s, uint32_t in_off,
                      uint32_t out_off, bool obs_split) {
    HostLayer h = make_layer(epi, K, K_s, N_c, (out + N_c - 1) / N_c, in_off, out_off);
    for (int n = 0; n < out; ++n) {
        for (int k = 0; k < in; ++k) {
            h.w[(size_t)n * K + k] = W[(size_t)n * in + k];
            if (obs_split) {   // operand columns [obs_hi | obs_lo | obs_hi] against [W_hi | W_hi | W_lo]
                uint32_t hb = (uint32_t)bf16_bits(W[(size_t)n * in + k]) << 16;
                float w_hi;
                memcpy(&w_hi, &hb, 4);
                h.w[(size_t)n * K + k] = w_hi;
                h.w[(size_t)n * K + in + k] = w_hi;
                h.w[(size_t)n * K + 2 * in + k] = W[(size_t)n * in + k] - w_hi;
            }
        }
        h.b[n] = B[n];
    }
    h.real_flops = 2.0 * out * in;
    return h;
}

// nn.LSTM cell, gate rows i,f,g,o of `hidden` each.  Chunk c holds units [c*U, (c+1)*U) of every gate it needs:
// stateless (h0 = c0 = 0): i,g,o (3 gates, W_hh and the f gate drop out); stateful: i,f,g,o over [W_ih | W_hh].
HostLayer lstm_layer(bool stateful, const float* W_ih, const float* W_hh, const float* b_ih, const float* b_hh, int in, int U,
                     uint32_t in_off, uint32_t out_off) {
    // ACB200_POLICY_WIDE=1 (experiment, off by default): the stateless layers issue full-width N = 256 instructions, the unused
    // forget gate carried as a fourth block of ZERO rows (33 % more multiply-adds on those two layers).  The idea: a K = 16 step
    // of an SM pair costs 126 cycles at N = 256 against 148-162 at N = 192.  Measured (r02, 262 144 aircraft): 0.5349 ms against
    // 0.5316 ms -- slower; those chunks are not bound by the instruction shape.  Results are identical either way.
    static const bool wide = getenv("ACB200_POLICY_WIDE") && getenv("ACB200_POLICY_WIDE")[0] == '1';
    const int gates = (stateful || wide) ? 4 : 3, K = stateful ? in + kHidden : in;
    static const int kStateless[3] = {0, 2, 3};
    HostLayer h = make_layer(stateful ? EPI_LSTM : EPI_LSTM0, K, 32, gates * U, kHidden / U, in_off, out_off);
    for (int c = 0; c < kHidden / U; ++c)
        for (int g = 0; g < gates; ++g)
            for (int u = 0; u < U; ++u) {
                if (!stateful && gates == 4 && g == 1) continue;                       // the unused forget gate: zero rows, zero bias
                const int src = ((stateful || gates == 4) ? g : kStateless[g]) * kHidden + c * U + u;
                const size_t dst = (size_t)(c * gates + g) * U + u;
                for (int k = 0; k < in; ++k) h.w[dst * K + k] = W_ih[(size_t)src * in + k];
                if (stateful)
                    for (int k = 0; k < kHidden; ++k) h.w[dst * K + in + k] = W_hh[(size_t)src * kHidden + k];
                const bool sigmoid_gate = ((stateful || gates == 4) ? g : kStateless[g]) != 2;        // i, f, o; g is the tanh gate
                h.b[dst] = (sigmoid_gate ? 0.5f : 1.0f) * (b_ih[src] + b_hh[src]);   // see sigmoid_half()
            }
    h.real_flops = 2.0 * (stateful ? 4 : 3) * kHidden * K;          // algorithmic: the zero forget-gate block is not counted
    return h;
}

int finish_policy(acb200_ctx* ctx, acb200_policy* p, std::vector<HostLayer>& layers, acb200_policy** out);

int build_policy(acb200_ctx* ctx, int kind, std::vector<HostLayer>& layers, acb200_policy** out) {
    acb200_policy* p = new acb200_policy();
    p->kind = kind;
    // Default: the SM-pair kernel (cta_group::2 MMAs, tensor-map weight loads onto the leader's barrier, half the weight
    // traffic per SM): 0.513 ms against 0.561 ms per 262 144-aircraft forward.  ACB200_POLICY_CTA=1 selects
    // the single-block kernel (same arithmetic, bit-identical results; tests/test_gpu_policy.py compares the two).
    if (const char* v = getenv("ACB200_POLICY_CTA")) p->cta = atoi(v) == 1 ? 1 : 2;
    const int rc = finish_policy(ctx, p, layers, out);
    if (rc != ACB200_OK) acb200_policy_destroy(p);
    return rc;
}

int finish_policy(acb200_ctx* ctx, acb200_policy* p, std::vector<HostLayer>& layers, acb200_policy** out) {
    std::vector<uint16_t> packed;
    std::vector<float> bias;
    std::vector<size_t> layer_offset;
    p->params.n_layers = (int)layers.size();
    p->params.recurrent_layer = -1;
    for (size_t l = 0; l < layers.size(); ++l) {
        HostLayer& h = layers[l];
        LayerDesc& d = h.d;
        // stage depth: as many K slabs as one ring slot holds of this block's share of the columns (at most 8 MMAs)
        d.K_s = 16;
        for (int ks = 16; ks <= d.K && ks <= 128; ks += 16)
            if (d.K % ks == 0 && (uint32_t)(d.N_c / p->cta) * ks * 2u <= kStageBytes) d.K_s = ks;
        d.box_rows = (d.N_c / p->cta) * d.K_s * 2 / 128;
        layer_offset.push_back(packed.size() * 2);
        if (d.K % 16 || d.K % d.K_s || d.K_s % 16 || d.N_c % 16 || d.N_c > 256 ||
            (uint32_t)(d.N_c / p->cta) * d.K_s * 2u > kStageBytes || (d.N_c / p->cta) % 8 ||
            ((d.N_c / p->cta) * d.K_s * 2) % 128 || d.box_rows > 256)
            return pfail(ctx, ACB200_EVALUE, "policy: bad layer geometry");
        d.bias_off = (int)bias.size();
        bias.insert(bias.end(), h.b.begin(), h.b.end());
        if (d.epi == EPI_LSTM) p->params.recurrent_layer = (int)l;
        if ((d.epi == EPI_LSTM && d.N_c != 256) || (d.epi == EPI_LSTM0 && (d.N_c != 256 && d.N_c % (48 * kParts))) ||
            (d.epi == EPI_RELU && d.N_c % (16 * kParts)))
            return pfail(ctx, ACB200_EVALUE, "policy: chunk width does not split into 16-column groups over the epilogue warps");
        const int n_half = d.N_c / p->cta;      // a CTA pair splits every stage's columns in two halves, leader first
        for (int c = 0; c < d.n_chunks; ++c)
            for (int ks = 0; ks < d.K / d.K_s; ++ks)
                for (int r = 0; r < p->cta; ++r)
                    for (int kk = 0; kk < d.K_s / 8; ++kk)
                        for (int n = r * n_half; n < (r + 1) * n_half; ++n)
                            for (int j = 0; j < 8; ++j)
                                packed.push_back(bf16_bits(h.w[(size_t)(c * d.N_c + n) * d.K + ks * d.K_s + kk * 8 + j]));
        p->params.L[l] = d;
        p->flops_per_aircraft += h.real_flops;
    }
    if (p->params.recurrent_layer >= 0) {
        // the kernel loads h_prev behind the features (X + kHOff) while layer rec-2's last epilogue and layer rec-1 run
        const int r = p->params.recurrent_layer;
        if (r < 2 || layers[r - 2].d.epi != EPI_LSTM0 || layers[r - 1].d.in_off < kXBytes ||
            layers[r - 1].d.out_off + (uint32_t)layers[r - 1].d.N_c * layers[r - 1].d.n_chunks * 256u > kHOff ||
            layers[r].d.in_off != 0 || layers[r].d.out_off != kXBytes)
            return pfail(ctx, ACB200_EVALUE, "policy: layer table does not leave X's h region free ahead of the recurrent layer");
    }
    while (bias.size() % 4) bias.push_back(0.f);   // bulk copies move multiples of 16 bytes
    if (bias.size() > (size_t)kMaxBias) return pfail(ctx, ACB200_EVALUE, "policy: too many bias values");
    p->params.n_bias = (int)bias.size();
    p->weight_bytes = packed.size() * 2;
    p->device = ctx->device;
    cudaSetDevice(ctx->device);
    cudaError_t e;
    if ((e = cudaMalloc(&p->d_weights, p->weight_bytes)) != cudaSuccess ||
        (e = cudaMalloc(&p->d_bias, bias.size() * sizeof(float))) != cudaSuccess ||
        (e = alloc_error_flag(&p->d_error, &p->h_error)) != cudaSuccess ||
        (e = cudaMemcpy(p->d_weights, packed.data(), p->weight_bytes, cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(p->d_bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(policy_forward_kernel<1, kStagesDefault>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes(kStagesDefault))) != cudaSuccess ||
        (e = cudaFuncSetAttribute(policy_forward_kernel<2, kStagesDefault>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes(kStagesDefault))) != cudaSuccess ||
        (e = cudaFuncSetAttribute(policy_forward_lean_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes(2))) != cudaSuccess) {
        return pfail(ctx, ACB200_ECUDA, std::string("policy upload: ") + cudaGetErrorString(e));
    }
    if (p->cta == 2) {
        // one tiled tensor map per layer: its packed weights seen as rows of 128 bytes, box = one block's half stage
        typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if ((e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q)) != cudaSuccess || !fn)
            return pfail(ctx, ACB200_ECUDA, "policy: cuTensorMapEncodeTiled is not available");
        for (size_t l = 0; l < layers.size(); ++l) {
            const size_t end = l + 1 < layers.size() ? layer_offset[l + 1] : p->weight_bytes;
            const cuuint64_t dims[2] = {64, (end - layer_offset[l]) / 128};
            const cuuint64_t strides[1] = {128};
            const cuuint32_t box[2] = {64, (cuuint32_t)p->params.L[l].box_rows};
            const cuuint32_t estr[2] = {1, 1};
            const CUresult r = ((EncodeTiled)fn)(&p->params.wmap[l], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p->d_weights + layer_offset[l],
                                                 dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return pfail(ctx, ACB200_ECUDA, "policy: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
        }
    }
    p->params.weights = p->d_weights;
    p->params.bias = p->d_bias;
    p->params.error_flag = p->d_error;
    *out = p;
    return ACB200_OK;
}

}  // namespace

extern "C" {

int acb200_policy_create_lstm(acb200_ctx* ctx, const acb200_lstm_policy_weights* w, acb200_policy** out) {
    if (!ctx || !w || !out) return ACB200_ENULL;
    const float* const* ptrs = reinterpret_cast<const float* const*>(w);
    for (size_t i = 0; i < sizeof(*w) / sizeof(float*); ++i)
        if (!ptrs[i]) return pfail(ctx, ACB200_ENULL, "policy: NULL weight pointer");
    const uint32_t X = 0, Y = kXBytes, X2 = kHOff;
    std::vector<HostLayer> L;
    L.push_back(dense_layer(EPI_RELU, w->embed_w, w->embed_b, 128, ACB200_POLICY_OBS, 128, 64, 64, X, Y, true));
    L.push_back(lstm_layer(false, w->l0_w_ih, nullptr, w->l0_b_ih, w->l0_b_hh, 128, 64, Y, X));
    L.push_back(lstm_layer(false, w->l1_w_ih, nullptr, w->l1_b_ih, w->l1_b_hh, 256, 64, X, Y));
    L.push_back(dense_layer(EPI_RELU, w->proj_w, w->proj_b, 128, 256, 128, 256, 64, Y, X, false));
    L.push_back(lstm_layer(true, w->actor_w_ih, w->actor_w_hh, w->actor_b_ih, w->actor_b_hh, 128, 64, X, Y));
    L.push_back(dense_layer(EPI_RELU, w->pi0_w, w->pi0_b, 128, 256, 128, 256, 64, Y, X, false));
    L.push_back(dense_layer(EPI_RELU, w->pi1_w, w->pi1_b, 64, 128, 64, 128, 128, X, X2, false));
    L.push_back(dense_layer(EPI_HEAD, w->action_w, w->action_b, ACB200_POLICY_ACT, 64, 16, 64, 64, X2, X, false));
    return build_policy(ctx, ACB200_POLICY_LSTM, L, out);
}

int acb200_policy_create_mlp(acb200_ctx* ctx, const acb200_mlp_policy_weights* w, acb200_policy** out) {
    if (!ctx || !w || !out) return ACB200_ENULL;
    const float* const* ptrs = reinterpret_cast<const float* const*>(w);
    for (size_t i = 0; i < sizeof(*w) / sizeof(float*); ++i)
        if (!ptrs[i]) return pfail(ctx, ACB200_ENULL, "policy: NULL weight pointer");
    const uint32_t X = 0, Y = kXBytes, X2 = kHOff;
    std::vector<HostLayer> L;
    L.push_back(dense_layer(EPI_RELU, w->pi0_w, w->pi0_b, 256, ACB200_POLICY_OBS, 256, 64, 32, X, Y, true));
    L.push_back(dense_layer(EPI_RELU, w->pi1_w, w->pi1_b, 128, 256, 128, 256, 64, Y, X, false));
    L.push_back(dense_layer(EPI_RELU, w->pi2_w, w->pi2_b, 64, 128, 64, 128, 128, X, X2, false));
    L.push_back(dense_layer(EPI_HEAD, w->action_w, w->action_b, ACB200_POLICY_ACT, 64, 16, 64, 64, X2, X, false));
    return build_policy(ctx, ACB200_POLICY_MLP, L, out);
}

void acb200_policy_destroy(acb200_policy* p) {
    if (!p) return;
    cudaSetDevice(p->device);
    cudaFree(p->d_weights);
    cudaFree(p->d_bias);
    if (p->h_error) { if (g_last_error_flag == p->h_error) g_last_error_flag = nullptr; cudaFreeHost(p->h_error); }
    delete p;
}

int acb200_policy_state_bytes(const acb200_policy* p, int64_t n, size_t* h_bytes, size_t* c_bytes) {
    if (!p) return ACB200_ENULL;
    if (n < 0) return ACB200_EVALUE;
    const size_t tiles = (size_t)(((n + kTile - 1) / kTile + 1) & ~(int64_t)1);   // whole SM pairs
    const bool rec = p->params.recurrent_layer >= 0;
    if (h_bytes) *h_bytes = rec ? tiles * kHTileBytes : 0;
    if (c_bytes) *c_bytes = rec ? tiles * (size_t)kHidden * kTile * sizeof(float) : 0;
    return ACB200_OK;
}

}  // extern "C"

// One forward of n aircraft on `stream` with a weight ring of `stages` slots (3: the standalone forward; 2: beside an env block).
static int policy_forward_impl(acb200_ctx* ctx, acb200_policy* p, int64_t n, const float* obs, const uint8_t* episode_start,
                               const uint8_t* episode_start_b, void* h_state, void* c_state, float* action_out, cudaStream_t stream,
                               int stages) {
    if (!ctx || !p || !obs || !action_out) return pfail(ctx, ACB200_ENULL, "policy_forward: NULL argument");
    if (n < 0) return pfail(ctx, ACB200_EVALUE, "policy_forward: n < 0");
    if (n == 0) return ACB200_OK;
    const bool rec = p->params.recurrent_layer >= 0;
    if (rec && (!h_state || !c_state)) return pfail(ctx, ACB200_ENULL, "policy_forward: recurrent policy needs h_state and c_state");
    if (((uintptr_t)obs & 7) || ((uintptr_t)action_out & 15) || (rec && (((uintptr_t)h_state & 15) || ((uintptr_t)c_state & 3))))
        return pfail(ctx, ACB200_EVALUE, "policy_forward: misaligned buffer");
    if (p->cta != 2) stages = kStagesDefault;            // the 2-slot variant exists for the SM-pair kernel only
    const uint32_t smem = smem_bytes(stages);
    cudaSetDevice(ctx->device);
    PolicyParams P = p->params;
    P.obs = obs; P.episode_start = episode_start; P.episode_start_b = episode_start_b;
    P.h_state = static_cast<uint8_t*>(h_state); P.c_state = static_cast<float*>(c_state);
    P.action = action_out; P.n = n;
    static const int debug = getenv("ACB200_POLICY_DEBUG") ? atoi(getenv("ACB200_POLICY_DEBUG")) : 0;
    P.debug = debug;
    static long long* trace = nullptr;
    if ((debug & 4) && !trace) { cudaMalloc(&trace, 2048 * sizeof(long long)); cudaMemset(trace, 0, 2048 * sizeof(long long)); }
    P.trace = trace;
    const int64_t tiles = (n + kTile - 1) / kTile;
    const int64_t groups = (tiles + p->cta - 1) / p->cta;            // an odd last tile gets an all-padding partner
    cudaError_t e;
    if (p->resident == 0) {          // persistent grid: as many blocks / clusters as the device holds at once
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        p->resident = sms > 0 ? sms / p->cta : 1;
        if (p->cta == 2) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(2 * (unsigned)p->resident); q.blockDim = dim3(kThreads); q.dynamicSmemBytes = smem_bytes(kStagesDefault);
            cudaLaunchAttribute a[1];
            a[0].id = cudaLaunchAttributeClusterDimension;
            a[0].val.clusterDim.x = 2; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
            q.attrs = a; q.numAttrs = 1;
            int clusters = 0;
            if (cudaOccupancyMaxActiveClusters(&clusters, policy_forward_kernel<2, kStagesDefault>, &q) == cudaSuccess && clusters > 0)
                p->resident = clusters;
            else
                cudaGetLastError();
        }
    }
    unsigned n_blocks = (unsigned)(groups < p->resident ? groups : p->resident);
    static const int grid_cap = getenv("ACB200_POLICY_GRID") ? atoi(getenv("ACB200_POLICY_GRID")) : 0;   // development aid: fewer
    if (grid_cap > 0 && (unsigned)grid_cap < n_blocks) n_blocks = (unsigned)grid_cap;                   // blocks, same per-tile time?

    if (p->cta == 2) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * n_blocks);
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        e = (stages == 2) ? cudaLaunchKernelEx(&cfg, policy_forward_lean_kernel, P)
                          : cudaLaunchKernelEx(&cfg, policy_forward_kernel<2, kStagesDefault>, P);
    } else {
        policy_forward_kernel<1, kStagesDefault><<<n_blocks, kThreads, smem, stream>>>(P);
        e = cudaGetLastError();
    }
    ctx->launches += 1;
    if (e != cudaSuccess) return pfail(ctx, ACB200_ECUDA, std::string("policy_forward_kernel: ") + cudaGetErrorString(e));
    if ((debug & 4) && getenv("ACB200_POLICY_TRACE")) {   // development aid: dump block 0's time line
        static int dumped = 0;
        if (dumped++ == 3) {
            std::vector<long long> h(2048);
            cudaStreamSynchronize(stream);
            cudaMemcpy(h.data(), trace, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
            FILE* f = fopen(getenv("ACB200_POLICY_TRACE"), "w");
            if (f) { for (size_t i = 0; i < h.size(); ++i) if (h[i]) fprintf(f, "%zu %lld\n", i, h[i]); fclose(f); }
        }
    }
    return ACB200_OK;
}

extern "C" {

int acb200_policy_forward(acb200_ctx* ctx, acb200_policy* p, int64_t n, const float* obs, const uint8_t* episode_start,
                          const uint8_t* episode_start_b, void* h_state, void* c_state, float* action_out) {
    if (!ctx) return ACB200_ENULL;
    static const int ring = getenv("ACB200_POLICY_RING") ? atoi(getenv("ACB200_POLICY_RING")) : kStagesDefault;   // A/B aid: 2 or 3
    return policy_forward_impl(ctx, p, n, obs, episode_start, episode_start_b, h_state, c_state, action_out, ctx->stream, ring == 2 ? 2 : kStagesDefault);
}

// K closed-loop steps (policy forward -> env step with that action) for the whole fleet in ONE call (BASELINE config C4:
// "policy forward fused with env step and reward").  The fleet is cut into `parts` ranges (multiples of 256 aircraft: whole
// SM-pair tiles of the recurrent state), each with its own stream: within a part the steps are ordered, forward -> env ->
// forward ...; between parts nothing is, so part A's env step (FP32 / SFU pipes, 19 KB of shared memory per block) runs on
// the same SMs, at the same time, as part B's policy forward (tensor pipe + MUFU; persistent, 204 KB with the 2-slot ring) --
// the policy kernel leaves the FP32 pipes ~85 % idle, which is where the env step goes.  Everything is enqueued behind the
// context's stream and that stream continues after all parts: the call is asynchronous and can be captured in a CUDA graph.
int acb200_policy_env_rollout(acb200_ctx* ctx, acb200_policy* p, const acb200_env* env, const acb200_env_config* cfg, void* h_state,
                              void* c_state, float* action, int32_t steps, int32_t parts) {
    if (!ctx || !p || !env || !cfg || !action) return pfail(ctx, ACB200_ENULL, "policy_env_rollout: NULL argument");
    if (steps < 0 || parts < 0 || parts > acb200_ctx::kPipe) return pfail(ctx, ACB200_EVALUE, "policy_env_rollout: steps < 0 or parts outside [0, 3]");
    if (cfg->controller != ACB200_CTRL_EXTERNAL && cfg->controller != ACB200_CTRL_RESIDUAL)
        return pfail(ctx, ACB200_EVALUE, "policy_env_rollout: the env must take its action from the policy (controller external or residual)");
    const int64_t n = env->fleet.n;
    if (n == 0 || steps == 0) return ACB200_OK;
    if (parts == 0) parts = n >= 65536 ? 2 : 1;
    const bool rec = p->params.recurrent_layer >= 0;
    int64_t per = ((n + parts - 1) / parts + 255) / 256 * 256;
    cudaSetDevice(ctx->device);
    cudaError_t e = cudaSuccess;
    if (parts == 1) {
        for (int s = 0; s < steps; ++s) {
            int rc = policy_forward_impl(ctx, p, n, env->obs, env->terminated, env->truncated, h_state, c_state, action, ctx->stream, kStagesDefault);
            if (!rc) rc = acb200_env_step(ctx, env, cfg, action);
            if (rc) return rc;
        }
        return ACB200_OK;
    }
    if ((e = acb200_ctx_pipe_streams(ctx)) != cudaSuccess) return pfail(ctx, ACB200_ECUDA, std::string("rollout streams: ") + cudaGetErrorString(e));
    e = cudaEventRecord(ctx->fork, ctx->stream);
    int used = 0;
    for (int k = 0; k < parts && e == cudaSuccess; ++k) {
        const int64_t first = (int64_t)k * per, count = (n - first < per) ? (n - first) : per;
        if (count <= 0) break;
        used = k + 1;
        const cudaStream_t st = ctx->pipe[k];
        if ((e = cudaStreamWaitEvent(st, ctx->fork, 0)) != cudaSuccess) break;
        const int64_t tile = first / kTile;
        void* hk = rec ? (void*)((uint8_t*)h_state + tile * (int64_t)kHTileBytes) : nullptr;
        void* ck = rec ? (void*)((float*)c_state + tile * (int64_t)kHidden * kTile) : nullptr;
        for (int s = 0; s < steps; ++s) {
            int rc = policy_forward_impl(ctx, p, count, env->obs + first * ACB200_POLICY_OBS, env->terminated + first, env->truncated + first, hk,
                                         ck, action + first * ACB200_POLICY_ACT, st, 2);
            if (!rc) rc = acb200_env_step_range_on(ctx, env, cfg, action, first, count, st);
            if (rc) return rc;
        }
        if ((e = cudaEventRecord(ctx->pipe_done[k], st)) != cudaSuccess) break;
    }
    for (int k = 0; k < used && e == cudaSuccess; ++k) e = cudaStreamWaitEvent(ctx->stream, ctx->pipe_done[k], 0);
    return e == cudaSuccess ? ACB200_OK : pfail(ctx, ACB200_ECUDA, std::string("policy_env_rollout: ") + cudaGetErrorString(e));
}

int acb200_policy_state_export(acb200_ctx* ctx, const acb200_policy* p, int64_t n, const void* h_state, const void* c_state,
                               float* h_out, float* c_out) {
    if (!ctx || !p || !h_state || !c_state) return ACB200_ENULL;
    if (p->params.recurrent_layer < 0) return pfail(ctx, ACB200_ENOSUP, "policy has no recurrent state");
    if (n <= 0) return n < 0 ? ACB200_EVALUE : ACB200_OK;
    cudaSetDevice(ctx->device);
    const int64_t total = n * kHidden;
    state_export_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(n, static_cast<const __nv_bfloat16*>(h_state),
                                                                                 static_cast<const float*>(c_state), h_out, c_out);
    ctx->launches += 1;
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? ACB200_OK : pfail(ctx, ACB200_ECUDA, std::string("state_export_kernel: ") + cudaGetErrorString(e));
}

int acb200_policy_state_import(acb200_ctx* ctx, const acb200_policy* p, int64_t n, const float* h_in, const float* c_in,
                               void* h_state, void* c_state) {
    if (!ctx || !p || !h_state || !c_state) return ACB200_ENULL;
    if (p->params.recurrent_layer < 0) return pfail(ctx, ACB200_ENOSUP, "policy has no recurrent state");
    if (n <= 0) return n < 0 ? ACB200_EVALUE : ACB200_OK;
    cudaSetDevice(ctx->device);
    const int64_t total = n * kHidden;
    state_import_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(n, h_in, c_in, static_cast<__nv_bfloat16*>(h_state),
                                                                                 static_cast<float*>(c_state));
    ctx->launches += 1;
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? ACB200_OK : pfail(ctx, ACB200_ECUDA, std::string("state_import_kernel: ") + cudaGetErrorString(e));
}

int acb200_policy_info(const acb200_policy* p, double* flops_per_aircraft, size_t* weight_bytes, int* smem_bytes) {
    if (!p) return ACB200_ENULL;
    if (flops_per_aircraft) *flops_per_aircraft = p->flops_per_aircraft;
    if (weight_bytes) *weight_bytes = p->weight_bytes;
    if (smem_bytes) *smem_bytes = (int)acbp::smem_bytes(kStagesDefault);
    return ACB200_OK;
}

}  // extern "C"
