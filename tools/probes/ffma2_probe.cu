// ffma2_probe.cu -- issue/pipe throughput of Blackwell's packed FP32 instructions (FFMA2) against scalar FFMA, and
// of the mixes the fused substep is made of (FMA-pipe + ALU-pipe FMNMX + MUFU).  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probes/ffma2_probe tools/probes/ffma2_probe.cu
// Prints one JSON object.  Measurement tool only (not on the product path).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define ITERS 4096
template <int MODE> __global__ void __launch_bounds__(256) probe(float* out, float a, float b) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    float2 x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = make_float2(tid * 1e-9f + i * 0.01f, tid * 2e-9f - i * 0.01f);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    for (int it = 0; it < ITERS; ++it) {
        const float2 y2 = make_float2(x[7].x * 0.5f, x[6].y * 0.5f);   // a per-thread register pair that changes every pass (mode 13)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) {          // 16 scalar FFMA
                x[i].x = fmaf(x[i].x, a, b);
                x[i].y = fmaf(x[i].y, a, b);
            } else if (MODE == 1) {   // 8 FFMA2 (same flops)
                x[i] = __ffma2_rn(x[i], a2, b2);
            } else if (MODE == 2) {   // 8 FFMA2 + 16 FMNMX
                x[i] = __ffma2_rn(x[i], a2, b2);
                x[i].x = fminf(x[i].x, 3.0e38f);
                x[i].y = fminf(x[i].y, 3.0e38f);
            } else if (MODE == 3) {   // 16 FFMA + 16 FMNMX
                x[i].x = fminf(fmaf(x[i].x, a, b), 3.0e38f);
                x[i].y = fminf(fmaf(x[i].y, a, b), 3.0e38f);
            } else if (MODE == 4) {   // 8 FFMA2 + 4 MUFU
                x[i] = __ffma2_rn(x[i], a2, b2);
                if (i % 2 == 0) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i].x));
            } else if (MODE == 5) {   // 16 FFMA + 4 MUFU
                x[i].x = fmaf(x[i].x, a, b);
                x[i].y = fmaf(x[i].y, a, b);
                if (i % 2 == 0) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i].x));
            } else if (MODE == 6) {   // 8 FFMA2 + 8 FMNMX + 2 MUFU  (close to the substep's mix per 16 flop-pairs)
                x[i] = __ffma2_rn(x[i], a2, b2);
                x[i].x = fminf(x[i].x, 3.0e38f);
                if (i % 4 == 0) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i].y));
            } else if (MODE == 8) {   // 8 FFMA2, three distinct 64-bit register operands each (register-file bandwidth)
                x[i] = __ffma2_rn(x[(i + 3) & 7], x[(i + 5) & 7], x[i]);
            } else if (MODE == 9) {   // 16 FFMA, three distinct register operands each
                x[i].x = fmaf(x[(i + 3) & 7].x, x[(i + 5) & 7].x, x[i].x);
                x[i].y = fmaf(x[(i + 3) & 7].y, x[(i + 5) & 7].y, x[i].y);
            } else if (MODE == 10) {  // 8 FFMA2 (3 register operands) + 8 FMNMX + 2 MUFU
                x[i] = __ffma2_rn(x[(i + 3) & 7], x[(i + 5) & 7], x[i]);
                x[i].x = fminf(x[i].x, 3.0e38f);
                if (i % 4 == 0) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i].y));
            } else if (MODE == 11) {  // 8 FFMA2, three register pairs, neighbours (another register-bank pattern than mode 8)
                x[i] = __ffma2_rn(x[(i + 1) & 7], x[(i + 2) & 7], x[i]);
            } else if (MODE == 12) {  // 16 FFMA, three registers of mixed parity
                x[i].x = fmaf(x[(i + 3) & 7].x, x[(i + 5) & 7].y, x[i].x);
                x[i].y = fmaf(x[(i + 3) & 7].y, x[(i + 5) & 7].x, x[i].y);
            } else if (MODE == 13) {  // 8 FFMA2, three register pairs, the multiplier shared by the eight (operand reuse cache)
                x[i] = __ffma2_rn(x[(i + 3) & 7], y2, x[i]);
            } else if (MODE == 14) {  // 8 FFMA2, two register pairs + the accumulator itself (x = x * y + x')
                x[i] = __ffma2_rn(x[i], x[(i + 3) & 7], x[i]);
            } else if (MODE == 7) {   // the same mix, scalar
                x[i].x = fminf(fmaf(x[i].x, a, b), 3.0e38f);
                x[i].y = fmaf(x[i].y, a, b);
                if (i % 4 == 0) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i].y));
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i].x + x[i].y;
    if (s == 12345.678f) out[tid] = s;
}

template <int MODE> static double run(int sms, float* d, int blocks_per_sm = 8, int threads = 256) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sms * blocks_per_sm;
    probe<MODE><<<blocks, threads>>>(d, 0.999f, 1e-3f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) probe<MODE><<<blocks, threads>>>(d, 0.999f, 1e-3f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    // fma lanes-ops per second (one FFMA lane-op = 2 flop)
    return 5.0 * blocks * (double)threads * ITERS * 16.0 / (ms * 1e-3);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    float* d; cudaMalloc(&d, 1 << 26);
    const int sms = p.multiProcessorCount;
    double r[8];
    r[0] = run<0>(sms, d); r[1] = run<1>(sms, d); r[2] = run<2>(sms, d); r[3] = run<3>(sms, d);
    r[4] = run<4>(sms, d); r[5] = run<5>(sms, d); r[6] = run<6>(sms, d); r[7] = run<7>(sms, d);
    // three-register-operand forms, and the substep-like mix at the pair kernel's occupancy (12 warps per SM)
    const double r8 = run<8>(sms, d), r9 = run<9>(sms, d), r10 = run<10>(sms, d);
    const double m12 = run<6>(sms, d, 3, 128), m12r = run<10>(sms, d, 3, 128), m16 = run<6>(sms, d, 4, 128), m24s = run<7>(sms, d, 6, 128);
    printf("{\"register_patterns\": {\"ffma2_neighbours\": %.4g, \"ffma_mixed_parity\": %.4g, \"ffma2_shared_multiplier\": %.4g, "
           "\"ffma2_two_pairs\": %.4g}}\n", run<11>(sms, d), run<12>(sms, d), run<13>(sms, d), run<14>(sms, d));
    printf("{\"three_reg\": {\"ffma2\": %.4g, \"ffma\": %.4g, \"mix_ffma2\": %.4g}, \"occupancy\": {\"mix_ffma2_12warps\": %.4g, "
           "\"mix_ffma2_3reg_12warps\": %.4g, \"mix_ffma2_16warps\": %.4g, \"mix_ffma_24warps\": %.4g}}\n", r8, r9, r10, m12, m12r, m16, m24s);
    printf("{\"sms\": %d, \"fma_lane_ops_per_s\": {\"ffma\": %.4g, \"ffma2\": %.4g, \"ffma2+fmnmx\": %.4g, \"ffma+fmnmx\": %.4g, "
           "\"ffma2+mufu\": %.4g, \"ffma+mufu\": %.4g, \"mix_ffma2\": %.4g, \"mix_ffma\": %.4g}, "
           "\"tflops_ffma\": %.2f, \"tflops_ffma2\": %.2f}\n",
           sms, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], 2 * r[0] / 1e12, 2 * r[1] / 1e12);
    return 0;
}
