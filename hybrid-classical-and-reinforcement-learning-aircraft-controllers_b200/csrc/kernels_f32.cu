// kernels_f32.cu -- THROUGHPUT build of the kernels: float physics / reward (every fused multiply-add written
// out: fmaf, __ffma2_rn, Math<float>::mad; short-range polynomial transcendentals) + the same bit-exact float PID.
// Compiled with -fmad=false as well (since round 2): the compiler contracts nothing on its own, so two kernels that inline
// the same function in different surroundings (one aircraft per thread / a pair per thread / TMA-staged) round it the same
// way -- their bit-identity tests had started to fail on contraction choices that moved with unrelated edits.
#define ACB_PAIR_KERNEL 1
#include "kernels.cuh"

namespace acb {
template struct Launch<float>;
}
