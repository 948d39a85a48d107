// kernels_f32.cu -- THROUGHPUT build of the kernels: float physics/reward (FMA contraction on,
// short-range polynomial transcendentals) + the same bit-exact float PID.
#define ACB_PAIR_KERNEL 1
#include "kernels.cuh"

namespace acb {
template struct Launch<float>;
}
