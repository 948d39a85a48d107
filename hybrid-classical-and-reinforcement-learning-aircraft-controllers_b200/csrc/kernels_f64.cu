// kernels_f64.cu -- PARITY build of the kernels: double physics/reward + float PID.
// Compiled with -fmad=false so that no a*b+c is contracted: the reference's numpy/C++ arithmetic has
// no FMA and 1e-9 parity over 10k steps has less than a decade of margin (F10).
#include "kernels.cuh"

namespace acb {
template struct Launch<double>;
}
