/* CPU sweep behind csrc/pid.cuh's PidDivFused: for 1 500 divisors d in [2^-20, 2^21) (random mantissas + the extreme ones)
 * and ALL 2^23 numerators of one binade, q = fma(x - (x r) d, r, x r) with r = (float)(1 / (double)d) equals x / d bit for bit.
 *   gcc -O2 -fopenmp -mfma -ffp-contract=off tools/probes/pid_div_check.c -o /tmp/pid_div_check -lm && /tmp/pid_div_check
 * prints bad=0 (8 cores: 1.2 s).  The window in the exponent of x (2^-100 .. 2^100) was found with the same loop over every
 * positive float for 18 divisors (0.01, 0.02, 0.001, 1/60 ...): no mismatch inside it, under-/overflow of the remainder
 * outside.  With r off by one ulp the same sweep finds 1 096 mismatches in 1.3e10 quotients: it does detect. */
#include <stdio.h>
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
static inline float asf(uint32_t u){float f; memcpy(&f,&u,4); return f;}
static inline uint32_t asu(float f){uint32_t u; memcpy(&u,&f,4); return u;}
int main(){
  srand(12345);
  unsigned long bad=0; int nd=0;
  #pragma omp parallel for reduction(+:bad)
  for (int t=0;t<1500;++t){
    uint32_t m;
    if (t<8) { uint32_t sp[8]={0x7fffff,0x7ffffe,0x000001,0x400000,0x3fffff,0x555555,0x2aaaaa,0x000000}; m=sp[t]; }
    else { m = ((uint32_t)rand_r((unsigned*)&(unsigned){t*7919u+1}) ^ ((uint32_t)t*2654435761u)) & 0x7fffff; }
    int e = 107 + (t % 41);
    float d=asf(((uint32_t)e<<23)|m);
    float r=(float)(1.0/(double)d);
    for (uint32_t u=0x3f800000u; u<0x40000000u; ++u){
      float x=asf(u);
      float q0=x*r; float rem=fmaf(-q0,d,x); float q=fmaf(rem,r,q0);
      if (asu(q)!=asu(x/d)) bad++;
    }
  }
  printf("bad=%lu\n", bad);
}
