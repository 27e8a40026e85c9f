// Exhaustive host check of the constant-divisor sequence K2 uses for levels == 255 (kernels.cuh:
// divideByLevels): every float with 1e-20 < |x| < 1e20, and +0, against x / 255.0f.
//   gcc -O2 -ffp-contract=off -fopenmp -o /tmp/div255 scripts/div255_exhaustive.c -lm && /tmp/div255
// Result (this container, 2026-11): unsafe 2048841785 bad(one step) 0 bad(two steps) 0
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <omp.h>
int main(){
  const float d=255.0f; const float r=1.0f/d;
  uint64_t bad1=0,bad2=0,unsafe=0;
  #pragma omp parallel for reduction(+:bad1,bad2,unsafe) schedule(static)
  for(uint64_t u=0;u<(1ull<<32);++u){
    uint32_t b=(uint32_t)u; float x; memcpy(&x,&b,4);
    if(isnan(x)||isinf(x)) continue;
    float ax=fabsf(x);
    int safe=(ax>1e-20f&&ax<1e20f)||b==0;
    if(!safe){unsafe++;continue;}
    float want=x/d;
    float q0=x*r; float e=fmaf(-q0,d,x); float q1=fmaf(e,r,q0);
    float e2=fmaf(-q1,d,x); float q2=fmaf(e2,r,q1);
    uint32_t w,a1,a2; memcpy(&w,&want,4); memcpy(&a1,&q1,4); memcpy(&a2,&q2,4);
    if(w!=a1) bad1++; if(w!=a2) bad2++;
  }
  printf("unsafe %llu bad(one step) %llu bad(two steps) %llu\n",(unsigned long long)unsafe,(unsigned long long)bad1,(unsigned long long)bad2);
  return 0;
}
