#include <cuda_runtime.h>
typedef unsigned long long f2;
__device__ __forceinline__ f2 pack(float lo, float hi){f2 r; asm("mov.b64 %0, {%1, %2};":"=l"(r):"f"(lo),"f"(hi)); return r;}
__device__ __forceinline__ f2 mul2(f2 a,f2 b){f2 d; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ f2 add2(f2 a,f2 b){f2 d; asm("add.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ f2 sub2(f2 a,f2 b){f2 d; asm("sub.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__global__ void k(const f2* a, const f2* b, const f2* c, f2* o, f2* o2, f2*o3){
  int i = threadIdx.x;
  o[i] = add2(mul2(a[i], b[i]), c[i]);
  o2[i] = sub2(c[i], mul2(a[i], b[i]));
  o3[i] = sub2(a[i], b[i]);
}
__global__ void ks(const float* a, const float* b, const float* c, float* o){
  int i = threadIdx.x;
  o[i] = __fadd_rn(__fmul_rn(a[i], b[i]), c[i]);
}
