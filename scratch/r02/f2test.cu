#include <cuda_runtime.h>
#include <stdint.h>
typedef unsigned long long f2;
__device__ __forceinline__ f2 pack(float lo, float hi){f2 r; asm("mov.b64 %0, {%1, %2};":"=l"(r):"f"(lo),"f"(hi)); return r;}
__device__ __forceinline__ void unpack(f2 v,float&lo,float&hi){asm("mov.b64 {%0, %1}, %2;":"=f"(lo),"=f"(hi):"l"(v));}
__device__ __forceinline__ f2 fma2(f2 a,f2 b,f2 c){f2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ f2 mul2(f2 a,f2 b){f2 d; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ f2 sub2(f2 a,f2 b){f2 d; asm("sub.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ f2 neg2(f2 a){ return a ^ 0x8000000080000000ull; }
__global__ void k(const float4* a, const float4* b, float4* o, float c){
  int i = threadIdx.x;
  float4 x=a[i], y=b[i];
  f2 x0=pack(x.x,x.y), x1=pack(x.z,x.w), y0=pack(y.x,y.y), y1=pack(y.z,y.w);
  f2 cc=pack(c,c);
  f2 r0=fma2(neg2(x0), y0, cc), r1=fma2(x1, neg2(y1), cc);
  r0 = sub2(r0, mul2(x0, cc));
  float4 w; unpack(r0,w.x,w.y); unpack(r1,w.z,w.w);
  o[i]=w;
}
