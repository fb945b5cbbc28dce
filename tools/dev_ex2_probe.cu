// Probe: does sm_100a have a PACKED half-precision exp2 (two results per MUFU op)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -cubin -o ex2_probe.cubin ex2_probe.cu && cuobjdump -sass ex2_probe.cubin
#include <stdint.h>
__global__ void k(const float2* x, uint32_t* y, uint32_t* z) {
  float2 v = x[threadIdx.x];
  uint32_t h, e, b, eb;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(v.y), "f"(v.x));
  asm("ex2.approx.f16x2 %0, %1;" : "=r"(e) : "r"(h));          // (.ftz is rejected by ptxas for f16x2)
  y[threadIdx.x] = e;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(b) : "f"(v.y), "f"(v.x));
  asm("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(eb) : "r"(b));
  z[threadIdx.x] = eb;
}
