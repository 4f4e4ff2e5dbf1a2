// does sm_100a have packed fp32 FMA (fma.rn.f32x2)?  development probe
#include <cstdint>
__global__ void k(const float2* a, const float2* b, const float2* c, float2* o)
{
    const int i = threadIdx.x;
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<const unsigned long long*>(a + i);
    rb = *reinterpret_cast<const unsigned long long*>(b + i);
    rc = *reinterpret_cast<const unsigned long long*>(c + i);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    *reinterpret_cast<unsigned long long*>(o + i) = rd;
}
