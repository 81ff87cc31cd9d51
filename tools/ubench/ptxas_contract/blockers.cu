#include <cstdint>
__global__ void k_ftzmul(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* o)
{
    unsigned long long p, r;
    asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(p) : "l"(a[threadIdx.x]), "l"(b[threadIdx.x]));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(c[threadIdx.x]));
    o[threadIdx.x] = r;
}
__global__ void k_ftzadd(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* o)
{
    unsigned long long p, r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(p) : "l"(a[threadIdx.x]), "l"(b[threadIdx.x]));
    asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(c[threadIdx.x]));
    o[threadIdx.x] = r;
}
__global__ void k_volatile(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* o)
{
    unsigned long long p, r;
    asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(p) : "l"(a[threadIdx.x]), "l"(b[threadIdx.x]));
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(c[threadIdx.x]));
    o[threadIdx.x] = r;
}
__global__ void k_twouse(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* o)
{   // two products added to each other: which one is fused?
    unsigned long long p, q, r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(p) : "l"(a[threadIdx.x]), "l"(b[threadIdx.x]));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(q) : "l"(c[threadIdx.x]), "l"(b[threadIdx.x]));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(q));
    o[threadIdx.x] = r;
}
