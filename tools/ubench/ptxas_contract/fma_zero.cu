#include <cstdint>
__global__ void k1(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* o)
{   // mul.rn.f32x2 feeding add.rn.f32x2
    unsigned long long p, r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(p) : "l"(a[threadIdx.x]), "l"(b[threadIdx.x]));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(c[threadIdx.x]));
    o[threadIdx.x] = r;
}
__global__ void k2(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* o)
{   // product as fma(a, b, -0) feeding add.rn.f32x2
    unsigned long long p, r, nz = 0x8000000080000000ull;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(p) : "l"(a[threadIdx.x]), "l"(b[threadIdx.x]), "l"(nz));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(c[threadIdx.x]));
    o[threadIdx.x] = r;
}
