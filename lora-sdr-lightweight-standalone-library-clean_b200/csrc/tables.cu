// tables.cu -- plan tables built ON THE DEVICE at context creation (the reference builds the
// same tables in lora_demod_init / init / demodulate with libm calls):
//   twiddles  kissfft.hh:24-29     tw[i] = cexpf(j * float(i) * phinc), phinc = float(-2*pi/N)
//   Hann      LoRaDemod.cpp:18-24  0.5f - 0.5f*cosf(2*PI*i/(N-1))
//   chirps    ChirpGenerator.hpp:24-51 (serial float recurrence, one thread per table)
// cosf/sinf are the glibc restatements of glibc_math.cuh, so the tables are bit-identical
// to the reference's (tests/test_gpu_tables.py reads them back and compares).
#include "internal.h"
#include "glibc_math.cuh"

namespace lk {

__global__ void twiddle_hann_kernel(int sf, cf* tw, float* hann)
{
    const int N = 1 << sf;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float phinc = (float)(-2.0 * 0x1.921fb54442d18p+1 / (double)N);   // acos(-1.0) == pi
    const float ang = fx::fmul((float)i, phinc);
    float sn, cs;
    gm::sincosf_g(ang, &sn, &cs);
    tw[i] = fx::mk(cs, sn);
    const float PI_F = 3.14159265358979323846f;
    const float arg = __fdiv_rn(fx::fmul(2.0f * PI_F, (float)i), fx::fsub((float)N, 1.0f));
    hann[i] = fx::fsub(0.5f, fx::fmul(0.5f, gm::cosf_g(arg)));
}

template <int SF>
__global__ void compact_kernel(const cf* tw, cf* ct, int n)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) ct[j] = fx::Plan<SF>::compact_entry(j, tw);
}

template <int SF> static void launch_compact(const cf* tw, cf* ct, int n, cudaStream_t s)
{
    compact_kernel<SF><<<(n + 127) / 128, 128, 0, s>>>(tw, ct, n);
}

int compact_size_of(int sf)
{
    switch (sf) {
    case 2: return fx::Plan<2>::compact_size();   case 3: return fx::Plan<3>::compact_size();
    case 4: return fx::Plan<4>::compact_size();   case 5: return fx::Plan<5>::compact_size();
    case 6: return fx::Plan<6>::compact_size();   case 7: return fx::Plan<7>::compact_size();
    case 8: return fx::Plan<8>::compact_size();   case 9: return fx::Plan<9>::compact_size();
    case 10: return fx::Plan<10>::compact_size(); case 11: return fx::Plan<11>::compact_size();
    case 12: return fx::Plan<12>::compact_size();
    default: return 0;
    }
}

cudaError_t launch_build_tables(int sf, cf* tw, cf* ct, int ct_size, float* hann, cudaStream_t stream)
{
    const int N = 1 << sf;
    twiddle_hann_kernel<<<(N + 127) / 128, 128, 0, stream>>>(sf, tw, hann);
    switch (sf) {
    case 2: launch_compact<2>(tw, ct, ct_size, stream); break;
    case 3: launch_compact<3>(tw, ct, ct_size, stream); break;
    case 4: launch_compact<4>(tw, ct, ct_size, stream); break;
    case 5: launch_compact<5>(tw, ct, ct_size, stream); break;
    case 6: launch_compact<6>(tw, ct, ct_size, stream); break;
    case 7: launch_compact<7>(tw, ct, ct_size, stream); break;
    case 8: launch_compact<8>(tw, ct, ct_size, stream); break;
    case 9: launch_compact<9>(tw, ct, ct_size, stream); break;
    case 10: launch_compact<10>(tw, ct, ct_size, stream); break;
    case 11: launch_compact<11>(tw, ct, ct_size, stream); break;
    case 12: launch_compact<12>(tw, ct, ct_size, stream); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// genChirp(buf, N, osr, NN = N*osr, f0 = 0, down = true, ampl = 1, phase = 0, bw_scale)
// ChirpGenerator.hpp:28-48; the call sites are phy.cpp:205-208 and tests/e2e_chain_test.cpp:83-87.
__global__ void chirp_table_kernel(int N, int osr, float bw_scale, cf* out)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const float PI_F = 3.14159265358979323846f;
    const float fMin = __fdiv_rn(fx::fmul(-PI_F, bw_scale), (float)osr);
    const float fMax = __fdiv_rn(fx::fmul(PI_F, bw_scale), (float)osr);
    const float fStep = __fdiv_rn(fx::fmul(2 * PI_F, bw_scale), (float)(N * osr * osr));
    float f = fx::fadd(fMin, 0.0f);
    float phase = 0.0f;
    const int NN = N * osr;
    for (int i = 0; i < NN; ++i) {
        f = fx::fadd(f, fStep);
        if (f > fMax) f = fx::fsub(f, fx::fsub(fMax, fMin));
        phase = fx::fsub(phase, f);
        float sn, cs;
        gm::sincosf_g(phase, &sn, &cs);
        out[i] = fx::mk(cs, sn);      // ampl == 1: 1.0f*x == x
    }
}

cudaError_t launch_build_chirp(int sf, int osr, float bw_scale, cf* out, cudaStream_t stream)
{
    chirp_table_kernel<<<1, 32, 0, stream>>>(1 << sf, osr, bw_scale, out);
    return cudaGetLastError();
}

} // namespace lk
