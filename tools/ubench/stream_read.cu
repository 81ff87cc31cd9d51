// stream_read.cu -- what HBM read rate does the demodulator's ACCESS PATTERN allow, with no arithmetic to speak of?
// One warp = one pipeline: a ring of S stages of B bytes filled by cp.async.bulk (one leader lane), consumed by the
// warp with 64-bit shared-memory loads and one add per element.  Two orders over the same buffer:
//   packet       warp w walks its own packet (PKT contiguous bytes) in B-byte steps, then the packet n_warps further on
//                (the certified demodulator at SF7-9: 2368 .. 4736 independent sequential streams)
//   interleaved  step i of warp w reads chunk i * n_warps + w (all warps sweep the buffer together, like a copy kernel)
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o stream_read stream_read.cu
// Run:   ./stream_read            (sweeps B, S, warps per SM, both orders; prints GB/s)
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t ph)
{
    asm volatile("{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(b), "r"(ph) : "memory");
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int S>
__global__ void k(const unsigned char* __restrict__ buf, size_t n_chunks, uint32_t B, uint32_t chunks_per_pkt, int interleaved,
                  float* out)
{
    extern __shared__ __align__(128) unsigned char sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
    unsigned char* ring = sm + (size_t)warp * S * B;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + (size_t)W * S * B) + warp * S;
    if (lane == 0) {
        for (int s = 0; s < S; ++s) mbar_init(s32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const size_t n_warps = (size_t)gridDim.x * W, w = (size_t)blockIdx.x * W + warp;
    auto chunk_of = [&](size_t i) -> size_t {               // chunk index of this warp's i-th step
        if (interleaved) return i * n_warps + w;
        const size_t pk = i / chunks_per_pkt, c = i % chunks_per_pkt;
        return (pk * n_warps + w) * chunks_per_pkt + c;
    };
    size_t n_steps;
    if (interleaved) n_steps = (n_chunks > w) ? (n_chunks - w + n_warps - 1) / n_warps : 0;
    else {
        const size_t n_pk = n_chunks / chunks_per_pkt;
        n_steps = ((n_pk > w) ? (n_pk - w + n_warps - 1) / n_warps : 0) * chunks_per_pkt;
    }
    if (lane == 0)
        for (int s = 0; s < S - 1 && (size_t)s < n_steps; ++s) {
            mbar_expect(s32(&bars[s]), B);
            bulk(s32(ring + (size_t)s * B), buf + chunk_of(s) * B, B, s32(&bars[s]));
        }
    float acc0 = 0.f, acc1 = 0.f;
    int stage = 0;
    uint32_t phase = 0;
    for (size_t i = 0; i < n_steps; ++i) {
        __syncwarp();                                  // everyone has finished the slot that is refilled now
        if (lane == 0 && i + S - 1 < n_steps) {
            int st = stage + S - 1; if (st >= S) st -= S;
            mbar_expect(s32(&bars[st]), B);
            bulk(s32(ring + (size_t)st * B), buf + chunk_of(i + S - 1) * B, B, s32(&bars[st]));
        }
        mbar_wait(s32(&bars[stage]), (phase >> stage) & 1u);
        phase ^= 1u << stage;
        const float2* p = reinterpret_cast<const float2*>(ring + (size_t)stage * B);
        for (uint32_t j = lane; j < B / 8; j += 32) { const float2 v = p[j]; acc0 += v.x; acc1 += v.y; }
        stage = stage + 1 == S ? 0 : stage + 1;
    }
    if (acc0 + acc1 == 123.456f) out[0] = acc0;
}

int main()
{
    const size_t bytes = (size_t)32 << 30;
    unsigned char* buf; float* out;
    cudaMalloc(&buf, bytes); cudaMalloc(&out, 4);
    cudaMemset(buf, 0, bytes);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const uint32_t PKT = 66 * 128 * 8;                // SF7 cfg-2 packet: 67584 bytes
    struct Case { uint32_t B; int S; int warps; int ctas; };
    std::vector<Case> cases;
    for (uint32_t B : {2048u, 4096u, 8192u})
        for (int S : {2, 3, 4})
            for (int wc : {8, 16, 24, 32}) {
                const int ctas = 2, warps = wc / ctas;
                const size_t smem = (size_t)warps * S * B + warps * S * 8;
                if (smem * ctas + 2048 * ctas > 227 * 1024) continue;
                cases.push_back({B, S, warps, ctas});
            }
    printf("%8s %2s %5s %12s %12s\n", "B", "S", "w/SM", "packet GB/s", "interl GB/s");
    for (const Case& c : cases) {
        double gbs[2] = {0, 0};
        for (int inter = 0; inter < 2; ++inter) {
            const uint32_t cpp = PKT / c.B ? PKT / c.B : 1;
            const size_t n_chunks = bytes / c.B;
            const size_t smem = (size_t)c.warps * c.S * c.B + c.warps * c.S * 8;
            auto launch = [&]() {
                const dim3 g(sms * c.ctas), b(c.warps * 32);
                switch (c.S) {
                case 2: cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); k<2><<<g, b, smem>>>(buf, n_chunks, c.B, cpp, inter, out); break;
                case 3: cudaFuncSetAttribute(k<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); k<3><<<g, b, smem>>>(buf, n_chunks, c.B, cpp, inter, out); break;
                default: cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); k<4><<<g, b, smem>>>(buf, n_chunks, c.B, cpp, inter, out); break;
                }
            };
            launch();
            cudaDeviceSynchronize();
            float best = 1e30f;
            for (int r = 0; r < 3; ++r) {
                cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (ms < best) best = ms;
            }
            const size_t read = inter ? n_chunks * c.B : (n_chunks / cpp) * cpp * (size_t)c.B;
            gbs[inter] = read / (best * 1e6);
        }
        printf("%8u %2d %5d %12.0f %12.0f\n", c.B, c.S, c.warps * c.ctas, gbs[0], gbs[1]);
        fflush(stdout);
    }
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(err)); return 1; }
    return 0;
}
