// dev microbenchmark: throughput of incoherent 32-byte gathers (one record per lane, every lane a
// different line) through the LSU path (LDG.256, 2x LDG.128) and the texture path (2x tex1Dfetch<uint4>),
// on an L2-resident table.  Prints lane-gathers per clock per SM.   nvcc -arch=sm_100a -O3 gather_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s; }

template <int MODE>
__global__ void __launch_bounds__(128) k(const uint4* __restrict__ tab, cudaTextureObject_t tex, uint32_t mask, int iters, uint32_t* sink)
{
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int i = 0; i < iters; ++i) {
        const uint32_t idx = (lcg(s) >> 4) & mask;  // record index (32 B records)
        if (MODE == 0) {
            uint32_t a, b, c, d, e, f, g, h;
            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(tab + 2 * (size_t)idx));
            acc += a ^ h;
        }
        else if (MODE == 1) {
            const uint4 x = __ldg(tab + 2 * (size_t)idx), y = __ldg(tab + 2 * (size_t)idx + 1);
            acc += x.x ^ y.w;
        }
        else if (MODE == 2) {
            const uint4 x = tex1Dfetch<uint4>(tex, 2 * idx), y = tex1Dfetch<uint4>(tex, 2 * idx + 1);
            acc += x.x ^ y.w;
        }
        else if (MODE == 3) {  // alternate LSU / TEX
            if (i & 1) {
                const uint4 x = tex1Dfetch<uint4>(tex, 2 * idx), y = tex1Dfetch<uint4>(tex, 2 * idx + 1);
                acc += x.x ^ y.w;
            }
            else {
                uint32_t a, b, c, d, e, f, g, h;
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(tab + 2 * (size_t)idx));
                acc += a ^ h;
            }
        }
        else if (MODE == 4) {  // 16 B only, LSU
            const uint4 x = __ldg(tab + 2 * (size_t)idx);
            acc += x.x;
        }
        else if (MODE == 5) {  // 16 B only, TEX
            const uint4 x = tex1Dfetch<uint4>(tex, 2 * idx);
            acc += x.x;
        }
        else if (MODE == 6) {  // 2 TEX : 1 LSU
            if (i % 3) {
                const uint4 x = tex1Dfetch<uint4>(tex, 2 * idx), y = tex1Dfetch<uint4>(tex, 2 * idx + 1);
                acc += x.x ^ y.w;
            }
            else {
                uint32_t a, b, c, d, e, f, g, h;
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(tab + 2 * (size_t)idx));
                acc += a ^ h;
            }
        }
        else if (MODE == 7) {  // 64-byte records (two adjacent sectors of one line): what a 4-wide node would cost
            const uint4* q = tab + 4 * (size_t)(idx >> 1);
            uint32_t a, b, c, d, e, f, g, h, a2, b2, c2, d2, e2, f2, g2, h2;
            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(q));
            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a2), "=r"(b2), "=r"(c2), "=r"(d2), "=r"(e2), "=r"(f2), "=r"(g2), "=r"(h2) : "l"(q + 2));
            acc += a ^ h ^ a2 ^ h2;
        }
        else if (MODE == 8) {  // 128-byte records (a whole line)
            const uint4* q = tab + 8 * (size_t)(idx >> 2);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t a, b, c, d, e, f, g, h;
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(q + 2 * j));
                acc += a ^ h;
            }
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <int MODE>
void run(const char* name, const uint4* tab, cudaTextureObject_t tex, uint32_t mask, uint32_t* sink, int sms, double ghz)
{
    const int iters = 2000, blocks = sms * 12;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k<MODE><<<blocks, 128>>>(tab, tex, mask, 200, sink);
    cudaEventRecord(a);
    k<MODE><<<blocks, 128>>>(tab, tex, mask, iters, sink);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double gathers = (double)blocks * 128 * iters;
    printf("%-28s records=%8u  %7.3f ms  %6.1f Ggathers/s  %.3f lane-gathers/clk/SM (at %.2f GHz)  err=%s\n", name, mask + 1, ms, gathers / ms / 1e6,
           gathers / (ms * 1e-3) / sms / (ghz * 1e9), ghz, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double ghz = clk / 1e6;
    for (uint32_t recs : {1u << 10, 1u << 20, 1u << 21}) {  // 32 KB (L1-resident), 32 MB, 64 MB (L2-resident)
        uint4* tab; cudaMalloc(&tab, (size_t)recs * 32); cudaMemset(tab, 1, (size_t)recs * 32);
        uint32_t* sink; cudaMalloc(&sink, 4);
        cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = tab; rd.res.linear.desc = cudaCreateChannelDesc<uint4>(); rd.res.linear.sizeInBytes = (size_t)recs * 32;
        cudaTextureDesc td = {}; td.readMode = cudaReadModeElementType;
        cudaTextureObject_t tex = 0; cudaCreateTextureObject(&tex, &rd, &td, nullptr);
        run<0>("LSU  LDG.256", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<1>("LSU  2x LDG.128", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<2>("TEX  2x tex1Dfetch<uint4>", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<3>("alternate LDG.256 / 2x tex", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<6>("1 LDG.256 : 2 (2x tex)", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<4>("LSU  LDG.128 (16 B)", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<5>("TEX  tex1Dfetch (16 B)", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<7>("LSU  2x LDG.256 (64 B record)", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        run<8>("LSU  4x LDG.256 (128 B line)", tab, tex, recs - 1, sink, p.multiProcessorCount, ghz);
        cudaDestroyTextureObject(tex); cudaFree(tab); cudaFree(sink);
    }
    return 0;
}
