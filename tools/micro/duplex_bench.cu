// dev microbenchmark: one bench.py e2e step's PCIe traffic (384 MB host->device, 576 MB device->host, at the same time), with N
// processes (one per GPU) running together: device->host by the copy engine or by SM stores into the pinned buffer.
//   nvcc -arch=sm_100a -O3 -o duplex_bench duplex_bench.cu ;  for g in 0..N-1: ./duplex_bench $g <unix start time> &
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>

__global__ void k_copy(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
static double now() { return std::chrono::duration<double>(std::chrono::system_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv)
{
    const int dev = argc > 1 ? atoi(argv[1]) : 0;
    const double start = argc > 2 ? atof(argv[2]) : 0.0;
    const size_t in_bytes = 384ull << 20, out_bytes = 576ull << 20;
    cudaSetDevice(dev);
    void *h_in, *h_out, *d_in, *d_out;
    cudaHostAlloc(&h_in, in_bytes, cudaHostAllocDefault);
    cudaHostAlloc(&h_out, out_bytes, cudaHostAllocDefault);  // plain pinned memory: device-accessible through UVA
    cudaMalloc(&d_in, in_bytes);
    cudaMalloc(&d_out, out_bytes);
    cudaMemset(d_out, 1, out_bytes);
    cudaStream_t s_in, s_out;
    cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking);
    const int chunks = 16;
    for (int mode = 0; mode < 3; ++mode) {  // 0: CE both ways; 1: CE in, SM stores out (296 blocks); 2: CE in, SM stores out (74 blocks)
        for (int rep = 0; rep < 2; ++rep) {
            while (now() < start + 4.0 * mode + 2.0 * rep) std::this_thread::sleep_for(std::chrono::microseconds(200));
            const double t0 = now();
            for (int c = 0; c < chunks; ++c) {  // chunked like the pipeline
                const size_t io = c * (in_bytes / chunks), oo = c * (out_bytes / chunks);
                cudaMemcpyAsync((char*)d_in + io, (char*)h_in + io, in_bytes / chunks, cudaMemcpyHostToDevice, s_in);
                if (mode == 0) cudaMemcpyAsync((char*)h_out + oo, (char*)d_out + oo, out_bytes / chunks, cudaMemcpyDeviceToHost, s_out);
                else k_copy<<<mode == 1 ? 296 : 74, 256, 0, s_out>>>((const uint4*)((char*)d_out + oo), (uint4*)((char*)h_out + oo), out_bytes / chunks / 16);
            }
            cudaStreamSynchronize(s_in);
            const double t_in = now() - t0;
            cudaStreamSynchronize(s_out);
            const double t = now() - t0;
            if (rep == 1) printf("gpu %d mode %d: step %.1f ms (h2d done at %.1f ms) -> %.0f Mrays/s per GPU\n", dev, mode, t * 1e3, t_in * 1e3, 16.777216 / t);
        }
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("gpu %d: %s\n", dev, cudaGetErrorString(e));
    return 0;
}
