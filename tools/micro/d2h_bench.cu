// dev microbenchmark: device-to-host bandwidth of the copy engine (cudaMemcpyAsync into pinned memory) against SM stores
// into mapped pinned memory, with N processes (one per GPU) running at the same time.
//   nvcc -arch=sm_100a -O3 -o d2h_bench d2h_bench.cu ;  for g in 0..7: ./d2h_bench $g <unix time to start> &
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
    const size_t bytes = 512ull << 20;
    cudaSetDevice(dev);
    void *h = nullptr, *hd = nullptr, *d = nullptr;
    cudaHostAlloc(&h, bytes, cudaHostAllocMapped);
    cudaHostGetDevicePointer(&hd, h, 0);
    cudaMalloc(&d, bytes);
    cudaMemset(d, 1, bytes);
    cudaStream_t s;
    cudaStreamCreate(&s);
    cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s);
    k_copy<<<148 * 4, 256, 0, s>>>((const uint4*)d, (uint4*)hd, bytes / 16);
    cudaStreamSynchronize(s);
    for (int mode = 0; mode < 3; ++mode) {  // 0: copy engine, 1: SM stores (592 blocks), 2: SM stores (148 blocks)
        while (now() < start + 3.0 * mode) std::this_thread::sleep_for(std::chrono::microseconds(200));
        const double t0 = now();
        const int reps = 6;
        for (int r = 0; r < reps; ++r) {
            if (mode == 0) cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s);
            else k_copy<<<mode == 1 ? 148 * 4 : 148, 256, 0, s>>>((const uint4*)d, (uint4*)hd, bytes / 16);
        }
        cudaStreamSynchronize(s);
        const double dt = now() - t0;
        printf("gpu %d mode %d (%s): %.1f GB/s\n", dev, mode, mode == 0 ? "copy engine" : (mode == 1 ? "SM stores x592 blocks" : "SM stores x148 blocks"), reps * bytes / dt / 1e9);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("gpu %d: %s\n", dev, cudaGetErrorString(e));
    return 0;
}
