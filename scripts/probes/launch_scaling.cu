// Probe: how does kernel-launch throughput scale with host threads launching into their own streams?
// nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o launch_scaling launch_scaling.cu -lpthread
#include <cstdio>
#include <thread>
#include <vector>
#include <chrono>
#include <cuda_runtime.h>
struct Args { double a[40]; };
__global__ void tiny(Args a, double * out) { if (a.a[0] < 0) out[0] = a.a[1]; }
int main()
{
  double * d; cudaMalloc(&d, 8);
  const int per_thread_streams = 4, launches = 20000;
  for (int T : {1, 2, 4, 8, 16})
  {
    std::vector<cudaStream_t> st(T * per_thread_streams);
    for (auto & s : st) cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t)
      th.emplace_back([&, t]() {
        cudaSetDevice(0);
        Args a; a.a[0] = 1;
        for (int i = 0; i < launches / T; ++i)
          tiny<<<1, 32, 0, st[t * per_thread_streams + i % per_thread_streams]>>>(a, d);
      });
    for (auto & x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    cudaDeviceSynchronize();
    auto t2 = std::chrono::steady_clock::now();
    printf("threads %2d: issue %.2f us/launch aggregate, drained %.2f us/launch\n", T,
           std::chrono::duration<double, std::micro>(t1 - t0).count() / launches,
           std::chrono::duration<double, std::micro>(t2 - t0).count() / launches);
    for (auto & s : st) cudaStreamDestroy(s);
  }
  return 0;
}
