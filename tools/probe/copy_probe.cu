// Practical ceiling of a memory-bound launch of K_post's size: a plain 128-bit grid-stride copy kernel (read X MB, write X MB) between
// rotating buffers (nothing L2-resident), launched from a CUDA graph and event-timed exactly like bench.py times the task kernels.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/copy_probe tools/probe/copy_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>

__global__ void __launch_bounds__(256) copy_kernel(const float4* __restrict__ src, float4* __restrict__ dst, size_t n4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) dst[i] = __ldg(src + i);
}

int main() {
  cudaStream_t st;
  cudaStreamCreate(&st);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("{");
  const int sizes_mb[] = {5, 10, 20, 40, 80, 160};
  for (int si = 0; si < 6; ++si) {
    const int mb = sizes_mb[si];
    const size_t bytes = (size_t)mb * 1000 * 1000, n4 = bytes / 16;
    const int pairs = 640 / mb < 2 ? 2 : 640 / mb;
    std::vector<float4*> src(pairs), dst(pairs);
    for (int p = 0; p < pairs; ++p) {
      cudaMalloc(&src[p], bytes);
      cudaMalloc(&dst[p], bytes);
      cudaMemsetAsync(src[p], 1, bytes, st);
    }
    for (int cta_per_sm = 4; cta_per_sm <= 8; cta_per_sm += 4) {
      const int grid = sms * cta_per_sm;
      cudaGraph_t graph;
      cudaGraphExec_t exec;
      cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal);
      for (int p = 0; p < pairs; ++p) copy_kernel<<<grid, 256, 0, st>>>(src[p], dst[p], n4);
      cudaStreamEndCapture(st, &graph);
      cudaGraphInstantiate(&exec, graph, 0);
      for (int w = 0; w < 3; ++w) cudaGraphLaunch(exec, st);
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
      const int reps = 20;
      cudaEventRecord(e0, st);
      for (int r = 0; r < reps; ++r) cudaGraphLaunch(exec, st);
      cudaEventRecord(e1, st);
      cudaStreamSynchronize(st);
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      const double us = ms * 1e3 / (reps * pairs);
      printf("%s\"%dMB+%dMB grid %dx%d\": {\"us\": %.2f, \"GB/s\": %.0f}", (si || cta_per_sm > 4) ? ", " : "", mb, mb, sms, cta_per_sm, us,
             2.0 * bytes / (us * 1e-6) / 1e9);
      cudaGraphExecDestroy(exec);
      cudaGraphDestroy(graph);
    }
    for (int p = 0; p < pairs; ++p) { cudaFree(src[p]); cudaFree(dst[p]); }
  }
  printf("}\n");
  return cudaGetLastError() != cudaSuccess;
}
