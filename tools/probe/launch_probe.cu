// Launch-floor probe: what does a kernel node cost inside a CUDA graph on this GPU, before it does any work?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/launch_probe tools/probe/launch_probe.cu
// Prints us per node for chains of 16 kernel nodes (replayed 400 times, CUDA events), for: empty kernels of the K_post / K_pre
// launch shapes, alternating shared-memory carve-outs, a one-round-trip "touch" kernel over a rotating > L2 footprint, and
// the same chains with programmatic dependent launch (griddepcontrol).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(128) k_null(int pdl) {
  if (pdl) { asm volatile("griddepcontrol.launch_dependents;"); asm volatile("griddepcontrol.wait;" ::: "memory"); }
}
// one dependent DRAM round trip per thread + one store (rotating footprint)
__global__ void __launch_bounds__(128) k_touch(const float4* __restrict__ src, float4* __restrict__ dst, size_t n, int pdl) {
  if (pdl) { asm volatile("griddepcontrol.launch_dependents;"); asm volatile("griddepcontrol.wait;" ::: "memory"); }
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}
// per-thread: `per` float4 loads in flight, then stores (streaming copy of a given size)
__global__ void __launch_bounds__(256) k_copy(const float4* __restrict__ src, float4* __restrict__ dst, size_t n, int pdl) {
  if (pdl) { asm volatile("griddepcontrol.launch_dependents;"); asm volatile("griddepcontrol.wait;" ::: "memory"); }
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) if (i + k * stride < n) v[k] = src[i + k * stride];
#pragma unroll
    for (int k = 0; k < 4; ++k) if (i + k * stride < n) dst[i + k * stride] = v[k];
  }
}

// ---- parameter-bank probe: what does touching L distinct 128-byte lines of a by-value parameter struct cost per launch? ----
struct BigParams { float v[352]; };              // 1408 bytes = 11 lines, the size of HaKukaParams
__constant__ BigParams c_user;                   // the same bytes in a user __constant__ bank (unchanged between launches)
template <int L>
__global__ void __launch_bounds__(128) k_param(const __grid_constant__ BigParams p, float* out) {
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < L; ++i) acc += p.v[32 * i + 1];        // compile-time offsets into the parameter bank, one per line
  if (acc == 12345.f) out[threadIdx.x] = acc;
}
template <int L>
__global__ void __launch_bounds__(128) k_const(float* out) {
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < L; ++i) acc += c_user.v[32 * i + 1];
  if (acc == 12345.f) out[threadIdx.x] = acc;
}
template <int L>
__global__ void __launch_bounds__(128) k_gmem(const float* __restrict__ tab, float* out) {
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < L; ++i) acc += __ldg(tab + 32 * i + 1);  // independent loads: one round trip
  if (acc == 12345.f) out[threadIdx.x] = acc;
}

template <typename F>
float time_chain(cudaStream_t st, int nodes, F launch_node) {
  cudaGraph_t g; cudaGraphExec_t ge;
  CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  for (int i = 0; i < nodes; ++i) launch_node(i);
  CK(cudaStreamEndCapture(st, &g));
  CK(cudaGraphInstantiate(&ge, g, 0));
  for (int i = 0; i < 5; ++i) CK(cudaGraphLaunch(ge, st));
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  const int reps = 400;
  CK(cudaEventRecord(a, st));
  for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ge, st));
  CK(cudaEventRecord(b, st));
  CK(cudaStreamSynchronize(st));
  float ms = 0; CK(cudaEventElapsedTime(&ms, a, b));
  CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
  return ms * 1e3f / (reps * nodes);
}

template <typename... KArgs, typename... Args>
void launch(void (*k)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  CK(cudaLaunchKernelEx(&cfg, k, static_cast<KArgs>(args)...));
}

int main() {
  cudaStream_t st; CK(cudaStreamCreate(&st));
  CK(cudaFuncSetAttribute(k_null, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  const size_t slab = 64u << 20;                 // 64 MB per slab, 16 slabs src + 16 dst = 2 GB footprint
  const int nslab = 16;
  float4 *src, *dst;
  CK(cudaMalloc(&src, slab * nslab)); CK(cudaMalloc(&dst, slab * nslab));
  CK(cudaMemset(src, 1, slab * nslab)); CK(cudaMemset(dst, 0, slab * nslab));
  {
    BigParams hp; for (int i = 0; i < 352; ++i) hp.v[i] = 0.001f * i;
    CK(cudaMemcpyToSymbol(c_user, &hp, sizeof(hp)));
    float* tab; CK(cudaMalloc(&tab, sizeof(hp))); CK(cudaMemcpy(tab, &hp, sizeof(hp), cudaMemcpyHostToDevice));
    float* out = (float*)dst;
    printf("---- parameter lines touched per launch (grid 148 x 128) ----\n");
    printf("param bank   1 line : %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_param<1>, 148, 128, 0, st, false, hp, out); }));
    printf("param bank   4 lines: %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_param<4>, 148, 128, 0, st, false, hp, out); }));
    printf("param bank  11 lines: %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_param<11>, 148, 128, 0, st, false, hp, out); }));
    printf("user const   1 line : %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_const<1>, 148, 128, 0, st, false, out); }));
    printf("user const  11 lines: %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_const<11>, 148, 128, 0, st, false, out); }));
    printf("global ldg  11 lines: %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_gmem<11>, 148, 128, 0, st, false, (const float*)tab, out); }));
    printf("param bank  11 lines, grid 1024: %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_param<11>, 1024, 128, 0, st, false, hp, out); }));
    printf("user const  11 lines, grid 1024: %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_const<11>, 1024, 128, 0, st, false, out); }));
  }
  for (int pdl = 0; pdl < 1; ++pdl) {
    printf("---- %s ----\n", pdl ? "programmatic dependent launch" : "plain stream order");
    for (int grid : {148, 592, 1024, 4096})
      printf("null  grid %4d smem 0      : %.2f us/node\n", grid, time_chain(st, 16, [&](int) { launch(k_null, grid, 128, 0, st, pdl, pdl); }));
    printf("null  grid  592 smem 48K    : %.2f us/node\n", time_chain(st, 16, [&](int) { launch(k_null, 592, 128, 48 * 1024, st, pdl, pdl); }));
    printf("null  592/48K alt 1024/0    : %.2f us/node\n", time_chain(st, 16, [&](int i) {
      if (i & 1) launch(k_null, 1024, 128, 0, st, pdl, pdl); else launch(k_null, 592, 128, 48 * 1024, st, pdl, pdl); }));
    for (int grid : {148, 1024})
      printf("touch grid %4d (1 trip)    : %.2f us/node\n", grid, time_chain(st, 16, [&](int i) {
        launch(k_touch, grid, 128, 0, st, pdl, (const float4*)((char*)src + slab * (i % nslab)), (float4*)((char*)dst + slab * (i % nslab)), (size_t)grid * 128, pdl); }));
    for (size_t mb : {2, 8, 20, 40}) {
      const size_t n = (mb << 20) / 16;
      printf("copy  %2zu MB r + %2zu MB w (1184 x 256): %.2f us/node\n", mb, mb, time_chain(st, 16, [&](int i) {
        launch(k_copy, 1184, 256, 0, st, pdl, (const float4*)((char*)src + slab * (i % nslab)), (float4*)((char*)dst + slab * (i % nslab)), n, pdl); }));
    }
  }
  return 0;
}
