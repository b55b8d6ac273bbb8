// Probe: how fast can a B200 touch the AllegroKuka post-step access pattern (no arithmetic)?
//   mode 0: gather only (palm row, 4 tip rows, object root row, goal row, init z) with 4-byte LDG, lane per float
//   mode 1: mode 0 + contiguous reads (dof 184 B, kp 48, scales 12, coef 16, state 53, acc 48 per env)
//   mode 2: mode 1 + writes (obs 468 + 145 B per env)
//   mode 3: writes only;  mode 4: contiguous reads only; mode 5: gather as whole 128-byte lines (LDG.128 x 8 lanes per line)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_probe gather_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
constexpr int NB = 27, NA = 4;
struct Bufs { const float *rb, *root, *goal, *init, *cont; float *out; float* sink; };

__global__ void __launch_bounds__(256) probe(Bufs b, int n, int mode, int cont_per_env, int out_per_env) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;
  float acc = 0.f;
  if (mode <= 2) {
    // 16 lanes per 13-float row: element index space [n * 8 rows * 16]; U independent loads in flight per thread
    const long total = (long)n * 8 * 16;
    constexpr int U = 8;
    for (long i0 = tid; i0 < total; i0 += (long)nthreads * U) {
      float v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + (long)u * nthreads;
        v[u] = 0.f;
        if (i >= total) continue;
        const int c = i & 15; const long r = i >> 4; const int row = r & 7; const long e = r >> 3;
        const float* src; int w = 13;
        if (row == 0) src = b.rb + (e * NB + 7) * 13;
        else if (row <= 4) { w = 7; src = b.rb + (e * NB + 7 + 4 * row) * 13; }
        else if (row == 5) src = b.root + (e * NA + 1) * 13;
        else if (row == 6) { w = 7; src = b.goal + e * 13; }
        else { w = 3; src = b.init + e * 13; }
        if (c < w) v[u] = __ldg(src + c);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc += v[u];
    }
  }
  if (mode == 5) {
    // sparse whole lines: the 128-byte lines that cover the needed rows, 8 lanes x float4 per line, U in flight
    const long total = (long)n * 7 * 2 * 8;
    constexpr int U = 4;
    for (long i0 = tid; i0 < total; i0 += (long)nthreads * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + (long)u * nthreads;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i >= total) continue;
        const int l = i & 7; const long r = i >> 3; const int half = r & 1; const long rr = r >> 1; const int row = rr % 7; const long e = rr / 7;
        const float* src;
        if (row == 0) src = b.rb + (e * NB + 7) * 13;
        else if (row <= 4) src = b.rb + (e * NB + 7 + 4 * row) * 13;
        else if (row == 5) src = b.root + (e * NA + 1) * 13;
        else src = b.goal + e * 13;
        const size_t byte = (size_t)((const char*)src - (const char*)0);
        const float4* line = (const float4*)((byte & ~(size_t)127) + (size_t)half * 128);
        const size_t endb = byte + (row == 0 || row == 5 ? 52 : 28) - 1;
        if (half == 1 && (endb & ~(size_t)127) == (byte & ~(size_t)127)) continue;
        v[u] = __ldg(line + l);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc += v[u].x + v[u].w;
    }
  }
  if (mode == 6 || mode == 7) {
    // mode 6: stream ALL of rb + root + goal; mode 7: per env the contiguous span of bodies 7..23 (16-byte aligned) + root + goal
    constexpr int U = 4;
    const long per_env = 55;
    const long total = (mode == 6) ? (long)n * NB * 52 / 16 : (long)n * per_env;
    for (long i0 = tid; i0 < total; i0 += (long)nthreads * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + (long)u * nthreads;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i >= total) continue;
        if (mode == 6) v[u] = __ldg((const float4*)b.rb + i);
        else {
          const long e = i / 55; const int q = i - e * 55;
          const size_t byte = ((size_t)e * NB * 52 + 364) & ~(size_t)15;
          v[u] = __ldg((const float4*)((const char*)b.rb + byte) + q);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc += v[u].x + v[u].w;
    }
    const long t2 = (long)n * (NA * 52 + 52) / 16;
    for (long i = tid; i < t2; i += nthreads) {
      const long nroot = (long)n * NA * 52 / 16; const float4 v = (i < nroot) ? __ldg((const float4*)b.root + i) : __ldg((const float4*)b.goal + (i - nroot));
      acc += v.x;
    }
  }
  if (mode == 1 || mode == 2 || mode == 4) {
    const long total = (long)n * cont_per_env / 16;   // bytes -> float4
    const float4* src = (const float4*)b.cont;
    for (long i = tid; i < total; i += nthreads) { const float4 v = __ldg(src + i); acc += v.x + v.y + v.z + v.w; }
  }
  if (mode == 2 || mode == 3) {
    const long total = (long)n * out_per_env / 16;
    float4* dst = (float4*)b.out;
    for (long i = tid; i < total; i += nthreads) dst[i] = make_float4(acc, 1.f, 2.f, 3.f);
  }
  if (acc == 123.456f) *b.sink = acc;
}

int main(int argc, char** argv) {
  const int n = 32768, sets = 6;       // rotate over `sets` copies so nothing is L2-resident
  const int cont = 368, outw = 624;    // bytes per env: cont 361 B -> 364... rounded to 368, out 613 -> 624 (16-byte multiples)
  float *rb, *root, *goal, *init, *contb, *out, *sink;
  CK(cudaMalloc(&rb, (size_t)sets * n * NB * 52)); CK(cudaMalloc(&root, (size_t)sets * n * NA * 52));
  CK(cudaMalloc(&goal, (size_t)sets * n * 52)); CK(cudaMalloc(&init, (size_t)sets * n * 52));
  CK(cudaMalloc(&contb, (size_t)sets * n * cont)); CK(cudaMalloc(&out, (size_t)sets * n * outw)); CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(rb, 0, (size_t)sets * n * NB * 52)); CK(cudaMemset(root, 0, (size_t)sets * n * NA * 52));
  CK(cudaMemset(goal, 0, (size_t)sets * n * 52)); CK(cudaMemset(init, 0, (size_t)sets * n * 52)); CK(cudaMemset(contb, 0, (size_t)sets * n * cont));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grids[] = {148 * 8, 148 * 16};
  for (int mode = 6; mode <= 7; ++mode)
    for (int g : grids) {
      const int reps = 60;
      for (int w = 0; w < 2; ++w)
        for (int i = 0; i < reps; ++i) {
          if (w == 1 && i == 0) CK(cudaEventRecord(e0));
          const int s = i % sets;
          Bufs b{rb + (size_t)s * n * NB * 13, root + (size_t)s * n * NA * 13, goal + (size_t)s * n * 13, init + (size_t)s * n * 13,
                 contb + (size_t)s * n * cont / 4, out + (size_t)s * n * outw / 4, sink};
          probe<<<g, 256>>>(b, n, mode, cont, outw);
        }
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      printf("mode %d grid %5d : %.2f us per launch\n", mode, g, ms * 1e3 / reps);
    }
  return 0;
}
