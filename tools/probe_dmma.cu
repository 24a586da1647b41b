// Microbenchmark: FP64 tensor-core (mma.sync m8n8k4 f64, SASS DMMA) throughput on sm_100a against the DFMA pipe, alone,
// fed from shared memory with the record-gather pattern of the Schur pair slots, and mixed with DFMA warps.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/probe_dmma tools/probe_dmma.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

template <int NACC>
__global__ void __launch_bounds__(512) k_dmma(int iters, double* sink) {
  double c[NACC][2];
#pragma unroll
  for (int k = 0; k < NACC; ++k) c[k][0] = c[k][1] = 0.0;
  const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < NACC; ++k) dmma(c[k][0], c[k][1], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < NACC; ++k) s += c[k][0] + c[k][1];
  if (s == 123.456) sink[0] = s;
}

// DFMA reference: 8 chains per thread
__global__ void __launch_bounds__(512) k_dfma(int iters, double* sink) {
  double a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = 1.0 + k + threadIdx.x * 1e-9;
  const double m = 1.0 - 1e-12, c = 1e-13;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] = a[k] * m + c;
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  if (s == 123.456) sink[0] = s;
}

// warps with (warp id % mix) == 0 run DFMA, the others DMMA: do the two share one pipe?
__global__ void __launch_bounds__(512) k_mixed(int iters, int mix, double* sink) {
  const int wid = threadIdx.x >> 5;
  double s = 0.0;
  if (wid % mix == 0) {
    double a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = 1.0 + k + threadIdx.x * 1e-9;
    const double m = 1.0 - 1e-12, c = 1e-13;
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = a[k] * m + c;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k];
  } else {
    double c[4][2];
#pragma unroll
    for (int k = 0; k < 4; ++k) c[k][0] = c[k][1] = 0.0;
    const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int k = 0; k < 4; ++k) dmma(c[k][0], c[k][1], a, b);
#pragma unroll
    for (int k = 0; k < 4; ++k) s += c[k][0] + c[k][1];
  }
  if (s == 123.456) sink[0] = s;
}

// Schur pair-slot pattern: every warp walks "entries"; per entry it gathers two 18-double records (6x3 W blocks, padded
// record stride REC doubles) from shared memory as m8n8k4 fragments (lane l -> W[l/4][l%4], zero outside 6x3) and
// issues one DMMA into one of NCH accumulator chains (NCH slots in flight per warp).
template <int NCH, int REC>
__global__ void __launch_bounds__(512) k_dmma_gather(int entries, int n_rec, double* sink) {
  extern __shared__ double rec[];
  for (int i = threadIdx.x; i < n_rec * REC; i += blockDim.x) rec[i] = (i % REC) < 18 ? 1e-3 * (i % 97) : 0.0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int row = lane >> 2, col = lane & 3;
  const bool in = row < 6 && col < 3;
  const int off = in ? 3 * row + col : (REC > 18 ? 18 : 0);   // REC > 18: padding lanes read the record's zero pad (no branch)
  double c[NCH][2];
#pragma unroll
  for (int k = 0; k < NCH; ++k) c[k][0] = c[k][1] = 0.0;
  unsigned st = 12345u + 977u * (blockIdx.x * 16 + wid);
  for (int e = 0; e < entries; e += NCH) {
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      st = st * 1664525u + 1013904223u;                 // same sequence in every lane: warp-uniform record ids
      const int ri = (st >> 8) & (n_rec - 1), rj = (st >> 20) & (n_rec - 1);   // n_rec is a power of two
      double a, b;
      if (REC > 18) { a = rec[ri * REC + off]; b = rec[rj * REC + off]; }
      else { a = in ? rec[ri * REC + off] : 0.0; b = in ? rec[rj * REC + off] : 0.0; }
      dmma(c[k][0], c[k][1], a, b);
    }
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < NCH; ++k) s += c[k][0] + c[k][1];
  if (s == 123.456) sink[0] = s;
}

template <class F>
static float time_ms(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  return best;
}

int main() {
  int sm = 0;
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
  double* sink;
  cudaMalloc(&sink, 64);
  const int iters = 4096;
  printf("SMs %d\n", sm);
  for (int threads : {128, 256, 512}) {
    for (int bps : {1, 2, 4}) {
      if (threads * bps > 2048) continue;
      const int blocks = sm * bps;
      float ms = time_ms([&] { k_dfma<<<blocks, threads>>>(iters, sink); });
      printf("DFMA            threads %4d x %d/SM : %7.2f TFLOP/s\n", threads, bps, 2.0 * 32 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12);
      ms = time_ms([&] { k_dmma<1><<<blocks, threads>>>(iters * 4, sink); });
      printf("DMMA m8n8k4 x1  threads %4d x %d/SM : %7.2f TFLOP/s\n", threads, bps, 512.0 * iters * 4 * (double)blocks * (threads / 32) / (ms * 1e-3) / 1e12);
      ms = time_ms([&] { k_dmma<4><<<blocks, threads>>>(iters, sink); });
      printf("DMMA m8n8k4 x4  threads %4d x %d/SM : %7.2f TFLOP/s\n", threads, bps, 512.0 * iters * 4 * (double)blocks * (threads / 32) / (ms * 1e-3) / 1e12);
      ms = time_ms([&] { k_dmma<8><<<blocks, threads>>>(iters, sink); });
      printf("DMMA m8n8k4 x8  threads %4d x %d/SM : %7.2f TFLOP/s\n", threads, bps, 512.0 * iters * 8 * (double)blocks * (threads / 32) / (ms * 1e-3) / 1e12);
    }
  }
  for (int mix : {2, 4}) {
    const int blocks = sm, threads = 512;
    float ms = time_ms([&] { k_mixed<<<blocks, threads>>>(iters, mix, sink); });
    const double wf = (threads / 32) / (double)mix, wm = threads / 32 - wf;
    printf("mixed 1/%d DFMA warps: DFMA %6.2f + DMMA %6.2f TFLOP/s (same launch)\n", mix,
           2.0 * 32 * iters * blocks * wf * 32 / (ms * 1e-3) / 1e12, 512.0 * iters * 4 * blocks * wm / (ms * 1e-3) / 1e12);
  }
  {
    const int blocks = sm, threads = 512, entries = 1 << 16;
    cudaFuncSetAttribute(k_dmma_gather<1, 18>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_dmma_gather<2, 18>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_dmma_gather<4, 18>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_dmma_gather<4, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int n_rec = 1024;
    auto rep = [&](const char* name, float ms) {
      const double ent = (double)entries * blocks * (threads / 32);
      printf("%s: %7.2f G entries/s  (%.2f clk/entry/SM at 1.965 GHz; useful %6.2f TFLOP/s at 216 flop/entry)\n", name,
             ent / (ms * 1e-3) / 1e9, 1.965e9 * sm / (ent / (ms * 1e-3)), 216.0 * ent / (ms * 1e-3) / 1e12);
    };
    rep("gather+DMMA 1 chain  stride 18", time_ms([&] { k_dmma_gather<1, 18><<<blocks, threads, n_rec * 18 * 8>>>(entries, n_rec, sink); }));
    rep("gather+DMMA 2 chains stride 18", time_ms([&] { k_dmma_gather<2, 18><<<blocks, threads, n_rec * 18 * 8>>>(entries, n_rec, sink); }));
    rep("gather+DMMA 4 chains stride 18", time_ms([&] { k_dmma_gather<4, 18><<<blocks, threads, n_rec * 18 * 8>>>(entries, n_rec, sink); }));
    cudaFuncSetAttribute(k_dmma_gather<1, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_dmma_gather<2, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_dmma_gather<8, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    rep("gather+DMMA 1 chain  stride 20", time_ms([&] { k_dmma_gather<1, 20><<<blocks, threads, n_rec * 20 * 8>>>(entries, n_rec, sink); }));
    rep("gather+DMMA 2 chains stride 20", time_ms([&] { k_dmma_gather<2, 20><<<blocks, threads, n_rec * 20 * 8>>>(entries, n_rec, sink); }));
    rep("gather+DMMA 8 chains stride 20", time_ms([&] { k_dmma_gather<8, 20><<<blocks, threads, n_rec * 20 * 8>>>(entries, n_rec, sink); }));
    rep("gather+DMMA 4 chains stride 20, 256 thr", time_ms([&] { k_dmma_gather<4, 20><<<blocks, 256, n_rec * 20 * 8>>>(entries * 2, n_rec, sink); }));
    rep("gather+DMMA 4 chains stride 20", time_ms([&] { k_dmma_gather<4, 20><<<blocks, threads, n_rec * 20 * 8>>>(entries, n_rec, sink); }));
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
