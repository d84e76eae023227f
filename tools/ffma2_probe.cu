// ffma2_probe.cu -- throughput of packed fp32 (FFMA2 / FADD2 / FMUL2) against scalar FFMA / FADD on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_probe tools/ffma2_probe.cu && /tmp/ffma2_probe
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(512) probe(float* out, float a, float b) {
  float2 acc[8];
#pragma unroll
  for (int i = 0; i < 8; i++) acc[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  const float2 pa = make_float2(a, a * 1.0001f), pb = make_float2(b, b * 0.9999f);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) {  // scalar FFMA, 3 registers
        acc[i].x = fmaf(acc[i].x, pa.x, pb.x);
        acc[i].y = fmaf(acc[i].y, pa.y, pb.y);
      } else if (MODE == 1) {  // packed
        acc[i] = __ffma2_rn(acc[i], pa, pb);
      } else if (MODE == 2) {  // scalar FADD
        acc[i].x = acc[i].x + pb.x;
        acc[i].y = acc[i].y + pb.y;
      } else if (MODE == 3) {
        acc[i] = __fadd2_rn(acc[i], pb);
      } else if (MODE == 4) {  // scalar FMUL
        acc[i].x = acc[i].x * pa.x;
        acc[i].y = acc[i].y * pa.y;
      } else {
        acc[i] = __fmul2_rn(acc[i], pa);
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, float* d) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int grid = 148 * 4;
  probe<MODE><<<grid, 512>>>(d, 1.0000001f, 1e-9f);
  cudaEventRecord(e0);
  for (int r = 0; r < 5; r++) probe<MODE><<<grid, 512>>>(d, 1.0000001f, 1e-9f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double ops = 5.0 * grid * 512.0 * ITERS * 16.0;  // scalar fp32 operations (an fma counted once)
  printf("%-12s %8.3f ms  %7.2f Tops/s  (%5.1f per clk per SM at 1.9 GHz)\n", name, ms / 5, ops / (ms * 1e-3) / 1e12,
         ops / (ms * 1e-3) / 148 / 1.9e9);
}

int main() {
  float* d;
  cudaMalloc(&d, 148 * 4 * 512 * sizeof(float));
  run<0>("FFMA", d);
  run<1>("FFMA2", d);
  run<2>("FADD", d);
  run<3>("FADD2", d);
  run<4>("FMUL", d);
  run<5>("FMUL2", d);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
