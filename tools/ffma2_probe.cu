// Micro-benchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2, sm_100) issue rate, latency, and co-issue with FMNMX.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_probe tools/ffma2_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 4096;

template <int MODE>
__global__ void __launch_bounds__(448, 1) k(float* out, float a, float b) {
  float2 x[8];
#pragma unroll
  for (int i = 0; i < 8; i++) x[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  float2 A = make_float2(a, a * 0.999f), B = make_float2(b, b * 1.001f);
  float m = 0.f;
  for (int it = 0; it < kIters; it++) {
    if (MODE == 0) {  // 16 independent scalar FFMA
#pragma unroll
      for (int i = 0; i < 8; i++) { x[i].x = fmaf(x[i].x, A.x, B.x); x[i].y = fmaf(x[i].y, A.y, B.y); }
    } else if (MODE == 1) {  // 8 independent FFMA2 (same flops)
#pragma unroll
      for (int i = 0; i < 8; i++) x[i] = __ffma2_rn(x[i], A, B);
    } else if (MODE == 2) {  // 8 FFMA2 + 8 FMNMX
#pragma unroll
      for (int i = 0; i < 8; i++) { x[i] = __ffma2_rn(x[i], A, B); x[i].x = fminf(x[i].x, 1e30f); }
    } else if (MODE == 3) {  // 16 FFMA + 8 FMNMX
#pragma unroll
      for (int i = 0; i < 8; i++) { x[i].x = fminf(fmaf(x[i].x, A.x, B.x), 1e30f); x[i].y = fmaf(x[i].y, A.y, B.y); }
    } else if (MODE == 4) {  // one dependent FFMA chain (latency)
#pragma unroll
      for (int i = 0; i < 16; i++) x[0].x = fmaf(x[0].x, A.x, B.x);
    } else if (MODE == 5) {  // one dependent FFMA2 chain (latency)
#pragma unroll
      for (int i = 0; i < 16; i++) x[0] = __ffma2_rn(x[0], A, B);
    } else if (MODE == 6) {  // FADD2 + FMUL2 mix
#pragma unroll
      for (int i = 0; i < 8; i++) { x[i] = __fadd2_rn(x[i], A); x[i] = __fmul2_rn(x[i], B); }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; i++) m += x[i].x + x[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = m;
}

template <int MODE>
void run(const char* name, double flops_per_iter, int threads) {
  float* out;
  cudaMalloc(&out, 148 * 448 * sizeof(float));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, threads>>>(out, 0.9999f, 1e-4f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int r = 0; r < 5; r++) k<MODE><<<148, threads>>>(out, 0.9999f, 1e-4f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
  double total = flops_per_iter * kIters * 148.0 * threads;
  double cyc = ms * 1e-3 * 1.965e9 / kIters;
  printf("%-34s threads %3d  %8.3f ms  %7.2f TFLOP/s  %6.2f cycles/iter\n", name, threads, ms, total / ms * 1e-9, cyc);
  cudaFree(out);
}

int main() {
  for (int threads : {448, 128}) {
    run<0>("16 FFMA", 32, threads);
    run<1>("8 FFMA2", 32, threads);
    run<2>("8 FFMA2 + 8 FMNMX", 32, threads);
    run<3>("16 FFMA + 8 FMNMX", 32, threads);
    run<6>("8 FADD2 + 8 FMUL2", 32, threads);
  }
  run<4>("dependent FFMA x16 (1 warp)", 32, 32);
  run<5>("dependent FFMA2 x16 (1 warp)", 64, 32);
  return 0;
}
