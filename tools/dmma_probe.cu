// Register-resident DMMA microbenchmark: what does mma.sync.aligned.m8n8k4.f64 sustain on B200 when
// nothing but the tensor pipe is in the way?  Each warp keeps NACC independent 8x8 accumulators (2 doubles per lane each)
// and issues ITER rounds of NACC dependent-free DMMAs; operands live in registers.  Compared with a plain DFMA loop of the
// same flop count.  Build + run:  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/dmma tools/dmma_probe.cu && /tmp/dmma
// ncu:  ncu --metrics sm__inst_executed_pipe_tensor_op_dmma.sum,sm__pipe_tensor_op_dmma_cycles_active.avg.pct_of_peak_sustained_active /tmp/dmma
#include <cstdio>
#include <cuda_runtime.h>

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma(int iters, double* out, double seed) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = seed * i; c[i][1] = seed + i; }
  double a = seed + threadIdx.x * 1e-3, b = seed - threadIdx.x * 1e-3;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dfma(int iters, double* out, double seed) {
  double c[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) c[i] = seed * i;
  double a = seed + threadIdx.x * 1e-3, b = 1.0 - 1e-9 * threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i] = fma(c[i], b, a);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f();  // warm-up
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  f();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * 256 * sms * 8);
  const int iters = 20000;
  printf("{\"sms\": %d", sms);
  {
    const int grid = sms * 4;  // 4 CTAs x 8 warps per SM
    float ms = time_ms([&] { k_dmma<8><<<grid, 256>>>(iters, out, 1.0); });
    const double flop = 2.0 * 8 * 8 * 4 * 8.0 * iters * (double)grid * 8;  // 512 flop per DMMA, NACC per round, 8 warps per CTA
    printf(", \"dmma_m8n8k4_tflops\": %.2f, \"dmma_ms\": %.3f", flop / ms / 1e9, ms);
    ms = time_ms([&] { k_dmma<8><<<sms, 32>>>(iters, out, 1.0); });  // one warp per SM: per-warp issue rate
    printf(", \"dmma_one_warp_per_sm_tflops\": %.2f", 2.0 * 256 * 8.0 * iters * sms / ms / 1e9);
  }
  {
    const int grid = sms * 4;
    float ms = time_ms([&] { k_dfma<16><<<grid, 256>>>(iters, out, 1.0); });
    const double flop = 2.0 * 16.0 * iters * (double)grid * 256;
    printf(", \"dfma_tflops\": %.2f, \"dfma_ms\": %.3f", flop / ms / 1e9, ms);
  }
  printf("}\n");
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { fprintf(stderr, "%s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
