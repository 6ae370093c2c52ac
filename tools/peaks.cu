// peaks.cu — measured denominators that MEASURED_PEAKS.json does not hold: FP32 FMA throughput and
// L2→SM read bandwidth of this GPU.  Built by __graft_entry__.build() into tools/peaks; bench.py runs it once and
// puts the numbers beside the roofline fractions.  Prints one JSON line.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__global__ void __launch_bounds__(256) k_fma(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void __launch_bounds__(256) k_read(const float4* __restrict__ in, size_t n, int reps, float* out) {
  float acc = 0.f;
  for (int r = 0; r < reps; r++)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
      float4 v = __ldcg(in + i);       // L2 (bypass L1)
      acc += v.x + v.y + v.z + v.w;
    }
  if (acc == 12345.678f) out[0] = acc;
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) { float ms; cudaEventElapsedTime(&ms, a, b); return ms; }

int main() {
  int dev = 0, sms = 0, clock_khz = 0;
  if (cudaSetDevice(dev) != cudaSuccess) { printf("{\"error\": \"no CUDA device\"}\n"); return 1; }
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, dev);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float* out; cudaMalloc(&out, (size_t)sms * 16 * 256 * sizeof(float));
  const int iters = 4096;
  double best_tf = 0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    k_fma<<<sms * 16, 256>>>(out, iters, 1.0001f, 0.5f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    double flops = (double)sms * 16 * 256 * iters * 64 * 2;
    double tf = flops / (time_ms(e0, e1) * 1e-3) / 1e12;
    if (rep > 0 && tf > best_tf) best_tf = tf;
  }
  const size_t bytes = 48u << 20;                 // 48 MiB: resident in the 126 MB L2
  float4* buf; cudaMalloc(&buf, bytes); cudaMemset(buf, 0, bytes);
  double best_l2 = 0;
  for (int rep = 0; rep < 5; rep++) {
    const int reps = 20;
    cudaEventRecord(e0);
    k_read<<<sms * 8, 256>>>(buf, bytes / 16, reps, out);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    double gbs = (double)bytes * reps / (time_ms(e0, e1) * 1e-3) / 1e9;
    if (rep > 0 && gbs > best_l2) best_l2 = gbs;
  }
  cudaError_t err = cudaGetLastError();
  printf("{\"fp32_fma_tflops\": %.2f, \"l2_read_gbs\": %.1f, \"sms\": %d, \"sm_clock_mhz_attr\": %.0f, \"how\": \"8 independent FFMA chains/thread, 16x256 threads/SM, best of 4; __ldcg float4 reads of a 48 MiB L2-resident buffer x20, best of 4\", \"cuda_error\": \"%s\"}\n",
         best_tf, best_l2, sms, clock_khz / 1000.0, cudaGetErrorString(err));
  return 0;
}
