// fp64_peak.cu -- measured FP64 peaks of the device the build kernels run on: DFMA (vector pipe), DMMA m8n8k4
// (tensor pipe) and both issued together.  The roofline of the fused build is FP64 work, not HBM,
// and MEASURED_PEAKS.json has no FP64 entry, so this writes one (profiles/rNN_fp64_peak.json).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/fp64_peak tools/fp64_peak.cu && /tmp/fp64_peak
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// mode 0: NF independent DFMA chains per thread; mode 1: NM independent DMMA accumulators per warp; mode 2: both interleaved
template <int MODE>
__global__ void __launch_bounds__(256) peak_kernel(double* out, int iters, double a, double b) {
  double f[8], c[8][2];
#pragma unroll
  for (int i = 0; i < 8; i++) { f[i] = threadIdx.x + i; c[i][0] = i; c[i][1] = -i; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (MODE == 0 || MODE == 2) f[i] = fma(f[i], a, b);
        if (MODE == 1 || MODE == 2) dmma884(c[i][0], c[i][1], a, b);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += f[i] + c[i][0] + c[i][1];
  if (s == 123.456) out[0] = s;
}

template <int MODE>
static int run(const char* name, int sms, double* d_out, double flop_per_thread_iter_fma, double flop_per_warp_iter_mma) {
  const int iters = 4096, ctas = sms * 8, threads = 256;
  peak_kernel<MODE><<<ctas, threads>>>(d_out, 64, 1.0000001, 1e-9);
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    CK(cudaEventRecord(e0));
    peak_kernel<MODE><<<ctas, threads>>>(d_out, iters, 1.0000001, 1e-9);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  const double nthreads = (double)ctas * threads;
  const double fl_fma = (MODE == 0 || MODE == 2) ? nthreads * iters * flop_per_thread_iter_fma : 0.0;
  const double fl_mma = (MODE == 1 || MODE == 2) ? nthreads / 32 * iters * flop_per_warp_iter_mma : 0.0;
  std::printf("  \"%s\": {\"ms\": %.4f, \"dfma_tflops\": %.3f, \"dmma_tflops\": %.3f, \"total_tflops\": %.3f},\n", name, best, fl_fma / best * 1e-9,
              fl_mma / best * 1e-9, (fl_fma + fl_mma) / best * 1e-9);
  return 0;
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  double* d_out;
  CK(cudaMalloc(&d_out, 64));
  int clk = 0; CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
  std::printf("{\n  \"device\": \"%s\", \"sms\": %d, \"sm_clock_khz\": %d,\n", p.name, p.multiProcessorCount, clk);
  // per inner iteration: 4 x 8 FMA per thread (2 flop each); 4 x 8 DMMA per warp (8*8*4*2 flop each)
  if (run<0>("dfma_only", p.multiProcessorCount, d_out, 4 * 8 * 2.0, 4 * 8 * 512.0)) return 1;
  if (run<1>("dmma_only", p.multiProcessorCount, d_out, 4 * 8 * 2.0, 4 * 8 * 512.0)) return 1;
  if (run<2>("dfma_and_dmma", p.multiProcessorCount, d_out, 4 * 8 * 2.0, 4 * 8 * 512.0)) return 1;
  std::printf("  \"how\": \"8 CTAs x 256 threads per SM, 8 independent chains per thread / accumulators per warp, best of 5, CUDA events\"\n}\n");
  return 0;
}
