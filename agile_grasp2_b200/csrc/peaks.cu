// Same-run FP32 / FP64 FMA peak of the device the path runs on ("FP64 peak from a same-run FMA
// microbenchmark").  The normals and hand-sweep kernels are bound by the SM's issue rate, not by HBM; their
// roofline fraction in bench.py is algorithmic flop / time against THESE measured numbers.
// Eight independent FMA chains per thread, 256 threads, eight CTAs per SM: enough parallelism to fill the pipe.
#include "dev_common.cuh"

namespace ag2 {
namespace {

template <typename T>
__global__ void __launch_bounds__(256) k_fma_peak(T* out, int iters, T x, T y) {
  T a0 = (T)threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
#pragma unroll 1
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      a0 = fma(a0, x, y), a1 = fma(a1, x, y), a2 = fma(a2, x, y), a3 = fma(a3, x, y);
      a4 = fma(a4, x, y), a5 = fma(a5, x, y), a6 = fma(a6, x, y), a7 = fma(a7, x, y);
    }
  }
  const T s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == (T)-1) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // never true: keeps the chains alive
}

template <typename T>
int measure(ag2_ctx* c, int iters, double* tflops) {
  cudaStream_t st = c->stream;
  const int blocks = c->sm_count * 8;
  DevBuf<T> sink;
  AG2_CUDA(c, sink.reserve((size_t)blocks * 256));
  cudaEvent_t e0, e1;
  AG2_CUDA(c, cudaEventCreate(&e0));
  AG2_CUDA(c, cudaEventCreate(&e1));
  double best = 0.0;
  for (int rep = 0; rep < 4; rep++) {  // the first run also warms the clocks up
    cudaEventRecord(e0, st);
    k_fma_peak<T><<<blocks, 256, 0, st>>>(sink.p, iters, (T)0.999, (T)0.001);
    cudaEventRecord(e1, st);
    AG2_LAUNCH_CHECK(c);
    AG2_CUDA(c, cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flop = 2.0 * 64.0 * (double)iters * 256.0 * (double)blocks;
    if (rep > 0 && ms > 0.f) best = std::max(best, flop / ((double)ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *tflops = best;
  return AG2_OK;
}

}  // namespace
}  // namespace ag2

extern "C" int ag2_measure_fp_peaks(ag2_ctx* c, double* fp32_tflops, double* fp64_tflops) {
  using namespace ag2;
  if (!c || !fp32_tflops || !fp64_tflops) return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  AG2_TRY(measure<float>(c, 4096, fp32_tflops));
  AG2_TRY(measure<double>(c, 2048, fp64_tflops));
  return AG2_OK;
}
