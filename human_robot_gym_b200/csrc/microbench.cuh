// Register-resident FP64 peak of the device this library runs on ("measure the FP64 FMA peak with a
// microbenchmark (not in MEASURED_PEAKS.json)"). The shield is compiled with -fmad=false for bit parity with the x86-64
// reference build, so its arithmetic is DADD + DMUL, not DFMA: both peaks are measured, and bench.py quotes each kernel's
// FP64 instruction rate against the add/mul one.
#pragma once
#include <cuda_runtime.h>

namespace hrg {

// MODE 0: DFMA chains (2 flop / instruction); MODE 1: DMUL + DADD pairs (1 flop / instruction, what -fmad=false emits);
// MODE 2: DADD only. 8 independent accumulators per thread hide the pipe latency; nothing touches memory inside the loop.
template <int MODE>
__global__ void __launch_bounds__(256) k_fp64_peak(double* __restrict__ out, int iters, double m, double c) {
  double a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = 1.0 + 1e-3 * (threadIdx.x + k);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (MODE == 0) a[k] = __fma_rn(a[k], m, c);
        if (MODE == 1) a[k] = __dadd_rn(__dmul_rn(a[k], m), c);
        if (MODE == 2) a[k] = __dadd_rn(a[k], c);
      }
    }
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Dependent-issue latency: ONE warp, ONE chain per thread, clock64 around `iters` x 16 dependent operations.
// MODE 0 DADD, 1 DMUL, 2 DFMA, 3 IEEE division, 4 sqrt, 5 DADD + two-instruction FP64 select (the sampler's recurrence step).
template <int MODE>
__global__ void k_fp64_latency(double* __restrict__ out, long long* __restrict__ cycles, int iters, double m, double c) {
  double a = 1.0 + 1e-3 * threadIdx.x;
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      if (MODE == 0) a = __dadd_rn(a, c);
      if (MODE == 1) a = __dmul_rn(a, m);
      if (MODE == 2) a = __fma_rn(a, m, c);
      if (MODE == 3) a = __ddiv_rn(m, a) + 1.0;
      if (MODE == 4) a = __dsqrt_rn(a) + 1.0;
      if (MODE == 5) {
        const double b = __dadd_rn(a, c);
        a = (b < 1e300) ? b : 0.0;
      }
    }
  }
  const long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) *cycles = t1 - t0;
}

}  // namespace hrg
