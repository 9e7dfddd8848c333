// Profiling aid, compiled only with -DHRG_TASK_TRACE (profiles/build_trace.sh -> profiles/_trace/libhrg_shield_trace.so;
// never part of libhrg_shield.so): per-task start / end times (%globaltimer, ns) and SM id of the replan-chain kernels, so
// that the tail of a kernel - the one warp everything waits for - can be told from its average.
#pragma once
#if defined(HRG_TASK_TRACE)
namespace hrg {
constexpr int kTraceTasks = 1 << 15;
__device__ unsigned long long g_task_trace[4][kTraceTasks][4];
__device__ __forceinline__ unsigned long long trace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned trace_smid() {
  unsigned s;
  asm volatile("mov.u32 %0, %smid;" : "=r"(s));
  return s;
}
__device__ __forceinline__ void trace_task(int kernel, int task, unsigned long long t0, unsigned long long extra) {
  if (task < 0 || task >= kTraceTasks) return;
  unsigned long long* r = g_task_trace[kernel][task];
  r[0] = t0;
  r[1] = trace_now();
  r[2] = trace_smid();
  r[3] = extra;
}
}  // namespace hrg
#define HRG_TRACE_T0() const unsigned long long trace_t0 = hrg::trace_now()
#define HRG_TRACE_TASK(K, TASK, EXTRA) \
  do {                                 \
    if ((threadIdx.x & 31) == 0) hrg::trace_task((K), (TASK), trace_t0, (EXTRA)); \
  } while (0)
#else
#define HRG_TRACE_T0() \
  do {                 \
  } while (0)
#define HRG_TRACE_TASK(K, TASK, EXTRA) \
  do {                                 \
  } while (0)
#endif
