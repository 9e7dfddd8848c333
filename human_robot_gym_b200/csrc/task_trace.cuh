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
// how often the exact QR restatement still runs: [0] k_replan_deep phase B, [1] inline (smallest_positive_real_root close
// call), [2] candidates settled by ltp_cand_rejected_from_rest
__device__ unsigned long long g_qr_count[4];
// whole-kernel spans of the other kernels of a tick: [id][0] = earliest block start, [id][1] = latest block end
__device__ unsigned long long g_kernel_span[16][2];
struct KernelSpan {
  int id;
  __device__ __forceinline__ explicit KernelSpan(int i) : id(i) {
    if (threadIdx.x == 0) atomicMin(&g_kernel_span[id][0], trace_now());
  }
  __device__ __forceinline__ ~KernelSpan() {
    if (threadIdx.x == 0) atomicMax(&g_kernel_span[id][1], trace_now());
  }
};
}  // namespace hrg
#define HRG_TRACE_SPAN(ID) hrg::KernelSpan trace_span_(ID)
#define HRG_TRACE_COUNT(I) atomicAdd(&hrg::g_qr_count[I], 1ull)
#define HRG_TRACE_T0() const unsigned long long trace_t0 = hrg::trace_now()
#define HRG_TRACE_TASK(K, TASK, EXTRA) \
  do {                                 \
    if ((threadIdx.x & 31) == 0) hrg::trace_task((K), (TASK), trace_t0, (EXTRA)); \
  } while (0)
#else
#define HRG_TRACE_COUNT(I) \
  do {                     \
  } while (0)
#define HRG_TRACE_SPAN(ID) \
  do {                     \
  } while (0)
#define HRG_TRACE_T0() \
  do {                 \
  } while (0)
#define HRG_TRACE_TASK(K, TASK, EXTRA) \
  do {                                 \
  } while (0)
#endif
