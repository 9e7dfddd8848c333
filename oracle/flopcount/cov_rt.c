// TEST INFRASTRUCTURE (oracle/): basic-block execution counter behind `g++ -fsanitize-coverage=trace-pc`.
//
// oracle/Makefile's `flopcount` target compiles the UNMODIFIED reference sources a second time with that flag, so that
// every basic block starts with a call to __sanitizer_cov_trace_pc(); this file is the callee (compiled WITHOUT the flag).
// It counts executions per call site; oracle/flopcount/count_flops.py joins the counts with `objdump -d` of the same
// library and sums the FP64 arithmetic instructions of each block times the number of times it ran. That is the
// "instrumented count of the oracle" SURVEY.md §8(d) / BASELINE.md §3 ask for (F_step), taken from the reference's own
// compiled code instead of an estimate. Single-threaded by design (run the reference with threads = 1).
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdint.h>
#include <string.h>

#define TAB_BITS 18
#define N_PHASE 4
#define TAB_SIZE (1u << TAB_BITS)

typedef struct { uintptr_t pc; uint64_t n; } Slot;
static Slot g_tab[N_PHASE][TAB_SIZE];
static int g_on, g_phase;

void __sanitizer_cov_trace_pc(void) {
  if (!g_on) return;
  uintptr_t pc = (uintptr_t)__builtin_return_address(0);
  uint32_t h = (uint32_t)((pc * 0x9E3779B97F4A7C15ull) >> (64 - TAB_BITS));
  Slot* tab = g_tab[g_phase];
  while (tab[h].pc != pc) {
    if (!tab[h].pc) { tab[h].pc = pc; break; }
    h = (h + 1) & (TAB_SIZE - 1);
  }
  tab[h].n++;
}

void cov_enable(int on) { g_on = on; }
// Counts are kept per phase (which API call of the tick is running): 0 .. N_PHASE-1.
void cov_phase(int p) { g_phase = p < 0 ? 0 : (p >= N_PHASE ? N_PHASE - 1 : p); }
void cov_reset(void) { memset(g_tab, 0, sizeof g_tab); }

// Load address of this shared object: return addresses minus this are the addresses `objdump -d` prints.
uint64_t cov_base(void) {
  Dl_info info;
  if (!dladdr((void*)&cov_base, &info)) return 0;
  return (uint64_t)(uintptr_t)info.dli_fbase;
}

// (return address - base, count) pairs of one phase into out[2 * cap]; returns how many there are (cap = 0 to size).
long cov_dump(int phase, uint64_t* out, long cap) {
  uint64_t base = cov_base();
  const Slot* tab = g_tab[phase < 0 || phase >= N_PHASE ? 0 : phase];
  long k = 0;
  for (uint32_t i = 0; i < TAB_SIZE; ++i) {
    if (!tab[i].pc) continue;
    if (k < cap) { out[2 * k] = (uint64_t)tab[i].pc - base; out[2 * k + 1] = tab[i].n; }
    ++k;
  }
  return k;
}
