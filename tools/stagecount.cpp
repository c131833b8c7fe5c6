// Development tool (host only, never shipped): the hostemu harness with the PTB_STAT hooks of traverse.h counting the
// stages rays run through.  Single-threaded use (emu_trace with threads = 1).
#include <cstdint>
uint64_t g_stat[8];
#define PTB_STAT(stage) (++g_stat[stage])
#include "../tests/hostemu/hostemu.cpp"
extern "C" void emu_stat_read(uint64_t* out8, int reset)
{
    for (int i = 0; i < 8; ++i) { out8[i] = g_stat[i]; if (reset) g_stat[i] = 0; }
}
