// host_timer.cpp -- TEST INFRASTRUCTURE ONLY.  Host implementation of the reference's
// software/timer.h:6-10 (TimerSetup/Start/Stop/Read/ReadSecs) over clock_gettime, for programs
// that link the reference's own SoftwareSpMV.cpp off the Zynq (oracle/_ref/spmvbench_ref).
#include <time.h>
static struct timespec g_t0, g_t1;
void TimerSetup() {}
void TimerStart() { clock_gettime(CLOCK_MONOTONIC, &g_t0); }
void TimerStop() { clock_gettime(CLOCK_MONOTONIC, &g_t1); }
static double secs() { return (double)(g_t1.tv_sec - g_t0.tv_sec) + 1e-9 * (double)(g_t1.tv_nsec - g_t0.tv_nsec); }
// the board returns ticks of the 333.333 MHz SCU timer (software/timer.c:24-31)
unsigned int TimerRead() { return (unsigned int)(secs() * 333333333.0); }
float TimerReadSecs() { return (float)secs(); }
