/* Host stand-in for the Zynq BSP header that /root/reference/software/timer.h:1 includes.
 * Only the type name is needed to compile SoftwareSpMV.cpp off-board. */
#ifndef ORACLE_STUB_XSCUTIMER_H
#define ORACLE_STUB_XSCUTIMER_H
typedef struct { int unused; } XScuTimer;
#endif
