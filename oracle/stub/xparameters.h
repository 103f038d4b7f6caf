/* Host stand-in for the Zynq BSP header that /root/reference/software/timer.h:2 includes. */
#ifndef ORACLE_STUB_XPARAMETERS_H
#define ORACLE_STUB_XPARAMETERS_H
#define XPAR_XSCUTIMER_0_DEVICE_ID 0
#endif
