// ref_tu.cpp -- builds the UNMODIFIED reference CPU path into oracle/_ref/libref_{f64,u64}.so.
//
// TEST INFRASTRUCTURE ONLY.  No reference source is copied: the three .cpp files are
// #included from where they lie (-I/root/reference/software) and wrapped in a C ABI
// so tests/ and bench.py can call them through ctypes.
//
// The reference hard-wires `typedef double SpMVData;` (software/SparseMatrix.h:6).  The
// uint64 build (-DREF_U64) re-types it without touching the files: all system headers
// the reference uses are included first, then `double` is redefined for the span of the
// reference sources only, so SoftwareSpMV.cpp:62 becomes wrapping u64 mul+add.
#include <assert.h>
#include <stdint.h>
#include <string.h>
#include <time.h>
#include <iostream>
#include <string>
#include <vector>

#include <xscutimer.h>
#include "xparameters.h"

#ifdef REF_U64
#define double unsigned long long
#define REF_FN(name) ref_u64_##name
#else
#define REF_FN(name) ref_f64_##name
#endif

#include "SpMV.cpp"
#include "SparseMatrix.cpp"
#include "SoftwareSpMV.cpp"

#ifdef REF_U64
#undef double
#endif

// ---- host implementation of software/timer.h:6-10 over clock_gettime ----
static struct timespec g_t0, g_t1;
void TimerSetup() {}
void TimerStart() { clock_gettime(CLOCK_MONOTONIC, &g_t0); }
void TimerStop() { clock_gettime(CLOCK_MONOTONIC, &g_t1); }
static double timer_secs()
{
    return (double)(g_t1.tv_sec - g_t0.tv_sec) + 1e-9 * (double)(g_t1.tv_nsec - g_t0.tv_nsec);
}
// the board returns ticks of the 333.333 MHz SCU timer (software/timer.c:24-31)
unsigned int TimerRead() { return (unsigned int)(timer_secs() * 333333333.0); }
float TimerReadSecs() { return (float)timer_secs(); }

// ---- loader: SparseMatrix members are protected (SparseMatrix.h:60-69) ----
struct HostSparseMatrix : public SparseMatrix {
    HostSparseMatrix(unsigned rows, unsigned cols, unsigned nz, SpMVIndex *ptr, SpMVIndex *ind,
                     SpMVData *data)
    {
        m_rows = rows; m_cols = cols; m_nz = nz;
        m_indPtrs = ptr; m_inds = ind; m_nzData = data;
    }
};

extern "C" {

// SoftwareSpMV::exec() on caller-owned x and y (y += A*x).  Returns loop seconds.
__attribute__((visibility("default")))
double REF_FN(spmv)(unsigned rows, unsigned cols, unsigned nz, unsigned *colptr, unsigned *rowind,
                    void *vals, void *x, void *y)
{
    HostSparseMatrix A(rows, cols, nz, colptr, rowind, (SpMVData *)vals);
    SoftwareSpMV s(&A, (SpMVData *)x, (SpMVData *)y);
    bool ok = s.exec();
    assert(ok); (void)ok;
    return timer_secs();
}

// SoftwareSpMV with its default vectors (x = 1, y = 0; SoftwareSpMV.cpp:23-39).
__attribute__((visibility("default")))
void REF_FN(spmv_default)(unsigned rows, unsigned cols, unsigned nz, unsigned *colptr,
                          unsigned *rowind, void *vals, void *y_out)
{
    HostSparseMatrix A(rows, cols, nz, colptr, rowind, (SpMVData *)vals);
    SoftwareSpMV s(&A);
    s.exec();
    memcpy(y_out, s.getY(), (size_t)rows * sizeof(SpMVData));
}

__attribute__((visibility("default")))
void REF_FN(mark_row_starts)(unsigned rows, unsigned nz, unsigned *inds, int reverse, int shift)
{
    HostSparseMatrix A(rows, 1, nz, 0, inds, 0);
    A.markRowStarts(reverse != 0, shift);
}

__attribute__((visibility("default")))
unsigned REF_FN(max_alive)(unsigned rows, unsigned nz, unsigned *inds)
{
    HostSparseMatrix A(rows, 1, nz, 0, inds, 0);
    return A.maxAlive();
}

__attribute__((visibility("default")))
unsigned REF_FN(max_col_span)(unsigned cols, unsigned nz, unsigned *colptr, unsigned *inds)
{
    HostSparseMatrix A(1, cols, nz, colptr, inds, 0);
    return A.maxColSpan();
}

__attribute__((visibility("default")))
void REF_FN(clear_row_markings)(unsigned nz, unsigned *inds, unsigned mask)
{
    HostSparseMatrix A(1, 1, nz, 0, inds, 0);
    A.clearRowMarkings(mask);
}

__attribute__((visibility("default")))
unsigned REF_FN(sizeof_data)() { return (unsigned)sizeof(SpMVData); }

} // extern "C"
