// CudaSpMV.h -- the new SpMV variant: a B200 behind the reference's operator interface.
//
// Selectable next to the FPGA variants of software/HWSpMVFactory.cpp:20-38 (see SpMVFactory.h); same
// setup / exec() / result-vector contract as software/HardwareSpMVNewCache.cpp:78-88: the constructor
// is "setupRegs + init" (upload the CSC arrays, transpose to CSR on the device, once), exec() is one
// blocking y += A*x whose result lands in the caller's y.  Compiled by g++: no CUDA headers, only the
// C ABI of include/b200spmv.h.  There is no CPU fallback: without a GPU the constructor reports the
// failure (ok() == false, exec() returns false, statInt("cudaError") != 0).
#ifndef B200_CUDASPMV_H_
#define B200_CUDASPMV_H_

#include <stdint.h>

#include <string>
#include <vector>

#include "SpMVInterface.h"

struct b200spmv_ctx;

struct CudaSpMVOptions {
    int device;          // CUDA device ordinal
    bool strictOrder;    // fp64 results memcmp-equal to SoftwareSpMV (row-sequential kernel)
    bool persistX;       // pin the gathered vector in L2 (the "vector cache"); false = A/B switch
    bool accumulate;     // y += A*x (SoftwareSpMV.cpp:62); false: y = A*x (what the FPGA variants compute)
    int slabs;           // column slabs, 0 = auto
    unsigned indexMask;  // 0x3FFFFFFF when the matrix went through SparseMatrix::markRowStarts
    bool requireAligned; // enforce the 64-byte rule of HardwareSpMV.cpp:6-16
    bool prep;           // conf token "prep": spmvbench calls measurePreprocessingTimes() after exec, like benchmarkSW
    int vcDepth;         // conf token "vcmodel<depth>": lines of the modelled vector cache (power of two), 0 = no model
    CudaSpMVOptions()
        : device(0), strictOrder(false), persistX(true), accumulate(true), slabs(0), indexMask(0xFFFFFFFFu),
          requireAligned(false), prep(false), vcDepth(0) {}
};

class CudaSpMV : public SpMV {
public:
    // x == 0 / y == 0: allocated and filled with ones / zeros like SoftwareSpMV.cpp:23-39
    CudaSpMV(SparseMatrix *A, SpMVData *x = 0, SpMVData *y = 0, const CudaSpMVOptions &opt = CudaSpMVOptions());
    virtual ~CudaSpMV();

    virtual bool exec();                                  // true on success (SoftwareSpMV.cpp:69)
    virtual unsigned int statInt(std::string name);       // unknown key -> 0 (HardwareSpMV.cpp:41-61)
    virtual std::vector<std::string> statKeys();

    void compareGolden(SpMVData *golden);                 // HardwareSpMV.cpp:37-39: memcmp over rows*8 bytes
    // SoftwareSpMV::measurePreprocessingTimes (SoftwareSpMV.cpp:72-94) on the device: maxColSpan, maxAlive and the
    // cold-miss-skip marking of the row-index stream, each timed (microseconds, upload included); fills the
    // SoftwareSpMV stat keys cmstime, maxAliveTime, maxColSpanTime, maxAlive, maxColSpan.  The matrix is not
    // modified (the reference marks in place and clears again).
    bool measurePreprocessingTimes();
    // The FPGA's direct-mapped vector cache over y, modelled for this matrix (csrc/vcache.cu): fills the NewCache
    // keys it can know (HardwareSpMVNewCache.cpp:189-204: readMisses, ocmDepth, cms) and the model's own counters.
    // cms = the matrix was row-start marked (conf token "cms") -> first touches skip the DRAM read.
    bool runVectorCacheModel();
    const CudaSpMVOptions &options() const { return m_opt; }
    bool execDeviceOnly(int iters, float *msEach = 0);    // kernels only, x / y already in HBM
    uint64_t stat64(const std::string &name);             // 64-bit statistics (ns timers, nnz > 2^32)
    bool ok() const { return m_ctx != 0 && m_ready; }
    std::string lastError() const;

protected:
    b200spmv_ctx *m_ctx;
    CudaSpMVOptions m_opt;
    bool m_ready, m_allocX, m_allocY;
    int m_golden;  // -1 = not compared, else memcmp result != 0
    int m_lastRc;
    unsigned m_cmsTime, m_maxAliveTime, m_maxColSpanTime, m_maxAlive, m_maxColSpan;
    uint64_t m_vc[7];  // reads, readMisses, conflictEvictions, coldMissSkips, ddrReads, flushWrites, hits
    bool m_vcDone;
};

#endif
