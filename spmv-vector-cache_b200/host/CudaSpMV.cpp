// CudaSpMV.cpp -- see CudaSpMV.h.  Host C++ over the C ABI (include/b200spmv.h); g++ only.
#include "CudaSpMV.h"

#include <string.h>

#include <iostream>

#include "../../include/b200spmv.h"

using namespace std;

#ifdef SPMV_DATA_U64
static const int kDtype = B200SPMV_U64;
#else
static const int kDtype = B200SPMV_F64;
#endif

CudaSpMV::CudaSpMV(SparseMatrix *A, SpMVData *x, SpMVData *y, const CudaSpMVOptions &opt)
    : SpMV(A, x, y), m_ctx(0), m_opt(opt), m_ready(false), m_allocX(false), m_allocY(false), m_golden(-1), m_lastRc(0)
{
    // default vectors like SoftwareSpMV.cpp:23-39: x = 1, y = 0, owned by this object
    if (m_x == 0) {
        m_x = new SpMVData[A->getCols()];
        for (SpMVIndex i = 0; i < A->getCols(); i++) m_x[i] = (SpMVData)1;
        m_allocX = true;
    }
    if (m_y == 0) {
        m_y = new SpMVData[A->getRows()];
        for (SpMVIndex i = 0; i < A->getRows(); i++) m_y[i] = (SpMVData)0;
        m_allocY = true;
    }
    m_lastRc = b200spmv_create(&m_ctx, opt.device);
    if (m_lastRc != 0) {
        cerr << "CudaSpMV: " << b200spmv_last_error(0) << endl;  // message + unusable object, like SparseMatrix.cpp:33-38
        m_ctx = 0;
        return;
    }
    if (opt.slabs > 0) b200spmv_set_option(m_ctx, "slabs", opt.slabs);
    if (opt.requireAligned) b200spmv_set_option(m_ctx, "require_aligned", 1);
    // m_x / m_y live as long as this object (SpMV.h:31-33): page-lock them in place on first use
    b200spmv_set_option(m_ctx, "pin_host", 1);
    m_lastRc = b200spmv_setup_csc(m_ctx, A->getRows(), A->getCols(), A->getNz(), A->getIndPtrs(), A->getInds(),
                                  A->getNzData(), kDtype, opt.indexMask);
    if (m_lastRc != 0) {
        cerr << "CudaSpMV: setup failed: " << b200spmv_last_error(m_ctx) << endl;
        return;
    }
    m_ready = true;
}

CudaSpMV::~CudaSpMV()
{
    if (m_ctx) b200spmv_destroy(m_ctx);
    if (m_allocX) delete[] m_x;
    if (m_allocY) delete[] m_y;
}

static int execFlags(const CudaSpMVOptions &o)
{
    return (o.accumulate ? B200SPMV_ACCUMULATE : 0) | (o.strictOrder ? B200SPMV_STRICT_ORDER : 0) |
           (o.persistX ? 0 : B200SPMV_NO_PERSIST);
}

bool CudaSpMV::exec()
{
    if (!ok()) return false;
    m_golden = -1;  // HardwareSpMV::resetAccelerator invalidates the last comparison (HardwareSpMV.cpp:30-34)
    m_lastRc = b200spmv_exec(m_ctx, m_x, m_y, execFlags(m_opt));
    if (m_lastRc != 0) cerr << "CudaSpMV: exec failed: " << b200spmv_last_error(m_ctx) << endl;
    return m_lastRc == 0;
}

bool CudaSpMV::execDeviceOnly(int iters, float *msEach)
{
    if (!ok()) return false;
    m_lastRc = b200spmv_exec_device(m_ctx, iters, execFlags(m_opt), msEach);
    return m_lastRc == 0;
}

void CudaSpMV::compareGolden(SpMVData *golden)
{
    m_golden = memcmp(golden, m_y, (size_t)m_A->getRows() * sizeof(SpMVData)) != 0 ? 1 : 0;
}

uint64_t CudaSpMV::stat64(const std::string &name)
{
    uint64_t v = 0;
    if (m_ctx && b200spmv_stat_u64(m_ctx, name.c_str(), &v) == 0) return v;
    return 0;
}

std::string CudaSpMV::lastError() const { return b200spmv_last_error(m_ctx); }

std::vector<std::string> CudaSpMV::statKeys()
{
    // the HardwareSpMV vocabulary (HardwareSpMV.cpp:54-61) first, then SoftwareSpMV's spmvtime, then ours
    static const char *k[] = {"diffFromGolden", "rows", "cols", "nz", "spmvtime", "xfertime", "setuptime",
                              "gpus", "slabs", "l2PersistKB", "mode", "cudaError"};
    return vector<string>(k, k + sizeof(k) / sizeof(k[0]));
}

unsigned int CudaSpMV::statInt(std::string name)
{
    if (name == "diffFromGolden") return m_golden < 0 ? 1u : (unsigned)m_golden;  // 1 until compared, like m_diffFromGolden
    if (name == "rows") return m_A->getRows();
    if (name == "cols") return m_A->getCols();
    if (name == "nz") return m_A->getNz();
    if (name == "spmvtime") return (unsigned)(stat64("spmvtime_ns") / 1000);   // microseconds, kernels only
    if (name == "xfertime") return (unsigned)(stat64("xfertime_ns") / 1000);
    if (name == "setuptime") return (unsigned)(stat64("setuptime_ns") / 1000);
    if (name == "gpus") return (unsigned)stat64("gpus");
    if (name == "slabs") return (unsigned)stat64("slabs");
    if (name == "l2PersistKB") return m_opt.persistX ? (unsigned)(stat64("l2_persist_bytes") / 1024) : 0u;
    if (name == "mode") return (unsigned)stat64("mode");
    if (name == "cudaError") return m_ctx ? (unsigned)stat64("cuda_error") : (unsigned)(-m_lastRc);
    return 0;
}
