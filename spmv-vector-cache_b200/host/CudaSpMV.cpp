// CudaSpMV.cpp -- see CudaSpMV.h.  Host C++ over the C ABI (include/b200spmv.h); g++ only.
#include "CudaSpMV.h"

#include <string.h>

#include <chrono>
#include <iostream>

#include "../../include/b200spmv.h"

using namespace std;

#ifdef SPMV_DATA_U64
static const int kDtype = B200SPMV_U64;
#else
static const int kDtype = B200SPMV_F64;
#endif

CudaSpMV::CudaSpMV(SparseMatrix *A, SpMVData *x, SpMVData *y, const CudaSpMVOptions &opt)
    : SpMV(A, x, y), m_ctx(0), m_opt(opt), m_ready(false), m_allocX(false), m_allocY(false), m_golden(-1), m_lastRc(0),
      m_cmsTime(0), m_maxAliveTime(0), m_maxColSpanTime(0), m_maxAlive(0), m_maxColSpan(0), m_vcDone(false)
{
    memset(m_vc, 0, sizeof m_vc);
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

static unsigned usSince(const chrono::steady_clock::time_point &t0)
{
    return (unsigned)chrono::duration_cast<chrono::microseconds>(chrono::steady_clock::now() - t0).count();
}

bool CudaSpMV::measurePreprocessingTimes()
{
    if (!ok()) return false;
    const unsigned nz = m_A->getNz();
    // a private copy of the row-index stream with any marker bits removed (clearRowMarkings, SparseMatrix.cpp:121-125)
    vector<SpMVIndex> inds(m_A->getInds(), m_A->getInds() + nz);
    for (unsigned i = 0; i < nz; i++) inds[i] &= 0x3FFFFFFFu;
    chrono::steady_clock::time_point t0 = chrono::steady_clock::now();
    m_lastRc = b200spmv_max_col_span(m_ctx, m_A->getCols(), nz, m_A->getIndPtrs(), &inds[0], &m_maxColSpan);
    m_maxColSpanTime = usSince(t0);
    if (m_lastRc != 0) return false;
    t0 = chrono::steady_clock::now();
    m_lastRc = b200spmv_max_alive(m_ctx, m_A->getRows(), nz, &inds[0], &m_maxAlive);  // leaves bits 31 / 30 set
    m_maxAliveTime = usSince(t0);
    if (m_lastRc != 0) return false;
    for (unsigned i = 0; i < nz; i++) inds[i] &= 0x3FFFFFFFu;
    t0 = chrono::steady_clock::now();
    m_lastRc = b200spmv_mark_row_starts(m_ctx, m_A->getRows(), nz, &inds[0], 0, 31);
    m_cmsTime = usSince(t0);
    return m_lastRc == 0;
}

bool CudaSpMV::runVectorCacheModel()
{
    if (!ok() || m_opt.vcDepth <= 0) return false;
    int lg = 0;
    while ((1 << lg) < m_opt.vcDepth) lg++;
    const unsigned nz = m_A->getNz();
    const bool cms = m_opt.indexMask != 0xFFFFFFFFu;
    vector<SpMVIndex> inds(m_A->getInds(), m_A->getInds() + nz);
    // with cold-miss skip the model reads bit 31 as "first touch of this row" (NoWMVectorCache.scala:240-252).  The
    // marks are (re)made on the device on this private copy -- the same bits SparseMatrix::markRowStarts sets
    // (main.cpp:228-229), whether or not the caller's matrix carries them
    for (unsigned i = 0; i < nz; i++) inds[i] &= 0x3FFFFFFFu;
    if (cms && (m_lastRc = b200spmv_mark_row_starts(m_ctx, m_A->getRows(), nz, &inds[0], 0, 31)) != 0) return false;
    m_lastRc = b200spmv_vector_cache_model(m_ctx, nz, &inds[0], lg, cms ? 1 : 0, m_vc);
    m_vcDone = m_lastRc == 0;
    return m_vcDone;
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
    // ... the preprocessing keys of SoftwareSpMV::statKeys (SoftwareSpMV.cpp:96-108; zero until
    // measurePreprocessingTimes has run) ...
    static const char *k[] = {"diffFromGolden", "rows", "cols", "nz", "spmvtime", "cmstime", "maxAliveTime", "maxColSpanTime",
                              "maxAlive", "maxColSpan", "xfertime", "setuptime", "gpus", "slabs", "l2PersistKB", "mode", "cudaError"};
    vector<string> keys(k, k + sizeof(k) / sizeof(k[0]));
    if (m_opt.vcDepth > 0) {
        // ... and for a "vcmodel" conf the NewCache keys a functional model can know (HardwareSpMVNewCache.cpp:189-204;
        // the cycle and stall counters of the RTL are not modelled) followed by the model's own counters
        static const char *v[] = {"readMisses", "ocmDepth", "cms", "reads", "hits", "conflictEvictions", "coldMissSkips",
                                  "ddrReads", "flushWrites"};
        keys.insert(keys.end(), v, v + sizeof(v) / sizeof(v[0]));
    }
    return keys;
}

unsigned int CudaSpMV::statInt(std::string name)
{
    if (name == "diffFromGolden") return m_golden < 0 ? 1u : (unsigned)m_golden;  // 1 until compared, like m_diffFromGolden
    if (name == "rows") return m_A->getRows();
    if (name == "cols") return m_A->getCols();
    if (name == "nz") return m_A->getNz();
    // microseconds on the compute stream between the first and the last kernel of the step; inside b200spmv_exec that
    // stream also waits for the slices of x to land, so this includes upload waits (execDeviceOnly: kernels only)
    if (name == "spmvtime") return (unsigned)(stat64("spmvtime_ns") / 1000);
    if (name == "cmstime") return m_cmsTime;
    if (name == "maxAliveTime") return m_maxAliveTime;
    if (name == "maxColSpanTime") return m_maxColSpanTime;
    if (name == "maxAlive") return m_maxAlive;
    if (name == "maxColSpan") return m_maxColSpan;
    if (m_opt.vcDepth > 0) {
        static const char *v[] = {"reads", "readMisses", "conflictEvictions", "coldMissSkips", "ddrReads", "flushWrites", "hits"};
        for (int i = 0; i < 7; i++) if (name == v[i]) return (unsigned)m_vc[i];
        if (name == "ocmDepth") return (unsigned)m_opt.vcDepth;
        if (name == "cms") return m_opt.indexMask != 0xFFFFFFFFu ? 1u : 0u;
    }
    if (name == "xfertime") return (unsigned)(stat64("xfertime_ns") / 1000);
    if (name == "setuptime") return (unsigned)(stat64("setuptime_ns") / 1000);
    if (name == "gpus") return (unsigned)stat64("gpus");
    if (name == "slabs") return (unsigned)stat64("slabs");
    if (name == "l2PersistKB") return m_opt.persistX ? (unsigned)(stat64("l2_persist_bytes") / 1024) : 0u;
    if (name == "mode") return (unsigned)stat64("mode");
    if (name == "cudaError") return m_ctx ? (unsigned)stat64("cuda_error") : (unsigned)(-m_lastRc);
    return 0;
}
