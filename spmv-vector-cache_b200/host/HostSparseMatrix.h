// HostSparseMatrix.h -- loads a matrix directory into a SparseMatrix on a 64-bit host.
//
// File-naming contract of software/main.cpp:26-37: <dir>/<name>/<name>-meta.bin, -indptr.bin,
// -inds.bin, -data.bin (+ golden.bin = A*1 for the fp64 directories, matrices/matrixutils.py:108-113).
// SparseMatrix::fromMemory (software/SparseMatrix.cpp:29-50) casts 32-bit base addresses to pointers
// and cannot be used off the Zynq; this subclass fills the protected members instead and keeps the
// reference's validation (zero rows / cols / nz rejected with the same message, :33-38) and its
// 64-byte alignment rule (software/HardwareSpMV.cpp:6-16, main.cpp malloc_aligned(64, ...)).
#ifndef B200_HOSTSPARSEMATRIX_H_
#define B200_HOSTSPARSEMATRIX_H_

#include <string>

#include "SpMVInterface.h"

void *b200_malloc_aligned(size_t alignment, size_t bytes);  // main.cpp:113-116 malloc_aligned
void b200_free_aligned(void *p);

class HostSparseMatrix : public SparseMatrix {
public:
    // returns 0 (after a message on cerr) when a file is missing or the metadata is invalid
    static HostSparseMatrix *fromDirectory(const std::string &root, const std::string &name);
    // adopt caller-built CSC arrays (copied into aligned storage)
    static HostSparseMatrix *fromArrays(unsigned rows, unsigned cols, unsigned nz, const SpMVIndex *colptr,
                                        const SpMVIndex *rowind, const SpMVData *vals);
    virtual ~HostSparseMatrix();
    // golden.bin next to the matrix files, 0 when absent; rows elements, owned by the matrix
    SpMVData *golden() const { return m_golden; }

protected:
    HostSparseMatrix() : m_golden(0) {}
    SpMVData *m_golden;
};

#endif
