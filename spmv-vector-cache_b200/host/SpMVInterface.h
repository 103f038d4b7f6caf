// SpMVInterface.h -- the operator interface of maltanar/spmv-vector-cache that CudaSpMV plugs in behind.
//
// Inside the reference tree, build with -DB200_USE_REFERENCE_HEADERS and this header simply pulls
// in the reference's own software/SpMV.h + software/SparseMatrix.h.  Standalone
// (this repository, the GPU box) it declares an interface-compatible mirror: same class names, same
// virtuals, same protected members, so CudaSpMV.cpp compiles unchanged in both settings.
//   SpMV          : software/SpMV.h:8-35        (ctor(A, x, y), exec(), statInt(), statKeys())
//   SparseMatrix  : software/SparseMatrix.h:18-70 (CSC container, u32 indices, 8-byte values)
//   SpMVData      : software/SparseMatrix.h:6 is `double`; -DSPMV_DATA_U64 selects the wrapping
//                   integer semiring of the *-uint64 matrix directories
#ifndef B200_SPMV_INTERFACE_H_
#define B200_SPMV_INTERFACE_H_

#ifdef B200_USE_REFERENCE_HEADERS
#include "SpMV.h"
#else

#include <string>
#include <vector>

typedef unsigned int SpMVIndex;
#ifdef SPMV_DATA_U64
typedef unsigned long long SpMVData;
#else
typedef double SpMVData;
#endif

// 28-byte header of <name>-meta.bin (software/SparseMatrix.h:8-16)
struct CompressedSparseMetadata {
    unsigned int numRows, numCols, numNZ, startingRow, indPtrBase, indBase, nzDataBase;
};

class SparseMatrix {
public:
    SparseMatrix() : m_rows(0), m_cols(0), m_nz(0), m_indPtrs(0), m_inds(0), m_nzData(0), m_rowStartsMarked(false) {}
    virtual ~SparseMatrix() {}
    bool isSquare() { return m_rows == m_cols; }
    void setName(std::string name) { m_name = name; }
    std::string getName() { return m_name; }
    unsigned int getRows() const { return m_rows; }
    unsigned int getCols() const { return m_cols; }
    unsigned int getNz() const { return m_nz; }
    SpMVIndex *getIndPtrs() const { return m_indPtrs; }
    SpMVIndex *getInds() const { return m_inds; }
    SpMVData *getNzData() const { return m_nzData; }

protected:
    unsigned int m_rows, m_cols, m_nz;
    SpMVIndex *m_indPtrs;  // column pointers [cols+1]
    SpMVIndex *m_inds;     // row indices [nz]
    SpMVData *m_nzData;    // values [nz]
    std::string m_name;
    bool m_rowStartsMarked;
};

class SpMV {
public:
    SpMV(SparseMatrix *A, SpMVData *x, SpMVData *y) : m_A(A), m_x(x), m_y(y) {}
    virtual ~SpMV() {}
    virtual bool exec() = 0;
    SparseMatrix *getA() const { return m_A; }
    SpMVData *getX() const { return m_x; }
    SpMVData *getY() const { return m_y; }
    virtual unsigned int statInt(std::string name) = 0;
    virtual std::vector<std::string> statKeys() = 0;

protected:
    SparseMatrix *m_A;
    SpMVData *m_x;
    SpMVData *m_y;
};

#endif  // B200_USE_REFERENCE_HEADERS
#endif
