// HostSparseMatrix.cpp -- see HostSparseMatrix.h.
#include "HostSparseMatrix.h"

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>

using namespace std;

void *b200_malloc_aligned(size_t alignment, size_t bytes)
{
    void *p = 0;
    if (posix_memalign(&p, alignment, bytes ? bytes : alignment) != 0) return 0;
    return p;
}
void b200_free_aligned(void *p) { free(p); }

// whole file into 64-byte aligned storage; *bytes = its size
static void *slurp(const string &path, size_t *bytes)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return 0;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    void *p = n >= 0 ? b200_malloc_aligned(64, (size_t)n) : 0;
    if (p && n > 0 && fread(p, 1, (size_t)n, f) != (size_t)n) { b200_free_aligned(p); p = 0; }
    fclose(f);
    if (p) *bytes = (size_t)n;
    return p;
}

HostSparseMatrix::~HostSparseMatrix()
{
    b200_free_aligned(m_indPtrs);
    b200_free_aligned(m_inds);
    b200_free_aligned(m_nzData);
    b200_free_aligned(m_golden);
}

HostSparseMatrix *HostSparseMatrix::fromDirectory(const string &root, const string &name)
{
    const string base = root + "/" + name + "/" + name;
    size_t nMeta = 0, nPtr = 0, nInd = 0, nData = 0, nGold = 0;
    CompressedSparseMetadata *md = (CompressedSparseMetadata *)slurp(base + "-meta.bin", &nMeta);
    if (!md || nMeta < sizeof(CompressedSparseMetadata)) {
        cerr << "Error: cannot read " << base << "-meta.bin" << endl;
        b200_free_aligned(md);
        return 0;
    }
    // same rule and wording as SparseMatrix::fromMemory (software/SparseMatrix.cpp:33-38)
    if (md->numCols == 0 || md->numRows == 0 || md->numNZ == 0) {
        cerr << "Error: matrix data at " << base << " is invalid" << endl;
        b200_free_aligned(md);
        return 0;
    }
    HostSparseMatrix *A = new HostSparseMatrix();
    A->m_rows = md->numRows;
    A->m_cols = md->numCols;
    A->m_nz = md->numNZ;
    b200_free_aligned(md);  // the three *Base fields are Zynq DRAM addresses: ignored here
    A->m_indPtrs = (SpMVIndex *)slurp(base + "-indptr.bin", &nPtr);
    A->m_inds = (SpMVIndex *)slurp(base + "-inds.bin", &nInd);
    A->m_nzData = (SpMVData *)slurp(base + "-data.bin", &nData);
    if (!A->m_indPtrs || !A->m_inds || !A->m_nzData || nPtr != ((size_t)A->m_cols + 1) * sizeof(SpMVIndex) ||
        nInd != (size_t)A->m_nz * sizeof(SpMVIndex) || nData != (size_t)A->m_nz * sizeof(SpMVData) ||
        A->m_indPtrs[0] != 0 || A->m_indPtrs[A->m_cols] != A->m_nz) {
        cerr << "Error: matrix data at " << base << " is invalid" << endl;
        delete A;
        return 0;
    }
    A->m_golden = (SpMVData *)slurp(root + "/" + name + "/golden.bin", &nGold);
    if (A->m_golden && nGold != (size_t)A->m_rows * sizeof(SpMVData)) { b200_free_aligned(A->m_golden); A->m_golden = 0; }
    A->m_rowStartsMarked = false;
    A->setName(name);
    return A;
}

HostSparseMatrix *HostSparseMatrix::fromArrays(unsigned rows, unsigned cols, unsigned nz, const SpMVIndex *colptr,
                                               const SpMVIndex *rowind, const SpMVData *vals)
{
    if (rows == 0 || cols == 0 || nz == 0 || !colptr || !rowind || !vals) {
        cerr << "Error: matrix data at " << (const void *)colptr << " is invalid" << endl;
        return 0;
    }
    HostSparseMatrix *A = new HostSparseMatrix();
    A->m_rows = rows; A->m_cols = cols; A->m_nz = nz;
    A->m_indPtrs = (SpMVIndex *)b200_malloc_aligned(64, ((size_t)cols + 1) * sizeof(SpMVIndex));
    A->m_inds = (SpMVIndex *)b200_malloc_aligned(64, (size_t)nz * sizeof(SpMVIndex));
    A->m_nzData = (SpMVData *)b200_malloc_aligned(64, (size_t)nz * sizeof(SpMVData));
    memcpy(A->m_indPtrs, colptr, ((size_t)cols + 1) * sizeof(SpMVIndex));
    memcpy(A->m_inds, rowind, (size_t)nz * sizeof(SpMVIndex));
    memcpy(A->m_nzData, vals, (size_t)nz * sizeof(SpMVData));
    A->m_rowStartsMarked = false;
    return A;
}
