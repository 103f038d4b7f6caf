// mtx2dir.cpp -- MatrixMarket file -> matrix directory (the reference's matrices/matrixutils.py
// prepareUFLMatrix / convertMatrix / makeGoldenResult, :76-113,186-260, as a host tool).
//   mtx2dir --mtx circuit204.mtx --root out --name circuit204            fp64 directory + golden.bin
//   mtx2dir --mtx circuit204.mtx --root out --name circuit204-uint64 --uint64   values = uint64 1, no golden
// golden.bin = A * ones is computed on the GPU by CudaSpMV in strict order (bit-equal to the reference's
// scipy product, which visits entries in the same CSC order); --no-golden skips it (no GPU needed).
#include <stdio.h>
#include <string.h>

#include <iostream>
#include <string>

#include "CudaSpMV.h"
#include "HostSparseMatrix.h"
#include "MatrixDirectory.h"

using namespace std;

int main(int argc, char **argv)
{
    string mtx, root = ".", name;
    bool u64 = false, golden = true;
    for (int i = 1; i < argc; i++) {
        const string a = argv[i];
        if (a == "--mtx" && i + 1 < argc) mtx = argv[++i];
        else if (a == "--root" && i + 1 < argc) root = argv[++i];
        else if (a == "--name" && i + 1 < argc) name = argv[++i];
        else if (a == "--uint64") u64 = true;
        else if (a == "--no-golden") golden = false;
        else { cerr << "usage: mtx2dir --mtx FILE --root DIR --name NAME [--uint64] [--no-golden]" << endl; return 2; }
    }
    if (mtx.empty() || name.empty()) { cerr << "mtx2dir: --mtx and --name are required" << endl; return 2; }
    CscArrays A;
    string err;
    if (!readMatrixMarket(mtx, &A, &err)) { cerr << "mtx2dir: " << mtx << ": " << err << endl; return 1; }
    if (!writeMatrixDirectory(root, name, A, u64 ? 1 : 0, &err)) { cerr << "mtx2dir: " << err << endl; return 1; }
    cout << "Rows = " << A.rows << endl << "Cols = " << A.cols << endl << "NonZ = " << A.nnz() << endl;  // matrixutils.py:252-254
    if (golden && !u64) {
#ifndef SPMV_DATA_U64
        HostSparseMatrix *M = HostSparseMatrix::fromArrays(A.rows, A.cols, A.nnz(), A.colptr.data(), A.rowind.data(), A.vals.data());
        if (!M) return 1;
        CudaSpMVOptions opt;
        opt.strictOrder = true;
        CudaSpMV spmv(M, 0, 0, opt);  // x = 1, y = 0 (matrixutils.py:105-106 makeUnitVector)
        if (!spmv.exec()) { cerr << "mtx2dir: golden.bin needs a GPU (or pass --no-golden): " << spmv.lastError() << endl; delete M; return 1; }
        if (!writeGolden(root, name, spmv.getY(), A.rows, &err)) { cerr << "mtx2dir: " << err << endl; delete M; return 1; }
        delete M;
#endif
    }
    return 0;
}
