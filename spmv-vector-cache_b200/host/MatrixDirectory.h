// MatrixDirectory.h -- MatrixMarket reader and matrix-directory writer.
//
// C++ statement of what matrices/matrixutils.py does with scipy: loadMatrix (:163-169, mmread + tocsc),
// convertMatrix (:186-260: <name>-meta.bin with the 7 u32 fields and the 64-byte aligned Zynq DRAM
// addresses counted from dramBase = 0x8000100, -indptr.bin, -inds.bin, -data.bin), toUInt64Matrix
// (:100-103: every value becomes uint64 1) and makeGoldenResult (:108-113: golden.bin = A * ones).
// The files it writes are byte-identical to the reference's fixtures (tests/test_matrix_directory.py).
// golden.bin is computed by the caller (mtx2dir uses CudaSpMV in strict order: there is no CPU SpMV
// in this package).
#ifndef B200_MATRIXDIRECTORY_H_
#define B200_MATRIXDIRECTORY_H_

#include <stdint.h>

#include <string>
#include <vector>

struct CscArrays {
    uint32_t rows, cols;
    std::vector<uint32_t> colptr;  // [cols+1]
    std::vector<uint32_t> rowind;  // [nnz], ascending inside a column
    std::vector<double> vals;      // [nnz]
    CscArrays() : rows(0), cols(0) {}
    uint32_t nnz() const { return (uint32_t)rowind.size(); }
};

// coordinate real | integer | pattern, general | symmetric | skew-symmetric; duplicates are summed and
// row indices sorted inside each column, like scipy's mmread(...).tocsc().  false + *err on failure.
bool readMatrixMarket(const std::string &path, CscArrays *out, std::string *err);

// dtype 0: fp64 values as read; 1: the -uint64 variant (all values 1).  Creates <root>/<name>/.
bool writeMatrixDirectory(const std::string &root, const std::string &name, const CscArrays &A, int dtype,
                          std::string *err);
bool writeGolden(const std::string &root, const std::string &name, const void *y, uint32_t rows, std::string *err);

#endif
