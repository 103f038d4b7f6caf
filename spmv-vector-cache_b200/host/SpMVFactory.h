// SpMVFactory.h -- selects an SpMV variant by configuration token, next to the FPGA variants.
//
// The reference picks its variant from a hardware signature register
// (software/HWSpMVFactory.cpp:20-38) after main.cpp:195-200 loaded the bitfile named by the conf
// token ("sw" = SoftwareSpMV, anything else = a bitfile).  On a GPU host the token selects directly:
//   "cuda"            CudaSpMV, y += A*x, x pinned in L2
//   "cuda-strict"     same, fp64 summed in the reference's order (memcmp-equal to SoftwareSpMV)
//   "cuda-nopersist"  same as "cuda" without the L2 access-policy window (A/B for the "vector cache")
//   "cuda-overwrite"  y = A*x (what the FPGA variants compute after their zero-fill)
//   "cudaN"           device ordinal N (e.g. "cuda3"); suffixes combine: "cuda3-strict"
// The four FPGA names ("BufferNone", "BufferSel", "NewCache", "BufferAll") are recognised and
// rejected with a message (they need the Zynq programmable logic); unknown tokens likewise, mirroring
// the reference's "Accelerator signature unrecognized!" + null (its assert(0) is left to the caller).
#ifndef B200_SPMVFACTORY_H_
#define B200_SPMVFACTORY_H_

#include <string>

#include "CudaSpMV.h"

class SpMVFactory {
public:
    static SpMV *make(const std::string &conf, SparseMatrix *A, SpMVData *x, SpMVData *y);
    static std::string name(const std::string &conf);  // "CUDA-B200" for the cuda tokens (HWSpMVFactory::name)
    static bool parse(const std::string &conf, CudaSpMVOptions *opt);  // false: not a cuda token
};

#endif
