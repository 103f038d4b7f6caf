// SpMVFactory.cpp -- see SpMVFactory.h.
#include "SpMVFactory.h"

#include <stdlib.h>

#include <iostream>

using namespace std;

static bool isFpgaName(const string &c)
{
    return c == "BufferNone" || c == "BufferSel" || c == "NewCache" || c == "BufferAll" || c == "OldCache";
}

bool SpMVFactory::parse(const string &conf, CudaSpMVOptions *opt)
{
    if (conf.compare(0, 4, "cuda") != 0) return false;
    CudaSpMVOptions o;
    size_t p = 4;
    if (p < conf.size() && conf[p] >= '0' && conf[p] <= '9') {
        char *end = 0;
        o.device = (int)strtol(conf.c_str() + p, &end, 10);
        p = (size_t)(end - conf.c_str());
    }
    while (p < conf.size()) {
        if (conf[p] != '-') return false;
        size_t q = conf.find('-', p + 1);
        if (q == string::npos) q = conf.size();
        const string tok = conf.substr(p + 1, q - p - 1);
        if (tok == "strict") o.strictOrder = true;
        else if (tok == "nopersist") o.persistX = false;
        else if (tok == "overwrite") o.accumulate = false;
        else if (tok == "cms") o.indexMask = 0x3FFFFFFFu;  // matrix went through markRowStarts (main.cpp:228-229)
        else if (tok == "prep") o.prep = true;             // measurePreprocessingTimes after exec (main.cpp:139)
        else if (tok.compare(0, 7, "vcmodel") == 0) {      // vcmodel<depth>: model the FPGA's vector cache of that many lines
            char *end = 0;
            const long d = strtol(tok.c_str() + 7, &end, 10);
            if (*end != 0 || d < 1 || d > (1L << 30) || (d & (d - 1)) != 0) return false;
            o.vcDepth = (int)d;
        }
        else return false;
        p = q;
    }
    if (opt) *opt = o;
    return true;
}

SpMV *SpMVFactory::make(const string &conf, SparseMatrix *A, SpMVData *x, SpMVData *y)
{
    CudaSpMVOptions opt;
    if (parse(conf, &opt)) return new CudaSpMV(A, x, y, opt);
    if (isFpgaName(conf))
        cout << "Accelerator " << conf << " requires the Zynq programmable logic; not available on this host" << endl;
    else
        cout << "Accelerator signature unrecognized! " << conf << endl;
    return 0;
}

string SpMVFactory::name(const string &conf)
{
    if (parse(conf, 0)) return "CUDA-B200";
    if (isFpgaName(conf)) return conf;
    return "<undefined>";
}
