// spmvbench.cpp -- host port of the reference's benchmark loop (software/main.cpp:102-144,146-264):
// confs x matrices, one CSV line per run with the variant's statKeys() plus the pseudo-keys
// "accType" and "matrix" (main.cpp:49-66), golden check inline (main.cpp:222-247).
//
//   spmvbench --root <matrices dir> --confs cuda,cuda-strict --matrices circuit204,i64k [--cms 1] [--iters N]
//   conf tokens: cuda[<device>][-strict][-nopersist][-overwrite][-cms][-prep][-vcmodel<depth>]
//     -prep            also run SoftwareSpMV::measurePreprocessingTimes' three passes on the GPU (cmstime, maxAlive...)
//     -vcmodel<depth>  also model the FPGA's vector cache of <depth> lines (NewCache stat keys readMisses, ...)
//   spmvbench --root <dir>          # no lists: the reference's prompts on stdin (confs ... q, CMS, matrices ... q)
//
// Built inside the reference tree (-DB200_USE_REFERENCE_HEADERS) the golden
// vector is SoftwareSpMV's result, exactly as main.cpp:222-224, and the "sw" conf runs benchmarkSW's
// loop.  Standalone there is no CPU implementation in this program (by design: no CPU fallback), the
// golden vector is the directory's golden.bin (A*1, fp64 directories) and "sw" is refused.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "HostSparseMatrix.h"
#include "SpMVFactory.h"
#ifdef B200_USE_REFERENCE_HEADERS
#include "SoftwareSpMV.h"
#endif

using namespace std;

static void printKeys(const vector<string> &keys)
{
    for (size_t i = 0; i < keys.size(); i++) cout << keys[i] << ",";
    cout << endl;
}

static void printResults(SpMV *spmv, const vector<string> &keys, const string &conf, const string &matrix)
{
    for (size_t i = 0; i < keys.size(); i++) {
        if (keys[i] == "matrix") cout << matrix << ",";
        else if (keys[i] == "accType") cout << SpMVFactory::name(conf) << ",";
        else cout << spmv->statInt(keys[i]) << ",";
    }
    cout << endl;
}

static vector<string> splitList(const string &s)
{
    vector<string> out;
    string tok;
    stringstream ss(s);
    while (getline(ss, tok, ',')) if (!tok.empty()) out.push_back(tok);
    return out;
}

static vector<string> promptList(const char *prompt)
{
    vector<string> out;
    string t;
    cout << prompt << endl;
    while (cin >> t && t != "q") out.push_back(t);
    return out;
}

int main(int argc, char **argv)
{
    string root = ".";
    vector<string> confs, ms;
    int cms = -1, iters = 0;
    for (int i = 1; i < argc; i++) {
        const string a = argv[i];
        if (a == "--root" && i + 1 < argc) root = argv[++i];
        else if (a == "--confs" && i + 1 < argc) confs = splitList(argv[++i]);
        else if (a == "--matrices" && i + 1 < argc) ms = splitList(argv[++i]);
        else if (a == "--cms" && i + 1 < argc) cms = atoi(argv[++i]);
        else if (a == "--iters" && i + 1 < argc) iters = atoi(argv[++i]);
        else { cerr << "usage: spmvbench --root DIR [--confs a,b --matrices m,n --cms 0|1 --iters N]" << endl; return 2; }
    }
    if (confs.empty()) confs = promptList("Enter list of configurations (cuda, cuda-strict, cuda-nopersist, sw), q to finalize: ");
    if (cms < 0) {
        if (ms.empty()) { cout << "Cold miss skip (0 to disable, 1 to enable): " << endl; cin >> cms; }
        else cms = 0;
    }
    if (ms.empty()) ms = promptList("Enter list of matrices, q to finalize: ");

    cout << "Benchmarking " << confs.size() << "x" << ms.size() << " confs x matrices..." << endl;
    cout << "=============================================================" << endl;
    bool keysBuilt = false;
    vector<string> keys;
    int failures = 0;
    for (size_t c = 0; c < confs.size(); c++) {
        string conf = confs[c];
        for (size_t m = 0; m < ms.size(); m++) {
            HostSparseMatrix *A = HostSparseMatrix::fromDirectory(root, ms[m]);
            if (!A) { failures++; continue; }
            SpMVData *x = (SpMVData *)b200_malloc_aligned(64, sizeof(SpMVData) * A->getCols());
            SpMVData *y = (SpMVData *)b200_malloc_aligned(64, sizeof(SpMVData) * A->getRows());
            for (SpMVIndex i = 0; i < A->getCols(); i++) x[i] = (SpMVData)1;
            for (SpMVIndex i = 0; i < A->getRows(); i++) y[i] = (SpMVData)0;
            SpMVData *golden = A->golden();
#ifdef B200_USE_REFERENCE_HEADERS
            SoftwareSpMV *check = new SoftwareSpMV(A, x);  // main.cpp:222-224
            check->exec();
            golden = check->getY();
            if (conf == "sw") {
                SoftwareSpMV *sw = new SoftwareSpMV(A, x, y);
                if (!keysBuilt) { keys = sw->statKeys(); keys.push_back("matrix"); printKeys(keys); keysBuilt = true; }
                sw->exec();
                printResults(sw, keys, conf, ms[m]);
                delete sw; delete check;
                b200_free_aligned(x); b200_free_aligned(y); delete A;
                continue;
            }
            if (cms) A->markRowStarts();  // main.cpp:228-229; the cuda variant then masks bits 31/30
#else
            if (conf == "sw") {
                cout << "sw: SoftwareSpMV is the reference's CPU code; build with -DB200_USE_REFERENCE_HEADERS inside its tree" << endl;
                b200_free_aligned(x); b200_free_aligned(y); delete A;
                failures++;
                break;
            }
#endif
            const string eff = (cms && conf.find("-cms") == string::npos) ? conf + "-cms" : conf;
            SpMV *spmv = SpMVFactory::make(eff, A, x, y);
            if (!spmv) { failures++; b200_free_aligned(x); b200_free_aligned(y); delete A; break; }
            if (!keysBuilt) {
                keys = spmv->statKeys();
                keys.push_back("accType");
                keys.push_back("matrix");
                printKeys(keys);
                keysBuilt = true;
            }
            if (!spmv->exec()) failures++;
            CudaSpMV *cu = dynamic_cast<CudaSpMV *>(spmv);
            if (cu && golden) cu->compareGolden(golden);
            if (cu && cu->options().prep && !cu->measurePreprocessingTimes()) failures++;  // main.cpp:139 (benchmarkSW)
            if (cu && cu->options().vcDepth > 0 && !cu->runVectorCacheModel()) failures++;
            printResults(spmv, keys, eff, ms[m]);
            if (cu && iters > 0) {  // device-resident timing, kernels only
                vector<float> t(iters);
                if (cu->execDeviceOnly(iters, &t[0])) {
                    float best = t[0];
                    for (int i = 1; i < iters; i++) if (t[i] < best) best = t[i];
                    cout << "# " << ms[m] << " " << eff << " device-only best " << best * 1e3f << " us over " << iters
                         << " runs, " << (double)A->getNz() / (best * 1e-3) / 1e9 << " GNNZ/s" << endl;
                }
            }
            delete spmv;
#ifdef B200_USE_REFERENCE_HEADERS
            delete check;
#endif
            b200_free_aligned(x);
            b200_free_aligned(y);
            delete A;
        }
    }
    cout << "=============================================================" << endl;
    cout << "Benchmarking complete" << endl;
    return failures ? 1 : 0;
}
