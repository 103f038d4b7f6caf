// MatrixDirectory.cpp -- see MatrixDirectory.h.
#include "MatrixDirectory.h"

#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <fstream>
#include <sstream>

using namespace std;

namespace {

struct Triplet {
    uint32_t col, row;
    double val;
    uint64_t seq;  // input order: duplicates are summed in the order they were read
};
bool tripletLess(const Triplet &a, const Triplet &b)
{
    if (a.col != b.col) return a.col < b.col;
    if (a.row != b.row) return a.row < b.row;
    return a.seq < b.seq;
}
string lower(string s)
{
    for (size_t i = 0; i < s.size(); i++) s[i] = (char)tolower(s[i]);
    return s;
}
bool fail(string *err, const string &msg)
{
    if (err) *err = msg;
    return false;
}
// matrixutils.py:171-178 alignedIncrement
uint32_t alignedIncrement(uint32_t base, uint32_t increment, uint32_t align)
{
    uint32_t res = base + increment;
    const uint32_t rem = res % align;
    if (rem != 0) res += align - rem;
    return res;
}
bool writeFile(const string &path, const void *p, size_t bytes, string *err)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) return fail(err, "cannot write " + path + ": " + strerror(errno));
    const bool ok = bytes == 0 || fwrite(p, 1, bytes, f) == bytes;
    fclose(f);
    return ok ? true : fail(err, "short write on " + path);
}

}  // namespace

bool readMatrixMarket(const string &path, CscArrays *out, string *err)
{
    ifstream in(path.c_str());
    if (!in) return fail(err, "cannot open " + path);
    string line;
    if (!getline(in, line)) return fail(err, "empty file");
    string banner, object, format, field, symmetry;
    {
        stringstream ss(line);
        ss >> banner >> object >> format >> field >> symmetry;
    }
    if (banner != "%%MatrixMarket" || lower(object) != "matrix") return fail(err, "not a MatrixMarket matrix file");
    format = lower(format); field = lower(field); symmetry = lower(symmetry);
    if (format != "coordinate") return fail(err, "only the coordinate format is supported");
    const bool pattern = field == "pattern";
    if (!pattern && field != "real" && field != "integer" && field != "double") return fail(err, "unsupported field " + field);
    const bool symm = symmetry == "symmetric", skew = symmetry == "skew-symmetric";
    if (!symm && !skew && symmetry != "general") return fail(err, "unsupported symmetry " + symmetry);
    while (getline(in, line)) if (!line.empty() && line[0] != '%') break;
    unsigned long long rows = 0, cols = 0, entries = 0;
    if (sscanf(line.c_str(), "%llu %llu %llu", &rows, &cols, &entries) != 3) return fail(err, "bad size line");
    if (rows == 0 || cols == 0 || rows > 0xFFFFFFFFull || cols > 0xFFFFFFFFull) return fail(err, "bad dimensions");
    vector<Triplet> t;
    t.reserve((size_t)entries * ((symm || skew) ? 2 : 1));
    for (unsigned long long k = 0; k < entries; k++) {
        unsigned long long r = 0, c = 0;
        double v = 1.0;
        if (!(in >> r >> c)) return fail(err, "file ends before all entries were read");
        if (!pattern && !(in >> v)) return fail(err, "missing value");
        if (r < 1 || r > rows || c < 1 || c > cols) return fail(err, "entry out of range");
        Triplet a = {(uint32_t)(c - 1), (uint32_t)(r - 1), v, t.size()};
        t.push_back(a);
        if ((symm || skew) && r != c) {
            Triplet b = {(uint32_t)(r - 1), (uint32_t)(c - 1), skew ? -v : v, t.size()};
            t.push_back(b);
        }
    }
    sort(t.begin(), t.end(), tripletLess);
    out->rows = (uint32_t)rows;
    out->cols = (uint32_t)cols;
    out->colptr.assign((size_t)cols + 1, 0);
    out->rowind.clear();
    out->vals.clear();
    for (size_t i = 0; i < t.size();) {
        size_t j = i;
        double sum = 0.0;
        while (j < t.size() && t[j].col == t[i].col && t[j].row == t[i].row) sum += t[j++].val;  // duplicates summed
        out->rowind.push_back(t[i].row);
        out->vals.push_back(sum);
        out->colptr[t[i].col + 1]++;
        i = j;
    }
    if (out->rowind.size() > 0xFFFFFFFFull) return fail(err, "more than 2^32-1 entries (numNZ is a u32, SparseMatrix.h:11)");
    for (size_t c = 0; c < cols; c++) out->colptr[c + 1] += out->colptr[c];
    return true;
}

bool writeMatrixDirectory(const string &root, const string &name, const CscArrays &A, int dtype, string *err)
{
    const string dir = root + "/" + name;
    if (mkdir(dir.c_str(), 0777) != 0 && errno != EEXIST) return fail(err, "cannot create " + dir + ": " + strerror(errno));
    const uint32_t nnz = A.nnz();
    // DRAM placement of matrixutils.py:186-245: metadata at dramBase, then each array 64-byte aligned
    const uint32_t dramBase = 0x8000100u, burstAlign = 64;
    uint32_t addr = alignedIncrement(dramBase, 28, burstAlign);
    const uint32_t indPtrBase = addr;
    addr = alignedIncrement(addr, (uint32_t)((A.cols + 1) * 4ull), burstAlign);
    const uint32_t indBase = addr;
    addr = alignedIncrement(addr, (uint32_t)(nnz * 4ull), burstAlign);
    const uint32_t nzDataBase = addr;
    const uint32_t meta[7] = {A.rows, A.cols, nnz, 0u /* startingRow */, indPtrBase, indBase, nzDataBase};
    const string base = dir + "/" + name;
    if (!writeFile(base + "-meta.bin", meta, sizeof meta, err)) return false;
    if (!writeFile(base + "-indptr.bin", A.colptr.data(), A.colptr.size() * 4, err)) return false;
    if (!writeFile(base + "-inds.bin", A.rowind.data(), (size_t)nnz * 4, err)) return false;
    if (dtype == 1) {
        vector<uint64_t> ones(nnz, 1ull);  // matrixutils.py:100-103 toUInt64Matrix
        return writeFile(base + "-data.bin", ones.data(), (size_t)nnz * 8, err);
    }
    return writeFile(base + "-data.bin", A.vals.data(), (size_t)nnz * 8, err);
}

bool writeGolden(const string &root, const string &name, const void *y, uint32_t rows, string *err)
{
    return writeFile(root + "/" + name + "/golden.bin", y, (size_t)rows * 8, err);
}
