// See wenobase_cache.h. Writer first (mirrors WENOBase::writeList), reader below.
#include "wenobase_cache.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>

namespace wenoext {

int32_t matrixKey(const double* A, size_t n) {
    double mx = 0;
    for (size_t i = 0; i < n; ++i) mx = std::max(mx, std::fabs(A[i]));
    int32_t key = 0;
    if (mx == 0) return key;
    const double mult = 1.0e6 / mx;
    for (size_t i = 0; i < n; ++i) key += (int32_t)(A[i] * mult);
    return key;
}

namespace {

[[noreturn]] void fail(const std::string& what) { throw std::runtime_error("WENOBase cache: " + what); }

// ---------------------------------------------------------------- writer ----
struct Out {
    FILE* f = nullptr;
    std::string path;
    int labelBytes;
    Out(const std::string& p, int lb) : path(p), labelBytes(lb) {
        f = fopen(p.c_str(), "wb");
        if (!f) fail("cannot create " + p);
    }
    ~Out() {
        if (f) fclose(f);
    }
    void text(const char* s) { fputs(s, f); }
    void num(long long v) { fprintf(f, "%lld", v); }
    void numNl(long long v) { fprintf(f, "%lld\n", v); }  // `os << v << endl`
    void raw(const void* p, size_t bytes) {                // OSstream::write(const char*, streamsize)
        fputc('(', f);
        if (bytes && fwrite(p, 1, bytes, f) != bytes) fail("short write to " + path);
        fputc(')', f);
    }
    // List<label>, contiguous, BINARY: "\nN\n" then the raw block if N > 0
    void labelList(const int32_t* v, size_t n) {
        fprintf(f, "\n%zu\n", n);
        if (!n) return;
        if (labelBytes == 4) {
            raw(v, n * 4);
        } else {
            std::vector<int64_t> w(v, v + n);
            raw(w.data(), n * 8);
        }
    }
    // head and tail of a List of a non-contiguous type
    void listBegin(size_t n) { fprintf(f, "\n%zu\n(", n); }
    void listEnd() { text("\n)\n"); }
    // blaze::DynamicMatrix<double,rowMajor>, BlazeIO.C:103-134
    void blazeRowMajor(const double* A, int rows, int cols, int simd) {
        const int spacing = simd > 1 ? (cols + simd - 1) / simd * simd : cols;
        numNl(rows);
        numNl(cols);
        numNl(spacing);
        std::vector<double> pad((size_t)rows * spacing, 0.0);
        for (int i = 0; i < rows; ++i) memcpy(&pad[(size_t)i * spacing], A + (size_t)i * cols, 8 * (size_t)cols);
        raw(pad.data(), pad.size() * 8);
    }
};

void writeStencilLists(const HostPlan& hp, const std::string& path, int lb, bool procMap) {
    Out o(path, lb);
    o.listBegin((size_t)hp.nCells);
    std::vector<int32_t> tmp;
    for (int c = 0; c < hp.nCells; ++c) {
        const int64_t s0 = hp.cellStencilStart[c], s1 = hp.cellStencilStart[c + 1];
        o.listBegin((size_t)(s1 - s0));
        for (int64_t s = s0; s < s1; ++s) {
            const int64_t e0 = hp.stencilEntryStart[s], e1 = hp.stencilEntryStart[s + 1];
            const int32_t* src = &hp.stencilsID[e0];
            if (procMap) {
                if (!hp.cellToProcMap.empty()) {
                    src = &hp.cellToProcMap[e0];
                } else {  // serial plan: every entry local, deleted slots keep their mark
                    tmp.assign((size_t)(e1 - e0), CELL_LOCAL);
                    if (hp.stencilsID[e0] == CELL_DELETED) tmp[0] = CELL_DELETED;
                    src = tmp.data();
                }
            }
            o.labelList(src, (size_t)(e1 - e0));
        }
        o.listEnd();
    }
    o.listEnd();
}

}  // namespace

void writeWenoBaseCache(const HostPlan& hp, const std::string& dir, const CacheFormat& fmt) {
    if (fmt.labelBytes != 4 && fmt.labelBytes != 8) fail("labelBytes must be 4 or 8");
    if (mkdir(dir.c_str(), 0777) != 0 && errno != EEXIST) fail("cannot create directory " + dir);
    const int lb = fmt.labelBytes;
    const int K = hp.nDvt;
    {
        Out o(dir + "/sendProcList", lb);
        o.labelList(hp.sendProcList.data(), hp.sendProcList.size());
    }
    {
        Out o(dir + "/receiveProcList", lb);
        o.labelList(hp.receiveProcList.data(), hp.receiveProcList.size());
    }
    {
        Out o(dir + "/DimLists", lb);
        o.listBegin((size_t)hp.nCells);
        for (int c = 0; c < hp.nCells; ++c) o.labelList(hp.dim, 3);
        o.listEnd();
    }
    writeStencilLists(hp, dir + "/StencilIDs", lb, false);
    writeStencilLists(hp, dir + "/CellToProcMap", lb, true);
    {
        // matrixDB::write (matrixDB.C:276-357): the database grouped by key in ascending key
        // order (a std::multimap keeps equal keys in insertion order), then per (cell, stencil)
        // a valid bit, the key and the position inside the key's group.
        Out o(dir + "/Pseudoinverses", lb);
        const int nU = (int)hp.matrixCols.size();
        std::multimap<int32_t, int32_t> db;
        std::vector<int32_t> keyOf(nU);
        for (int u = 0; u < nU; ++u) {
            keyOf[u] = matrixKey(&hp.matrixData[hp.matrixStart[u]], (size_t)(hp.matrixStart[u + 1] - hp.matrixStart[u]));
            db.emplace(keyOf[u], u);
        }
        std::vector<int32_t> posOf(nU, 0);
        o.numNl(nU);
        for (auto it = db.begin(); it != db.end();) {
            const int32_t key = it->first;
            const int count = (int)db.count(key);
            o.numNl(key);
            o.numNl(count);
            for (int i = 0; i < count; ++i, ++it) {
                posOf[it->second] = i;
                o.blazeRowMajor(&hp.matrixData[hp.matrixStart[it->second]], K, hp.matrixCols[it->second],
                                fmt.simdDoubles);
                o.text("\n");
            }
        }
        o.numNl(hp.nCells);
        for (int c = 0; c < hp.nCells; ++c) {
            const int64_t s0 = hp.cellStencilStart[c], s1 = hp.cellStencilStart[c + 1];
            o.numNl(s1 - s0);
            for (int64_t s = s0; s < s1; ++s) {
                const int32_t u = hp.stencilMatrix[s];
                o.numNl(u >= 0 ? 1 : 0);
                if (u < 0) continue;
                o.numNl(keyOf[u]);
                o.numNl(posOf[u]);
            }
        }
    }
    {
        Out o(dir + "/sendHaloCellIDList", lb);
        o.listBegin(hp.sendHaloCellIDList.size());
        for (const auto& l : hp.sendHaloCellIDList) o.labelList(l.data(), l.size());
        o.listEnd();
    }
    {
        Out o(dir + "/receiveHaloSize", lb);
        o.labelList(hp.receiveHaloSize.data(), hp.receiveHaloSize.size());
    }
    {
        Out o(dir + "/B", lb);
        o.listBegin((size_t)hp.nCells);
        for (int c = 0; c < hp.nCells; ++c) {
            o.text("\n");
            o.blazeRowMajor(&hp.B[(size_t)c * K * K], K, K, fmt.simdDoubles);
        }
        o.listEnd();
    }
    {
        // WENOBase.C:1300-1307: the count, then both List3D of every face back to back
        Out o(dir + "/intBasTrans", lb);
        const int n1 = hp.polOrder + 1;
        const size_t cube = (size_t)n1 * n1 * n1;
        o.numNl(hp.nFaces);
        for (size_t f = 0; f < (size_t)hp.nFaces; ++f)
            for (int side = 0; side < 2; ++side) {
                o.numNl(n1);
                o.numNl(n1);
                o.numNl(n1);
                o.raw(&hp.intBasTrans[(f * 2 + side) * cube], cube * 8);
            }
    }
    {
        Out o(dir + "/refFacAr", lb);
        fprintf(o.f, "\n%zu\n", hp.refFacAr.size());
        if (!hp.refFacAr.empty()) o.raw(hp.refFacAr.data(), hp.refFacAr.size() * 8);
    }
}


// ---------------------------------------------------------------- reader ----
namespace {

struct In {
    const unsigned char* p = nullptr;
    size_t n = 0, pos = 0;
    std::string path;
    int labelBytes = 4;
    explicit In(const std::string& file) : path(file) {
        int fd = open(file.c_str(), O_RDONLY);
        if (fd < 0) fail("cannot open " + file);
        struct stat st;
        if (fstat(fd, &st) != 0) {
            close(fd);
            fail("cannot stat " + file);
        }
        n = (size_t)st.st_size;
        if (n) {
            void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
            if (m == MAP_FAILED) {
                close(fd);
                fail("cannot map " + file);
            }
            p = (const unsigned char*)m;
        }
        close(fd);
    }
    ~In() {
        if (p) munmap((void*)p, n);
    }
    In(const In&) = delete;
    [[noreturn]] void bad(const std::string& what) const {
        fail(path + " at byte " + std::to_string(pos) + ": " + what);
    }
    void skipWs() {
        while (pos < n && (p[pos] == ' ' || p[pos] == '\n' || p[pos] == '\t' || p[pos] == '\r')) ++pos;
    }
    long long integer() {  // ISstream reads numbers as text in either format
        skipWs();
        size_t q = pos;
        bool neg = false;
        if (q < n && (p[q] == '-' || p[q] == '+')) neg = p[q++] == '-';
        if (q >= n || p[q] < '0' || p[q] > '9') bad("integer expected");
        long long v = 0;
        while (q < n && p[q] >= '0' && p[q] <= '9') v = v * 10 + (p[q++] - '0');
        pos = q;
        return neg ? -v : v;
    }
    long long count() {
        long long v = integer();
        if (v < 0) bad("negative size");
        return v;
    }
    bool boolean() {  // Ostream << bool writes 0/1; words accepted as Switch would
        skipWs();
        if (pos < n && p[pos] >= '0' && p[pos] <= '9') return integer() != 0;
        size_t q = pos;
        while (q < n && ((p[q] >= 'a' && p[q] <= 'z') || (p[q] >= 'A' && p[q] <= 'Z'))) ++q;
        std::string w((const char*)p + pos, q - pos);
        pos = q;
        if (w == "true" || w == "yes" || w == "on" || w == "y" || w == "t") return true;
        if (w == "false" || w == "no" || w == "off" || w == "n" || w == "f") return false;
        bad("bool expected");
    }
    void expect(char c) {
        skipWs();
        if (pos >= n || p[pos] != (unsigned char)c) bad(std::string("'") + c + "' expected");
        ++pos;
    }
    const unsigned char* raw(size_t bytes) {  // ISstream::read(char*, streamsize): '(' bytes ')'
        expect('(');
        if (bytes > n - pos) bad("raw block runs past the end of the file");
        const unsigned char* r = p + pos;
        pos += bytes;
        if (pos >= n || p[pos] != ')') bad("')' expected after a raw block (wrong label size?)");
        ++pos;
        return r;
    }
    // "N" + raw block if N > 0 (an empty contiguous list may also appear as "0()")
    void labelList(std::vector<int32_t>& out, bool append = false) {
        if (!append) out.clear();
        const size_t cnt = (size_t)count();
        if (cnt > n) bad("list longer than the file");
        if (!cnt) {
            size_t save = pos;
            skipWs();
            if (pos < n && p[pos] == '(') {
                ++pos;
                skipWs();
                if (pos < n && p[pos] == ')') {
                    ++pos;
                    return;
                }
            }
            pos = save;
            return;
        }
        const unsigned char* r = raw(cnt * (size_t)labelBytes);
        const size_t at = out.size();
        out.resize(at + cnt);
        if (labelBytes == 4) {
            memcpy(&out[at], r, cnt * 4);
        } else {
            for (size_t i = 0; i < cnt; ++i) {
                int64_t v;
                memcpy(&v, r + 8 * i, 8);
                if (v < INT32_MIN || v > INT32_MAX) bad("label does not fit 32 bits");
                out[at + i] = (int32_t)v;
            }
        }
    }
    // BlazeIO.C:30-62: rows, cols, spacing, rows*spacing doubles -> unpadded row-major
    void blazeRowMajor(std::vector<double>& out, int& rows, int& cols, int& spacing) {
        rows = (int)count();
        cols = (int)count();
        spacing = (int)count();
        if (spacing < cols) bad("matrix spacing smaller than its column count");
        const unsigned char* r = raw((size_t)rows * spacing * 8);
        const size_t at = out.size();
        out.resize(at + (size_t)rows * cols);
        for (int i = 0; i < rows; ++i) memcpy(&out[at + (size_t)i * cols], r + (size_t)i * spacing * 8, (size_t)cols * 8);
    }
    void end() {
        skipWs();
        if (pos != n) bad("trailing bytes");
    }
};

void readStencilLists(const std::string& path, int lb, std::vector<int64_t>& cellStart,
                      std::vector<int64_t>& entryStart, std::vector<int32_t>& ids) {
    In in(path);
    in.labelBytes = lb;
    cellStart.assign(1, 0);
    entryStart.assign(1, 0);
    ids.clear();
    const long long nCells = in.count();
    in.expect('(');
    for (long long c = 0; c < nCells; ++c) {
        const long long nSt = in.count();
        in.expect('(');
        for (long long s = 0; s < nSt; ++s) {
            in.labelList(ids, true);
            entryStart.push_back((int64_t)ids.size());
        }
        in.expect(')');
        cellStart.push_back((int64_t)entryStart.size() - 1);
    }
    in.expect(')');
    in.end();
}

void readLabelFile(const std::string& path, int lb, std::vector<int32_t>& out) {
    In in(path);
    in.labelBytes = lb;
    in.labelList(out);
    in.end();
}

void readAllLabelFiles(const std::string& dir, int lb, HostPlan& hp, std::vector<int32_t>& dimLists) {
    readLabelFile(dir + "/sendProcList", lb, hp.sendProcList);
    readLabelFile(dir + "/receiveProcList", lb, hp.receiveProcList);
    readLabelFile(dir + "/receiveHaloSize", lb, hp.receiveHaloSize);
    {
        In in(dir + "/DimLists");
        in.labelBytes = lb;
        dimLists.clear();
        const long long nc = in.count();
        in.expect('(');
        for (long long c = 0; c < nc; ++c) {
            const size_t before = dimLists.size();
            in.labelList(dimLists, true);
            if (dimLists.size() - before != 3) in.bad("a dimList entry must have 3 labels");
        }
        in.expect(')');
        in.end();
    }
    readStencilLists(dir + "/StencilIDs", lb, hp.cellStencilStart, hp.stencilEntryStart, hp.stencilsID);
    std::vector<int64_t> cs, es;
    readStencilLists(dir + "/CellToProcMap", lb, cs, es, hp.cellToProcMap);
    if (cs != hp.cellStencilStart || es != hp.stencilEntryStart) fail("StencilIDs and CellToProcMap differ in shape");
    {
        In in(dir + "/sendHaloCellIDList");
        in.labelBytes = lb;
        const long long np = in.count();
        in.expect('(');
        hp.sendHaloCellIDList.assign((size_t)np, {});
        for (long long q = 0; q < np; ++q) in.labelList(hp.sendHaloCellIDList[(size_t)q]);
        in.expect(')');
        in.end();
    }
}

}  // namespace

void readWenoBaseCache(const std::string& dir, int32_t myProc, HostPlan& hp, CacheFormat& fmt) {
    // label size: the raw blocks are framed by '(' and ')', so a wrong size fails to parse
    std::vector<int32_t> dimLists;
    if (fmt.labelBytes == 4 || fmt.labelBytes == 8) {
        readAllLabelFiles(dir, fmt.labelBytes, hp, dimLists);
    } else {
        try {
            readAllLabelFiles(dir, 4, hp, dimLists);
            fmt.labelBytes = 4;
        } catch (const std::runtime_error& e4) {
            try {
                readAllLabelFiles(dir, 8, hp, dimLists);
                fmt.labelBytes = 8;
            } catch (const std::runtime_error& e8) {
                fail(std::string("neither 32- nor 64-bit labels parse: [32] ") + e4.what() + " [64] " + e8.what());
            }
        }
    }
    hp.nCells = (int32_t)(hp.cellStencilStart.size() - 1);
    hp.nProcs = (int32_t)hp.sendProcList.size();
    hp.myProc = myProc;
    if (hp.receiveProcList.size() != (size_t)hp.nProcs || hp.receiveHaloSize.size() != (size_t)hp.nProcs ||
        hp.sendHaloCellIDList.size() != (size_t)hp.nProcs)
        fail("halo lists differ in length");
    if (dimLists.size() != (size_t)hp.nCells * 3) fail("DimLists does not have one entry per cell");
    for (int i = 0; i < 3; ++i) hp.dim[i] = hp.nCells ? dimLists[i] : 0;
    for (size_t i = 0; i < dimLists.size(); ++i)
        if (dimLists[i] != hp.dim[i % 3]) fail("DimLists entries differ between cells");  // WENOBase.C:1147-1154
    const int64_t nStencils = (int64_t)hp.stencilEntryStart.size() - 1;

    {
        // matrixDB::read (matrixDB.C:360-415)
        In in(dir + "/Pseudoinverses");
        const long long dbSize = in.count();
        hp.matrixStart.assign(1, 0);
        hp.matrixCols.clear();
        hp.matrixData.clear();
        std::map<int32_t, int32_t> firstOfKey, countOfKey;
        int rows = 0, cols = 0, spacing = 0;
        hp.nDvt = 0;
        for (long long i = 0; i < dbSize;) {
            const int32_t key = (int32_t)in.integer();
            const int cnt = (int)in.count();
            if (cnt == 0) in.bad("empty key group");
            if (firstOfKey.count(key)) in.bad("key group appears twice");
            firstOfKey[key] = (int32_t)hp.matrixCols.size();
            countOfKey[key] = cnt;
            for (int k = 0; k < cnt; ++k, ++i) {
                in.blazeRowMajor(hp.matrixData, rows, cols, spacing);
                if (hp.matrixCols.empty()) {
                    hp.nDvt = rows;
                    if (cols > 0 && spacing % 8 == 0) fmt.simdDoubles = 8;
                    else if (spacing % 4 == 0) fmt.simdDoubles = 4;
                    else if (spacing % 2 == 0) fmt.simdDoubles = 2;
                    else fmt.simdDoubles = 1;
                }
                if (rows != hp.nDvt) in.bad("matrices differ in row count");
                hp.matrixCols.push_back(cols);
                hp.matrixStart.push_back((int64_t)hp.matrixData.size());
            }
        }
        const long long nc = in.count();
        if (nc != hp.nCells) in.bad("LSmatrix and StencilIDs differ in cell count");
        hp.stencilMatrix.assign((size_t)nStencils, -1);
        hp.nMatricesTotal = 0;
        for (int c = 0; c < hp.nCells; ++c) {
            const long long nSt = in.count();
            if (nSt != hp.cellStencilStart[c + 1] - hp.cellStencilStart[c])
                in.bad("LSmatrix and StencilIDs differ in stencil count of cell " + std::to_string(c));
            for (long long s = 0; s < nSt; ++s) {
                if (!in.boolean()) continue;
                const int32_t key = (int32_t)in.integer();
                const long long posn = in.count();
                auto it = firstOfKey.find(key);
                if (it == firstOfKey.end() || posn >= countOfKey[key]) in.bad("matrix reference outside the database");
                hp.stencilMatrix[(size_t)(hp.cellStencilStart[c] + s)] = it->second + (int32_t)posn;
                ++hp.nMatricesTotal;
            }
        }
        in.end();
    }
    int K = hp.nDvt;  // 0 if the mesh has no live stencil at all: taken from B below
    // a live stencil's matrix has one column per entry after the cell itself (WENOBase.C:976-979)
    for (int64_t s = 0; s < nStencils; ++s) {
        const int64_t e0 = hp.stencilEntryStart[s], e1 = hp.stencilEntryStart[s + 1];
        if (e1 == e0) fail("empty stencil list");
        const bool deleted = hp.stencilsID[e0] == CELL_DELETED;
        const int32_t u = hp.stencilMatrix[s];
        if (!deleted && u >= 0 && hp.matrixCols[u] != e1 - e0 - 1)
            fail("stencil " + std::to_string(s) + ": matrix columns and stencil size disagree");
    }
    {
        In in(dir + "/B");
        const long long nc = in.count();
        if (nc != hp.nCells) in.bad("B does not have one matrix per cell");
        in.expect('(');
        hp.B.clear();
        hp.B.reserve((size_t)hp.nCells * K * K);
        int rows, cols, spacing;
        for (int c = 0; c < hp.nCells; ++c) {
            const size_t before = hp.B.size();
            in.blazeRowMajor(hp.B, rows, cols, spacing);
            if (K == 0 && rows > 0 && rows == cols) {
                K = hp.nDvt = rows;
                hp.B.insert(hp.B.begin(), (size_t)c * K * K, 0.0);  // default-constructed ones before it
            }
            if (rows == 0 && cols == 0) {  // cells the setup skipped keep a default-constructed matrix
                hp.B.resize(before + (size_t)K * K, 0.0);
            } else if (rows != K || cols != K) {
                in.bad("B matrix is not nDvt x nDvt");
            }
        }
        in.expect(')');
        in.end();
    }
    {
        In in(dir + "/intBasTrans");
        const long long nf = in.count();
        hp.intBasTrans.clear();
        hp.polOrder = -1;
        size_t cube = 0;
        for (long long f = 0; f < nf; ++f)
            for (int side = 0; side < 2; ++side) {
                const long long sx = in.count(), sy = in.count(), sz = in.count();
                if (hp.polOrder < 0) {
                    hp.polOrder = (int32_t)sx - 1;
                    cube = (size_t)(sx * sy * sz);
                    hp.intBasTrans.reserve((size_t)nf * 2 * cube);
                }
                if (sx != hp.polOrder + 1 || sy != sx || sz != sx) in.bad("List3D is not (polOrder+1)^3");
                const unsigned char* r = in.raw(cube * 8);
                const size_t at = hp.intBasTrans.size();
                hp.intBasTrans.resize(at + cube);
                memcpy(&hp.intBasTrans[at], r, cube * 8);
            }
        in.end();
        hp.nFaces = (int32_t)nf;
    }
    {
        In in(dir + "/refFacAr");
        const size_t nf = (size_t)in.count();
        hp.refFacAr.assign(nf, 0.0);
        if (nf) memcpy(hp.refFacAr.data(), in.raw(nf * 8), nf * 8);
        in.end();
        if ((long long)nf != hp.nFaces) fail("refFacAr and intBasTrans differ in face count");
    }
    // nDvt must be the one the order and the dimensions imply (WENOBase.C:1130-1140)
    {
        const int r = hp.polOrder;
        const int nd = (hp.dim[0] > 0) + (hp.dim[1] > 0) + (hp.dim[2] > 0);
        const int want = nd == 3 ? (r + 1) * (r + 2) * (r + 3) / 6 - 1 : (r + 1) * (r + 2) / 2 - 1;
        if (hp.nCells && !hp.matrixCols.empty() && K != want)
            fail("nDvt " + std::to_string(K) + " does not match order " + std::to_string(r));
        if (hp.matrixCols.empty()) hp.nDvt = want;
    }
}


// ------------------------------------------------- constant/polyMesh reader ----
namespace {

struct FoamIn : In {
    using In::In;
    void skip() {  // whitespace and C/C++ comments
        for (;;) {
            skipWs();
            if (pos + 1 < n && p[pos] == '/' && p[pos + 1] == '/') {
                while (pos < n && p[pos] != '\n') ++pos;
            } else if (pos + 1 < n && p[pos] == '/' && p[pos + 1] == '*') {
                pos += 2;
                while (pos + 1 < n && !(p[pos] == '*' && p[pos + 1] == '/')) ++pos;
                pos = std::min(n, pos + 2);
            } else {
                return;
            }
        }
    }
    static bool punct(unsigned char c) { return c == '{' || c == '}' || c == '(' || c == ')' || c == ';'; }
    // next token: a punctuation character, a "quoted string" or a word
    std::string token() {
        skip();
        if (pos >= n) bad("unexpected end of file");
        if (punct(p[pos])) return std::string(1, (char)p[pos++]);
        size_t q = pos;
        if (p[q] == '"') {
            ++q;
            while (q < n && p[q] != '"') ++q;
            q = std::min(n, q + 1);
        } else {
            while (q < n && !punct(p[q]) && p[q] != ' ' && p[q] != '\n' && p[q] != '\t' && p[q] != '\r') ++q;
        }
        std::string t((const char*)p + pos, q - pos);
        pos = q;
        return t;
    }
    char peek() {
        skip();
        return pos < n ? (char)p[pos] : '\0';
    }
    // `key value ... ;` entries of a dictionary body up to the closing brace; sub-dictionaries
    // and lists inside values are kept as flat text
    std::map<std::string, std::string> dictBody() {
        std::map<std::string, std::string> d;
        for (;;) {
            std::string key = token();
            if (key == "}") return d;
            std::string val;
            int depth = 0;
            bool subDict = false;  // `key { ... }` ends at its brace, without a semicolon
            for (;;) {
                std::string t = token();
                if (t == ";" && depth == 0) break;
                if (t == "{" && depth == 0 && val.empty()) subDict = true;
                if (t == "(" || t == "{") ++depth;
                if (t == ")" || t == "}") {
                    if (depth == 0) bad("unbalanced dictionary entry '" + key + "'");
                    --depth;
                }
                if (!val.empty()) val += ' ';
                val += t;
                if (subDict && depth == 0) break;
            }
            d[key] = val;
        }
    }
    std::map<std::string, std::string> foamFileHeader() {
        if (token() != "FoamFile") bad("FoamFile header expected");
        if (token() != "{") bad("'{' expected after FoamFile");
        return dictBody();
    }
};

// labelList file (owner / neighbour)
void readFoamLabelList(const std::string& path, std::vector<int32_t>& out, std::string* note) {
    FoamIn in(path);
    auto hdr = in.foamFileHeader();
    if (note) *note = hdr.count("note") ? hdr["note"] : "";
    const bool binary = hdr.count("format") && hdr["format"] == "binary";
    in.skip();
    if (binary) {
        int lb = 0;
        if (hdr.count("arch")) {
            const std::string& a = hdr["arch"];
            if (a.find("label=64") != std::string::npos) lb = 8;
            if (a.find("label=32") != std::string::npos) lb = 4;
        }
        const size_t save = in.pos;
        for (int tryLb : {lb ? lb : 4, lb ? lb : 8}) {
            try {
                in.pos = save;
                in.labelBytes = tryLb;
                in.labelList(out);
                in.skip();
                return;
            } catch (const std::runtime_error&) {
                if (tryLb == (lb ? lb : 8)) throw;
            }
        }
    }
    const long long cnt = in.count();
    out.clear();
    if (in.peek() == '{') {  // uniform list N{v}
        in.expect('{');
        const long long v = in.integer();
        in.skip();
        in.expect('}');
        out.assign((size_t)cnt, (int32_t)v);
        return;
    }
    in.skip();
    in.expect('(');
    out.reserve((size_t)cnt);
    for (long long i = 0; i < cnt; ++i) {
        in.skip();
        out.push_back((int32_t)in.integer());
    }
    in.skip();
    in.expect(')');
}

}  // namespace

void readPolyMeshAddressing(const std::string& dir, HostPlan& hp) {
    std::string note;
    readFoamLabelList(dir + "/owner", hp.owner, &note);
    readFoamLabelList(dir + "/neighbour", hp.neighbour, nullptr);
    hp.nFaces = (int32_t)hp.owner.size();
    hp.nInternalFaces = (int32_t)hp.neighbour.size();
    if (hp.nInternalFaces > hp.nFaces) fail("polyMesh: more neighbours than faces");
    // nCells: from the header note OpenFOAM writes ("nPoints:.. nCells:.. nFaces:.."), else max + 1
    int32_t nCells = -1;
    const size_t at = note.find("nCells:");
    if (at != std::string::npos) nCells = (int32_t)atol(note.c_str() + at + 7);
    int32_t mx = -1;
    for (int32_t c : hp.owner) mx = std::max(mx, c);
    for (int32_t c : hp.neighbour) mx = std::max(mx, c);
    if (nCells < 0) nCells = mx + 1;
    if (mx >= nCells) fail("polyMesh: owner/neighbour refer to cells beyond nCells");
    for (int32_t c : hp.owner)
        if (c < 0) fail("polyMesh: negative owner");
    hp.nCells = nCells;
    // mesh.cells() as primitiveMesh::calcCells builds it: owned faces, then neighboured ones
    hp.cellFaceStart.assign((size_t)nCells + 1, 0);
    for (int f = 0; f < hp.nFaces; ++f) ++hp.cellFaceStart[hp.owner[f] + 1];
    for (int f = 0; f < hp.nInternalFaces; ++f) ++hp.cellFaceStart[hp.neighbour[f] + 1];
    for (int c = 0; c < nCells; ++c) hp.cellFaceStart[c + 1] += hp.cellFaceStart[c];
    hp.cellFaces.assign((size_t)hp.cellFaceStart[nCells], 0);
    {
        std::vector<int32_t> fill(hp.cellFaceStart.begin(), hp.cellFaceStart.end() - 1);
        for (int f = 0; f < hp.nFaces; ++f) hp.cellFaces[fill[hp.owner[f]]++] = f;
        for (int f = 0; f < hp.nInternalFaces; ++f) hp.cellFaces[fill[hp.neighbour[f]]++] = f;
    }
    // boundary: N ( name { ... } ... )
    FoamIn in(dir + "/boundary");
    in.foamFileHeader();
    in.skip();
    const long long nP = in.count();
    in.skip();
    in.expect('(');
    hp.patches.clear();
    std::vector<std::string> nbrName;
    for (long long i = 0; i < nP; ++i) {
        Patch pt;
        pt.name = in.token();
        if (in.token() != "{") in.bad("'{' expected after patch name " + pt.name);
        auto d = in.dictBody();
        if (!d.count("nFaces") || !d.count("startFace")) in.bad("patch " + pt.name + " lacks nFaces/startFace");
        pt.size = (int32_t)atol(d["nFaces"].c_str());
        pt.start = (int32_t)atol(d["startFace"].c_str());
        const std::string type = d.count("type") ? d["type"] : "patch";
        std::string nb;
        if (type == "processor" || type == "processorCyclic") {
            pt.type = PATCH_PROCESSOR;
            if (!d.count("neighbProcNo")) in.bad("processor patch " + pt.name + " lacks neighbProcNo");
            pt.nbr = (int32_t)atol(d["neighbProcNo"].c_str());
        } else if (type == "cyclic") {
            pt.type = PATCH_CYCLIC;
            if (!d.count("neighbourPatch")) in.bad("cyclic patch " + pt.name + " lacks neighbourPatch");
            nb = d["neighbourPatch"];
        } else if (type == "empty") {
            pt.type = PATCH_EMPTY;
        } else {
            pt.type = PATCH_GENERIC;
        }
        if (pt.size < 0 || pt.start < hp.nInternalFaces || pt.start + pt.size > hp.nFaces)
            in.bad("patch " + pt.name + " lies outside the boundary faces");
        hp.patches.push_back(pt);
        nbrName.push_back(nb);
    }
    in.skip();
    in.expect(')');
    for (size_t i = 0; i < hp.patches.size(); ++i) {
        if (hp.patches[i].type != PATCH_CYCLIC) continue;
        for (size_t j = 0; j < hp.patches.size(); ++j)
            if (hp.patches[j].name == nbrName[i]) hp.patches[i].nbr = (int32_t)j;
        if (hp.patches[i].nbr < 0) fail("polyMesh: cyclic patch " + hp.patches[i].name + " has no partner " + nbrName[i]);
    }
}

}  // namespace wenoext
