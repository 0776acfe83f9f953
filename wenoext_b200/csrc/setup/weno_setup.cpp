// Restatement of WENOBase's preprocessing on a PolyMesh. See weno_setup.h.
#include "weno_setup.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <numeric>
#include <string>
#include <stdexcept>
#include <thread>
#include <chrono>
#include <cstdio>
#include <cstdlib>

#include "geometry_weno.h"
#include "pinv_device.h"
#include "svd.h"
#include "wenobase_cache.h"

namespace wenoext {

namespace {
using detail::CellSetup;

inline double ofSign(double s) { return s >= 0 ? 1.0 : -1.0; }  // OpenFOAM sign()

double factorial(int n) {
    double f = 1;
    for (int i = 2; i <= n; ++i) f *= i;
    return f;
}
inline double intPow(double b, unsigned e) {
    double t = 1.0;
    for (unsigned i = 0; i < e; ++i) t *= b;
    return t;
}

// WENOBase::calcGeom (WENOBase.C:600-643)
double calcGeom(const Vec3& x_ij, int n, int m, int o, const List3D& volMomJ,
                const List3D& volMomI) {
    double g = 0.0;
    for (int k = 0; k <= n; ++k)
        for (int l = 0; l <= m; ++l)
            for (int j = 0; j <= o; ++j)
                g += factorial(n) / (factorial(k) * factorial(n - k)) * factorial(m) /
                     (factorial(l) * factorial(m - l)) * factorial(o) /
                     (factorial(j) * factorial(o - j)) * intPow(x_ij.x, k) * intPow(x_ij.y, l) *
                     intPow(x_ij.z, j) * volMomJ(n - k, m - l, o - j);
    return g - volMomI(n, m, o);
}

// WENOBase::createStencilID + extendStencils + sortStencil for one cell
// (WENOBase.C:941-1027, :199-240, :244-331)
void collectCentral(const PolyMesh& gm, int g, int nDvt, int nStencils, double extendRatio,
                    const Mat3& JInv, const Vec3& refPoint, std::vector<int32_t>& central) {
    central.clear();
    central.push_back(g);
    for (int i = gm.cellCellStart[g]; i < gm.cellCellStart[g + 1]; ++i)
        central.push_back(gm.cellCells[i]);
    std::map<int, bool> stencilCells;
    stencilCells.insert({central[0], false});
    for (int32_t c : central) stencilCells.insert({c, true});
    size_t minStencilSize = 0;
    int iter = 0;
    while (minStencilSize < 1.2 * extendRatio * nDvt * nStencils) {
        std::map<int, bool> temp = stencilCells;
        for (const auto& it : stencilCells) {
            if (!it.second) continue;
            const int ci = it.first;
            for (int i = gm.cellCellStart[ci]; i < gm.cellCellStart[ci + 1]; ++i) {
                int nb = gm.cellCells[i];
                if (temp.find(nb) == temp.end()) {
                    central.push_back(nb);
                    temp.insert({nb, true});
                }
            }
            temp[ci] = false;
        }
        stencilCells.swap(temp);
        minStencilSize = central.size();
        if (++iter > 100) break;  // maxIter, WENOBase.C:996-1011
    }
    // sortStencil: stable sort by reference-space distance after the first 1.2*K
    const size_t n = central.size();
    std::vector<double> dist(n);
    const Vec3 tI = geom::transformPoint(JInv, gm.cellCentres[central[0]], refPoint);
    for (size_t i = 0; i < n; ++i)
        dist[i] = mag(geom::transformPoint(JInv, gm.cellCentres[central[i]], refPoint) - tI);
    std::vector<int> idx(n);
    std::iota(idx.begin(), idx.end(), 0);
    size_t first = std::min(n, (size_t)(1.2 * nDvt));
    std::vector<int32_t> copy(central);
    std::stable_sort(idx.begin() + first, idx.end(), [&](int i1, int i2) {
        if (std::fabs(dist[i1] - dist[i2]) < 1e-9) {
            const Vec3& p1 = gm.cellCentres[copy[i1]];
            const Vec3& p2 = gm.cellCentres[copy[i2]];
            if (std::fabs(p1.x - p2.x) < 1e-9)
                return std::fabs(p1.y - p2.y) < 1e-9 ? p1.z < p2.z : p1.y < p2.y;
            return p1.x < p2.x;
        }
        return dist[i1] < dist[i2];
    });
    for (size_t i = 1; i < n; ++i) central[i] = copy[idx[i]];
}

// WENOBase::splitStencil (WENOBase.C:43-196) in global ids
void splitStencil(const PolyMesh& gm, int g, int nDvt, double extendRatio, CellSetup& cs) {
    const int nf = gm.cellNFaces(g);
    const int32_t* fl = gm.cellFaceList(g);
    std::vector<std::vector<Mat3>> JQ;
    std::vector<std::array<int32_t, 3>> tris;
    const Vec3& C = gm.cellCentres[g];
    for (int i = 0; i < nf; ++i) {
        if (fl[i] >= gm.nInternalFaces) continue;
        gm.faceTriangles(fl[i], tris);
        std::vector<Mat3> q;
        for (const auto& t : tris) {
            Mat3 J;
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) J(r, c) = gm.points[t[c]][r] - C[r];
            q.push_back(geom::inv3(J));
        }
        JQ.push_back(std::move(q));
    }
    const int nSect = (int)JQ.size();
    cs.globalID.assign(nSect + 1, std::vector<int32_t>{(int32_t)g});
    cs.globalID[0] = cs.central;
    for (size_t cj = 1; cj < cs.central.size() && nSect > 0; ++cj) {
        bool attached = false;
        int actualFace = 0;
        while (!attached) {
            if (cs.globalID[actualFace + 1].size() < extendRatio * nDvt) {
                for (const Mat3& Ji : JQ[actualFace]) {
                    Vec3 t = geom::transformPoint(Ji, gm.cellCentres[cs.central[cj]], C);
                    if (ofSign(t.x) > 0 && ofSign(t.y) > 0 && ofSign(t.z) > 0) {
                        cs.globalID[actualFace + 1].push_back(cs.central[cj]);
                        attached = true;
                        break;
                    }
                }
            }
            if (actualFace < nSect - 1)
                actualFace++;
            else
                break;
        }
    }
    const size_t necSize = (size_t)std::floor(extendRatio * nDvt);
    for (auto& st : cs.globalID) {
        if (st.size() >= necSize)
            st.resize(necSize);
        else
            st.assign(1, CELL_DELETED);
    }
}

// WENOBase::calcMatrix (WENOBase.C:410-597) for one stencil. `rowCache` memoises
// the LS row of (cell I, cell J), which does not depend on the stencil.
void calcMatrix(const PolyMesh& gm, int g, int stencilI, int nDvt, const int dim[3],
                const SetupParams& prm, CellSetup& cs,
                std::map<int32_t, std::vector<double>>& rowCache) {
    std::vector<int32_t>& st = cs.globalID[stencilI];
    const int stencilSize = (int)st.size();
    int nCells = stencilSize - 1;
    const int P = prm.polOrder;
    if (prm.gpuPinvDevice >= 0 && !prm.bestConditioned) {
        // deferred: rows, matrices and pseudo-inverses of all stencils of the mesh are formed in one device
        // batch (finishPending -> wenoext_ls_pinv_batched)
        cs.Ainv[stencilI].clear();
        cs.pending[stencilI] = 1;
        return;
    }
    std::vector<double> A((size_t)nCells * nDvt, 0.0);
    const Vec3 tI = geom::transformPoint(cs.JInv, gm.cellCentres[g], cs.refPoint);
    List3D transVolMom;
    for (int cj = 1; cj <= nCells; ++cj) {
        auto it = rowCache.find(st[cj]);
        if (it == rowCache.end()) {
            const Vec3 tJ = geom::transformPoint(cs.JInv, gm.cellCentres[st[cj]], cs.refPoint);
            geom::transformIntegral(gm, st[cj], tJ, P, cs.JInv, cs.refPoint, cs.refDet,
                                    transVolMom);
            std::vector<double> row;
            for (int n = 0; n <= dim[0]; ++n)
                for (int m = 0; m <= dim[1]; ++m)
                    for (int l = 0; l <= dim[2]; ++l)
                        if ((n + m + l) <= P && (n + m + l) > 0) {
                            double v = calcGeom(tJ - tI, n, m, l, transVolMom, cs.volI);
                            row.push_back(std::fabs(v) < 1e-9 ? 0.0 : v);  // addCoeffs :646-670
                        }
            it = rowCache.emplace(st[cj], std::move(row)).first;
        }
        std::copy(it->second.begin(), it->second.end(), A.begin() + (size_t)(cj - 1) * nDvt);
    }
    SvdResult cur = svd(A, nCells, nDvt, prm.maxCondition);
    if (prm.bestConditioned) {
        bool haveBest = false;
        SvdResult best;
        for (nCells = nDvt + 1; nCells < stencilSize; ++nCells) {
            std::vector<double> As(A.begin(), A.begin() + (size_t)nCells * nDvt);
            cur = svd(As, nCells, nDvt, prm.maxCondition);
            if (!cur.converged) continue;
            if (!haveBest || cur.nZeros < best.nZeros ||
                (cur.nZeros == best.nZeros && cur.condition() < best.condition())) {
                best = cur;
                haveBest = true;
            }
        }
        if (haveBest) cur = best;
    }
    if (!cur.converged) throw std::runtime_error("calcMatrix: could not calculate SVD");
    if (prm.checkCondition && stencilI == 0 && cur.nZeros != 0) {
        st.assign(1, CELL_DELETED);
        cs.empty = true;  // stencilsID_[cell][0][0] = Cell::empty
        cs.Ainv[stencilI].clear();
        return;
    }
    cs.Ainv[stencilI] = cur.pseudoInverse();  // nDvt x cur.m
    if (cur.m != stencilSize - 1) st.resize(cur.m + 1);
}

}  // namespace

CellSetup detail::computeCell(const PolyMesh& gm, int g, int nDvt, const int dim[3],
                              const SetupParams& prm) {
    CellSetup cs;
    geom::initIntegrals(gm, g, prm.polOrder, cs.volI, cs.JInv, cs.refPoint, cs.refDet);
    int nStencils = 1;
    const int nf = gm.cellNFaces(g);
    const int32_t* fl = gm.cellFaceList(g);
    for (int i = 0; i < nf; ++i)
        if (fl[i] < gm.nInternalFaces) nStencils++;
    collectCentral(gm, g, nDvt, nStencils, prm.extendRatio, cs.JInv, cs.refPoint, cs.central);
    splitStencil(gm, g, nDvt, prm.extendRatio, cs);
    cs.Ainv.resize(cs.globalID.size());
    cs.pending.assign(cs.globalID.size(), 0);
    std::map<int32_t, std::vector<double>> rowCache;
    for (size_t s = 0; s < cs.globalID.size(); ++s)
        if (cs.globalID[s][0] != CELL_DELETED) calcMatrix(gm, g, (int)s, nDvt, dim, prm, cs, rowCache);
    // LSMatrixCheck (WENOBase.C:1327-1347)
    int valid = 0;
    for (const auto& st : cs.globalID)
        if (st[0] != CELL_DELETED) valid++;
    if (prm.checkCondition && valid == 0) cs.empty = true;
    return cs;
}

// matrixDB stand-in (libWENOEXT/WENOBase/matrixDB.C:75-199): tolerance-based
// de-duplication, same key, same search window, same comparison.
void detail::finishPending(std::vector<CellSetup>& cells, const PolyMesh& gm, const std::vector<int32_t>& cellGlobal,
                           int nDvt, const int dim[3], const SetupParams& prm) {
    // source mesh: points, fan triangles of every cell's faces (geom::cellTriangleList), centres, volumes
    LsBatchHost b;
    b.polOrder = prm.polOrder;
    b.nDvt = nDvt;
    for (int k = 0; k < 3; ++k) b.dim[k] = dim[k];
    b.points.resize((size_t)gm.points.size() * 3);
    for (size_t p = 0; p < gm.points.size(); ++p)
        for (int k = 0; k < 3; ++k) b.points[p * 3 + k] = gm.points[p][k];
    b.triStart.assign(1, 0);
    b.cellCentres.resize((size_t)gm.nCells * 3);
    std::vector<std::array<int32_t, 3>> tris;
    for (int c = 0; c < gm.nCells; ++c) {
        geom::cellTriangleList(gm, c, tris);
        for (const auto& t : tris) b.triPoints.insert(b.triPoints.end(), t.begin(), t.end());
        b.triStart.push_back((int64_t)b.triPoints.size() / 3);
        for (int k = 0; k < 3; ++k) b.cellCentres[(size_t)c * 3 + k] = gm.cellCentres[c][k];
    }
    b.cellVolumes = gm.cellVolumes;
    // one frame per cell with pending stencils; one pair per distinct stencil cell of that cell (the row of
    // (I, J) does not depend on the stencil); one matrix per pending stencil
    const int cube = (prm.polOrder + 1) * (prm.polOrder + 1) * (prm.polOrder + 1);
    std::vector<std::pair<int32_t, int32_t>> where;
    b.start.assign(1, 0);
    for (size_t c = 0; c < cells.size(); ++c) {
        const CellSetup& cs = cells[c];
        bool any = false;
        for (uint8_t pnd : cs.pending) any = any || pnd;
        if (!any) continue;
        const int32_t f = (int32_t)b.frameCell.size();
        b.frameCell.push_back(cellGlobal[c]);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) b.JInv.push_back(cs.JInv(i, j));
        for (int k = 0; k < 3; ++k) b.refPoint.push_back(cs.refPoint[k]);
        b.refDet.push_back(cs.refDet);
        if ((int)cs.volI.d.size() != cube) throw std::runtime_error("finishPending: volume moments of unexpected size");
        b.volI.insert(b.volI.end(), cs.volI.d.begin(), cs.volI.d.end());
        std::map<int32_t, int64_t> pairOf;
        for (size_t s = 0; s < cs.pending.size(); ++s) {
            if (!cs.pending[s]) continue;
            const auto& st = cs.globalID[s];
            for (size_t cj = 1; cj < st.size(); ++cj) {
                auto it = pairOf.find(st[cj]);
                if (it == pairOf.end()) {
                    it = pairOf.emplace(st[cj], (int64_t)b.pairCell.size()).first;
                    b.pairFrame.push_back(f);
                    b.pairCell.push_back(st[cj]);
                }
                b.rowPair.push_back(it->second);
            }
            b.rows.push_back((int32_t)st.size() - 1);
            b.start.push_back(b.start.back() + (int64_t)(st.size() - 1) * nDvt);
            where.push_back({(int32_t)c, (int32_t)s});
        }
    }
    if (where.empty()) return;
    const int32_t nMat = (int32_t)where.size();
    const std::vector<int64_t>& start = b.start;
    std::vector<double> P((size_t)start.back());
    std::vector<int32_t> nZeros(nMat), conv(nMat);
    std::string err;
    if (lsPinvBatchedDevice(prm.gpuPinvDevice, b, prm.maxCondition, P.data(), nZeros.data(), nullptr, conv.data(), err))
        throw std::runtime_error(err);
    for (int32_t i = 0; i < nMat; ++i) {
        CellSetup& cs = cells[where[i].first];
        const int s = where[i].second;
        cs.pending[s] = 0;
        // the decisions calcMatrix takes after the SVD (WENOBase.C:527-597)
        if (!conv[i]) throw std::runtime_error("calcMatrix: could not calculate SVD");
        if (prm.checkCondition && s == 0 && nZeros[i] != 0) {
            cs.globalID[s].assign(1, CELL_DELETED);
            cs.empty = true;
            cs.Ainv[s].clear();
            continue;
        }
        cs.Ainv[s].assign(P.begin() + start[i], P.begin() + start[i + 1]);
    }
    // LSMatrixCheck (WENOBase.C:1327-1347)
    for (CellSetup& cs : cells) {
        int valid = 0;
        for (const auto& st : cs.globalID)
            if (st[0] != CELL_DELETED) valid++;
        if (prm.checkCondition && valid == 0) cs.empty = true;
    }
}

namespace {
double maxMag(const std::vector<double>& A) {
    double m = 0;
    for (double v : A) m = std::max(m, std::fabs(v));
    return m;
}
int32_t hashMatrix(const std::vector<double>& A) { return matrixKey(A.data(), A.size()); }
}  // namespace

int32_t detail::MatrixPool::insertNew(int32_t key, const std::vector<double>& A) {
    int32_t id = (int32_t)plan->matrixCols.size();
    plan->matrixCols.push_back((int32_t)(A.size() / nDvt));
    plan->matrixData.insert(plan->matrixData.end(), A.begin(), A.end());
    plan->matrixStart.push_back((int64_t)plan->matrixData.size());
    db.emplace(key, id);
    return id;
}
int32_t detail::MatrixPool::similar(const std::vector<double>& A) {
    const int32_t key = hashMatrix(A);
    if (db.empty()) return insertNew(key, A);
    const double tol = 1e-9 * maxMag(A);
    auto itStart = db.lower_bound(key - 5);
    auto itEnd = db.lower_bound(key + 5);
    if (itStart == itEnd && itStart != db.begin()) --itStart;
    for (auto it = itStart; it != itEnd; ++it) {
        const int32_t id = it->second;
        const size_t sz = (size_t)(plan->matrixStart[id + 1] - plan->matrixStart[id]);
        if (sz != A.size()) continue;
        const double* cmp = &plan->matrixData[plan->matrixStart[id]];
        bool ok = true;
        for (size_t i = 0; i < sz; ++i) {
            if (std::fabs(A[i]) < 1e-15) continue;
            if (std::fabs(cmp[i] - A[i]) > tol) {
                ok = false;
                break;
            }
        }
        if (ok) return id;
    }
    return insertNew(key, A);
}

int detail::computeNDvt(const PolyMesh& m, int polOrder, int dim[3]) {
    // setDegreeOfFreedom (WENOBase.C:1130-1155)
    int nDvt;
    if (m.nSolutionD() == 3)
        nDvt = (int)((polOrder + 1.0) * (polOrder + 2.0) * (polOrder + 3.0) / 6.0 - 1.0);
    else
        nDvt = (int)((polOrder + 1.0) * (polOrder + 2.0) / 2.0 - 1.0);
    for (int i = 0; i < 3; ++i) dim[i] = m.solutionD[i] == 1 ? polOrder : 0;
    return nDvt;
}

namespace {
using detail::MatrixPool;
using detail::computeCell;
using detail::computeNDvt;
void copyMeshAddressing(const PolyMesh& lm, HostPlan& hp) {
    hp.nCells = lm.nCells;
    hp.nFaces = lm.nFaces();
    hp.nInternalFaces = lm.nInternalFaces;
    hp.owner = lm.owner;
    hp.neighbour = lm.neighbour;
    hp.cellFaceStart = lm.cellFaceStart;
    hp.cellFaces = lm.cellFaces;
    hp.patches = lm.patches;
    hp.cellCentres.resize((size_t)lm.nCells * 3);
    for (int c = 0; c < lm.nCells; ++c)
        for (int d = 0; d < 3; ++d) hp.cellCentres[(size_t)c * 3 + d] = lm.cellCentres[c][d];
    hp.faceAreaVec.resize((size_t)lm.nFaces() * 3);
    for (int f = 0; f < lm.nFaces(); ++f)
        for (int d = 0; d < 3; ++d) hp.faceAreaVec[(size_t)f * 3 + d] = lm.faceAreas[f][d];
}

}  // namespace

HostPlan buildPlanGeneric(const PolyMesh& gm, const std::vector<int32_t>& cellProcIn,
                          int32_t nProcs, int32_t myProc, const SetupParams& prm) {
    HostPlan hp;
    hp.polOrder = prm.polOrder;
    hp.nDvt = computeNDvt(gm, prm.polOrder, hp.dim);
    hp.nProcs = nProcs;
    hp.myProc = myProc;
    const int K = hp.nDvt;
    const bool parallel = nProcs > 1;

    // local mesh (globalfvMesh: local <-> global maps, globalfvMesh.H:112-136)
    std::vector<int32_t> cellProc = cellProcIn;
    if (cellProc.empty()) cellProc.assign(gm.nCells, myProc);
    std::vector<int32_t> g2l;
    LocalMesh L = extractLocalMesh(gm, cellProc, myProc, g2l);
    const PolyMesh& lm = L.mesh;
    copyMeshAddressing(lm, hp);
    // processor-local id of every global cell on its owner rank
    std::vector<int32_t> procLocalId(gm.nCells);
    {
        std::vector<int32_t> cnt(nProcs, 0);
        for (int c = 0; c < gm.nCells; ++c) procLocalId[c] = cnt[cellProc[c]]++;
    }

    const bool timing = getenv("WENOEXT_SETUP_TIMING") != nullptr;
    auto t0 = std::chrono::steady_clock::now();
    auto mark = [&](const char* what) {
        if (!timing) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[wenoext setup] %-28s %8.3f s\n", what, std::chrono::duration<double>(t1 - t0).count());
        t0 = t1;
    };
    mark("local mesh");
    // steps 1-4 per local cell, threaded; inserted into the pool in cell order
    const int nLocal = lm.nCells;
    std::vector<CellSetup> cells(nLocal);
    int nThreads = prm.nThreads > 0 ? prm.nThreads : (int)std::thread::hardware_concurrency();
    nThreads = std::max(1, std::min(nThreads, nLocal));
    {
        std::vector<std::thread> th;
        std::vector<std::string> err(nThreads);
        for (int t = 0; t < nThreads; ++t)
            th.emplace_back([&, t]() {
                try {
                    for (int c = t; c < nLocal; c += nThreads)
                        cells[c] = computeCell(gm, L.localToGlobalCell[c], K, hp.dim, prm);
                } catch (const std::exception& e) {
                    err[t] = e.what();
                }
            });
        for (auto& x : th) x.join();
        for (auto& e : err)
            if (!e.empty()) throw std::runtime_error(e);
    }

    mark(prm.gpuPinvDevice >= 0 ? "stencils" : "stencils + pseudo-inverses");
    if (prm.gpuPinvDevice >= 0) {
        detail::finishPending(cells, gm, L.localToGlobalCell, K, hp.dim, prm);
        mark("LS rows + pseudo-inverses (device batch)");
    }
    // correctParallelRun (WENOBase.C:1030-1126): halo ids in order of first
    // appearance over the (uncut) central lists
    hp.receiveProcList.assign(nProcs, -1);
    hp.sendProcList.assign(nProcs, -1);
    hp.receiveHaloSize.assign(nProcs, 0);
    hp.sendHaloCellIDList.assign(nProcs, {});
    hp.requestHaloCellID.assign(nProcs, {});
    std::vector<std::map<int32_t, int32_t>> newHaloID(nProcs);
    if (parallel)
        for (int c = 0; c < nLocal; ++c)
            for (int32_t gc : cells[c].central) {
                const int p = cellProc[gc];
                if (p == myProc) continue;
                hp.receiveProcList[p] = p;
                auto& mp = newHaloID[p];
                if (mp.find(gc) == mp.end()) {
                    int32_t id = (int32_t)mp.size();
                    mp.emplace(gc, id);
                    hp.requestHaloCellID[p].push_back(procLocalId[gc]);
                }
            }
    for (int p = 0; p < nProcs; ++p) hp.receiveHaloSize[p] = (int32_t)hp.requestHaloCellID[p].size();

    // patchNeighbourField addressing for the bounded limiter (WENOUpwindFit.C:130)
    {
        std::vector<int64_t> hoff(nProcs + 1, 0);
        for (int p = 0; p < nProcs; ++p) hoff[p + 1] = hoff[p] + hp.receiveHaloSize[p];
        const int nBnd = lm.nFaces() - lm.nInternalFaces;
        hp.patchNbrCell.assign(std::max(nBnd, 1), -1);
        for (const Patch& pt : lm.patches) {
            for (int i = 0; i < pt.size; ++i) {
                const int f = pt.start + i;
                if (pt.type == PATCH_CYCLIC) {
                    hp.patchNbrCell[f - lm.nInternalFaces] = lm.owner[lm.patches[pt.nbr].start + i];
                } else if (pt.type == PATCH_PROCESSOR) {
                    const int gf = L.localToGlobalFace[f];
                    const int go = gm.owner[gf], gn = gm.neighbour[gf];
                    const int remote = cellProc[go] == myProc ? gn : go;
                    const int p = cellProc[remote];
                    hp.patchNbrCell[f - lm.nInternalFaces] =
                        (int32_t)(nLocal + hoff[p] + newHaloID[p].at(remote));
                }
            }
        }
    }

    // flatten stencils + matrices
    MatrixPool pool;
    pool.plan = &hp;
    pool.nDvt = K;
    hp.matrixStart.assign(1, 0);
    hp.cellStencilStart.assign(1, 0);
    hp.stencilEntryStart.assign(1, 0);
    for (int c = 0; c < nLocal; ++c) {
        const CellSetup& cs = cells[c];
        for (size_t s = 0; s < cs.globalID.size(); ++s) {
            const auto& st = cs.globalID[s];
            if (st[0] == CELL_DELETED) {
                hp.stencilsID.push_back((s == 0 && cs.empty) ? CELL_EMPTY : CELL_DELETED);
                hp.cellToProcMap.push_back(CELL_DELETED);
                hp.stencilMatrix.push_back(-1);
            } else {
                for (size_t j = 0; j < st.size(); ++j) {
                    const int32_t gc = st[j];
                    const int p = cellProc[gc];
                    if (p == myProc) {
                        hp.stencilsID.push_back(g2l[gc]);
                        hp.cellToProcMap.push_back(CELL_LOCAL);
                    } else {
                        hp.stencilsID.push_back(newHaloID[p].at(gc));
                        hp.cellToProcMap.push_back(p);
                    }
                }
                if (s == 0 && cs.empty) hp.stencilsID[hp.stencilEntryStart.back()] = CELL_EMPTY;
                hp.stencilMatrix.push_back(pool.similar(cs.Ainv[s]));
                hp.nMatricesTotal++;
            }
            hp.stencilEntryStart.push_back((int64_t)hp.stencilsID.size());
        }
        hp.cellStencilStart.push_back((int64_t)hp.stencilMatrix.size());
    }

    mark("flatten + matrixDB");
    // step 5: B, intBasTrans, refFacAr on the local mesh (WENOBase.C:820-854); cells are
    // independent: a cell writes its own B, the side of each face it sits on, and refFacAr of
    // the faces it owns
    hp.B.assign((size_t)nLocal * K * K, 0.0);
    const int s1 = prm.polOrder + 1;
    hp.intBasTrans.assign((size_t)lm.nFaces() * 2 * s1 * s1 * s1, 0.0);
    hp.refFacAr.assign(lm.nFaces(), 0.0);
    {
        std::vector<std::thread> th;
        std::vector<std::string> err(nThreads);
        for (int t = 0; t < nThreads; ++t)
            th.emplace_back([&, t]() {
                try {
                    for (int c = t; c < nLocal; c += nThreads) {
                        const CellSetup& cs = cells[c];
                        std::vector<double> B = geom::getB(lm, c, prm.polOrder, K, cs.JInv, cs.refPoint, hp.dim);
                        std::copy(B.begin(), B.end(), hp.B.begin() + (size_t)c * K * K);
                        geom::surfIntTransCell(lm, c, prm.polOrder, cs.volI, cs.JInv, cs.refPoint,
                                               hp.intBasTrans.data(), hp.refFacAr.data());
                    }
                } catch (const std::exception& e) {
                    err[t] = e.what();
                }
            });
        for (auto& x : th) x.join();
        for (auto& e : err)
            if (!e.empty()) throw std::runtime_error(e);
    }
    mark("B + face integrals");
    return hp;
}

void setSendHalo(HostPlan& plan, int32_t from, const int32_t* ids, int64_t n) {
    plan.sendHaloCellIDList[from].assign(ids, ids + n);
    plan.sendProcList[from] = n > 0 ? from : -1;
}

}  // namespace wenoext
