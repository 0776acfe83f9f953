// Fast WENOBase setup for structured hex boxes (see box_fast.h).
//
// On a uniform box all cells are congruent, so the result of the reference's
// per-cell preprocessing (stencil lists as index offsets, pseudo-inverses, B,
// face integrals) depends only on where the cell sits relative to the box sides
// within the reach of its stencils. Each such CLASS is computed once with the
// generic restatement (weno_setup.cpp, which cites WENOBase.C step by step) on a
// small representative box with the same cell size and side types, and then
// instantiated for every cell of the class. This is what the reference's
// matrixDB de-duplication (matrixDB.C:75-150) converges to on such meshes: one
// representative matrix per class.
#include "box_fast.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <stdexcept>
#include <string>
#include <thread>

#include "geometry_weno.h"

namespace wenoext {

namespace {

using detail::CellSetup;

struct Axis {
    int n = 1;       // cells of the big box
    int m = 1;       // cells of the representative box
    int R = 0;       // clipping distance
    bool cyclic = false;
    // big index -> representative index
    int clip(int i) const {
        if (n == m) return i;
        if (cyclic) return R;
        if (i < R) return i;
        if (n - 1 - i < R) return m - 1 - (n - 1 - i);
        return R;
    }
    // neighbour index in the big box given the cell's big index and both
    // representative indices; -1 cannot happen for cells produced by the setup
    int apply(int iBig, int cSmall, int nbSmall) const {
        if (n == m) return nbSmall;
        int d = nbSmall - cSmall;
        if (cyclic) {
            if (d > m / 2) d -= m;
            if (d < -m / 2) d += m;
            int r = (iBig + d) % n;
            return r < 0 ? r + n : r;
        }
        return iBig + d;
    }
};

template <class F>
void parallelFor(int n, int nThreads, F&& f) {
    nThreads = std::max(1, std::min(nThreads, n));
    std::vector<std::thread> th;
    std::vector<std::string> err(nThreads);
    for (int t = 0; t < nThreads; ++t)
        th.emplace_back([&, t]() {
            try {
                const int64_t a = (int64_t)n * t / nThreads, b = (int64_t)n * (t + 1) / nThreads;
                for (int i = (int)a; i < (int)b; ++i) f(i);
            } catch (const std::exception& e) {
                err[t] = e.what();
            }
        });
    for (auto& x : th) x.join();
    for (auto& e : err)
        if (!e.empty()) throw std::runtime_error(e);
}

// per representative cell: everything the plan needs, in representative ids
struct ClassData {
    CellSetup cs;
    std::vector<double> B;                 // K x K
    std::array<std::vector<double>, 6> I;  // basis integrals per face direction (cube)
    std::array<double, 6> area{};          // reference-space face area per direction
    std::array<int, 6> hasFace{};
    std::vector<int32_t> poolId;           // per stencil: matrix pool id (-1 deleted), filled lazily
    bool pooled = false;
};

// face direction ids: 0 -x, 1 +x, 2 -y, 3 +y, 4 -z, 5 +z
int faceDirection(const PolyMesh& sm, const int mdim[3], int cell, int f) {
    if (f < sm.nInternalFaces) {
        const int other = sm.owner[f] == cell ? sm.neighbour[f] : sm.owner[f];
        const int d = other - cell, ad = d < 0 ? -d : d;
        // cells are numbered i-fastest; an axis with a single cell has no internal faces
        if (mdim[2] > 1 && ad == mdim[0] * mdim[1]) return d > 0 ? 5 : 4;
        if (mdim[1] > 1 && ad == mdim[0]) return d > 0 ? 3 : 2;
        if (mdim[0] > 1 && ad == 1) return d > 0 ? 1 : 0;
        throw std::runtime_error("box_fast: unexpected internal face");
    }
    for (size_t p = 0; p < sm.patches.size(); ++p)
        if (f >= sm.patches[p].start && f < sm.patches[p].start + sm.patches[p].size) return (int)p;
    throw std::runtime_error("box_fast: face without a patch");
}

}  // namespace

HostPlan buildPlanBoxFast(const BoxSpec& spec, const int32_t nProc[3], int32_t myProc,
                          const SetupParams& prm) {
    if (spec.tetSplit) throw std::runtime_error("box_fast: hex boxes only");
    const int nProcs = nProc[0] * nProc[1] * nProc[2];
    for (int a = 0; a < 3; ++a)
        if (spec.sideType[2 * a] == PATCH_CYCLIC && nProc[a] > 1)
            throw std::runtime_error("box_fast: decomposition across a cyclic direction is not supported");
    int nThreads = prm.nThreads > 0 ? prm.nThreads : (int)std::thread::hardware_concurrency();
    nThreads = std::max(1, nThreads);

    // ---- representative box and its classes (retry with a larger reach if needed) ----
    Axis ax[3];
    PolyMesh sm;
    std::vector<ClassData> cls;
    int K = 0, dim[3] = {0, 0, 0};
    int mdim[3] = {1, 1, 1};
    for (int R = prm.polOrder <= 2 ? 8 : 12;; R += 4) {
        BoxSpec ss = spec;
        for (int a = 0; a < 3; ++a) {
            ax[a].n = spec.n[a];
            ax[a].R = R;
            // cyclic sides are boundary faces for the stencil collection (cellCells()
            // does not cross them), so they clip like any other side
            ax[a].cyclic = false;
            ax[a].m = std::min(spec.n[a], 2 * R + 1);
            mdim[a] = ax[a].m;
            ss.n[a] = ax[a].m;
            ss.nDen[a] = spec.n[a];  // same point coordinates as the first layers of the big box
        }
        sm = makeBoxMesh(ss);
        K = detail::computeNDvt(sm, prm.polOrder, dim);
        cls.assign(sm.nCells, ClassData());
        SetupParams prmHost = prm;  // a few hundred class cells: not worth a device batch
        prmHost.gpuPinvDevice = -1;
        parallelFor(sm.nCells, nThreads,
                    [&](int c) { cls[c].cs = detail::computeCell(sm, c, K, dim, prmHost); });
        // the stencils of a clipped axis must stay strictly inside the reach
        bool ok = true;
        for (int c = 0; c < sm.nCells && ok; ++c) {
            const int ci[3] = {c % mdim[0], (c / mdim[0]) % mdim[1], c / (mdim[0] * mdim[1])};
            for (int32_t g : cls[c].cs.central) {
                const int gi[3] = {g % mdim[0], (g / mdim[0]) % mdim[1], g / (mdim[0] * mdim[1])};
                for (int a = 0; a < 3; ++a) {
                    if (ax[a].n == ax[a].m) continue;
                    int d = gi[a] - ci[a];
                    if (ax[a].cyclic) {
                        if (d > ax[a].m / 2) d -= ax[a].m;
                        if (d < -ax[a].m / 2) d += ax[a].m;
                    }
                    if (std::abs(d) > R - 1) ok = false;
                }
            }
        }
        if (ok) break;
        if (R > 64) throw std::runtime_error("box_fast: stencil reach does not converge");
    }
    // B and face integrals of every class (geometry of the representative box)
    const int s1 = prm.polOrder + 1, cube = s1 * s1 * s1;
    parallelFor(sm.nCells, nThreads, [&](int c) {
        ClassData& cd = cls[c];
        cd.B = geom::getB(sm, c, prm.polOrder, K, cd.cs.JInv, cd.cs.refPoint, dim);
        const int nf = sm.cellNFaces(c);
        const int32_t* fl = sm.cellFaceList(c);
        for (int i = 0; i < nf; ++i) {
            const int dir = faceDirection(sm, mdim, c, fl[i]);
            cd.I[dir].assign(cube, 0.0);
            cd.hasFace[dir] = 1;
            geom::surfIntTransFace(sm, c, fl[i], prm.polOrder, cd.cs.volI, cd.cs.JInv, cd.cs.refPoint,
                                   cd.I[dir].data(), cd.area[dir]);
        }
    });

    // ---- block decomposition (decomposePar simple): per axis cell -> (rank, local) ----
    std::vector<int> axProc[3], axLocal[3];
    int bs[3], b0[3], pc[3] = {myProc % nProc[0], (myProc / nProc[0]) % nProc[1],
                               myProc / (nProc[0] * nProc[1])};
    std::vector<int> axCount[3];
    for (int a = 0; a < 3; ++a) {
        axProc[a].resize(spec.n[a]);
        axLocal[a].resize(spec.n[a]);
        axCount[a].assign(nProc[a], 0);
        for (int i = 0; i < spec.n[a]; ++i) {
            const int p = (int)((int64_t)i * nProc[a] / spec.n[a]);
            axProc[a][i] = p;
            axLocal[a][i] = axCount[a][p]++;
        }
        bs[a] = axCount[a][pc[a]];
        b0[a] = 0;
        while (b0[a] < spec.n[a] && axProc[a][b0[a]] != pc[a]) b0[a]++;
    }
    const int nx = bs[0], ny = bs[1], nz = bs[2];
    const int nLocal = nx * ny * nz;
    auto lid = [&](int i, int j, int k) { return i + nx * (j + ny * k); };  // block-local ijk
    auto rankOf = [&](int gi, int gj, int gk) {
        return axProc[0][gi] + nProc[0] * (axProc[1][gj] + nProc[1] * axProc[2][gk]);
    };
    auto procLocalId = [&](int gi, int gj, int gk) {
        const int p[3] = {axProc[0][gi], axProc[1][gj], axProc[2][gk]};
        return axLocal[0][gi] +
               axCount[0][p[0]] * (axLocal[1][gj] + axCount[1][p[1]] * axLocal[2][gk]);
    };
    auto classOf = [&](int gi, int gj, int gk) {
        return ax[0].clip(gi) + mdim[0] * (ax[1].clip(gj) + mdim[1] * ax[2].clip(gk));
    };

    HostPlan hp;
    hp.polOrder = prm.polOrder;
    hp.nDvt = K;
    for (int a = 0; a < 3; ++a) hp.dim[a] = dim[a];
    hp.nProcs = nProcs;
    hp.myProc = myProc;
    hp.nCells = nLocal;
    const double h[3] = {(spec.hi[0] - spec.lo[0]) / spec.n[0], (spec.hi[1] - spec.lo[1]) / spec.n[1],
                         (spec.hi[2] - spec.lo[2]) / spec.n[2]};

    // ---- rank-local mesh addressing, in the order extractLocalMesh() produces ----
    // internal faces: upper-triangular order (per cell +x, +y, +z)
    struct FaceGen {
        int32_t own, nei;  // local cells (nei -1: boundary)
        int8_t dir;        // direction seen from the owner
    };
    std::vector<FaceGen> faces;
    faces.reserve((size_t)nLocal * 3 + 2 * ((size_t)nx * ny + (size_t)ny * nz + (size_t)nx * nz));
    for (int k = 0; k < nz; ++k)
        for (int j = 0; j < ny; ++j)
            for (int i = 0; i < nx; ++i) {
                const int c = lid(i, j, k);
                if (i + 1 < nx) faces.push_back({c, lid(i + 1, j, k), 1});
                if (j + 1 < ny) faces.push_back({c, lid(i, j + 1, k), 3});
                if (k + 1 < nz) faces.push_back({c, lid(i, j, k + 1), 5});
            }
    hp.nInternalFaces = (int32_t)faces.size();
    auto addPlane = [&](int dir) {  // boundary faces of the block side `dir`, owner ascending
        const int a = dir / 2, hi = dir & 1;
        const int fix = hi ? bs[a] - 1 : 0;
        for (int k = 0; k < nz; ++k)
            for (int j = 0; j < ny; ++j)
                for (int i = 0; i < nx; ++i) {
                    const int ijk[3] = {i, j, k};
                    if (ijk[a] == fix) faces.push_back({lid(i, j, k), -1, (int8_t)dir});
                }
    };
    static const char* names[6] = {"xlo", "xhi", "ylo", "yhi", "zlo", "zhi"};
    for (int sd = 0; sd < 6; ++sd) {
        const int a = sd / 2, hi = sd & 1;
        Patch p;
        p.name = names[sd];
        p.type = spec.sideType[sd];
        p.nbr = spec.sideType[sd] == PATCH_CYCLIC ? (sd ^ 1) : -1;
        p.start = (int32_t)faces.size();
        const bool onSide = hi ? (b0[a] + bs[a] == spec.n[a]) : (b0[a] == 0);
        if (onSide) addPlane(sd);
        p.size = (int32_t)faces.size() - p.start;
        hp.patches.push_back(p);
    }
    {
        // processor patches by ascending neighbour rank: -z, -y, -x, +x, +y, +z
        const int order[6] = {4, 2, 0, 1, 3, 5};
        const int stride[3] = {1, nProc[0], nProc[0] * nProc[1]};
        for (int o = 0; o < 6; ++o) {
            const int dir = order[o], a = dir / 2, hi = dir & 1;
            if (hi ? (pc[a] + 1 >= nProc[a]) : (pc[a] == 0)) continue;
            Patch p;
            p.type = PATCH_PROCESSOR;
            p.nbr = myProc + (hi ? stride[a] : -stride[a]);
            p.name = "procBoundary" + std::to_string(myProc) + "to" + std::to_string(p.nbr);
            p.start = (int32_t)faces.size();
            addPlane(dir);
            p.size = (int32_t)faces.size() - p.start;
            hp.patches.push_back(p);
        }
    }
    const int nF = (int)faces.size();
    hp.nFaces = nF;
    hp.owner.resize(nF);
    hp.neighbour.resize(hp.nInternalFaces);
    hp.faceAreaVec.assign((size_t)nF * 3, 0.0);
    for (int f = 0; f < nF; ++f) {
        hp.owner[f] = faces[f].own;
        if (f < hp.nInternalFaces) hp.neighbour[f] = faces[f].nei;
        const int a = faces[f].dir / 2;
        const double area = h[(a + 1) % 3] * h[(a + 2) % 3];
        hp.faceAreaVec[(size_t)f * 3 + a] = (faces[f].dir & 1) ? area : -area;
    }
    // cells(): owner faces ascending, then neighbour faces ascending
    {
        std::vector<int32_t> cnt(nLocal, 0);
        for (int f = 0; f < nF; ++f) cnt[hp.owner[f]]++;
        for (int f = 0; f < hp.nInternalFaces; ++f) cnt[hp.neighbour[f]]++;
        hp.cellFaceStart.assign(nLocal + 1, 0);
        for (int c = 0; c < nLocal; ++c) hp.cellFaceStart[c + 1] = hp.cellFaceStart[c] + cnt[c];
        hp.cellFaces.assign(hp.cellFaceStart[nLocal], 0);
        std::vector<int32_t> pos(hp.cellFaceStart.begin(), hp.cellFaceStart.end() - 1);
        for (int f = 0; f < nF; ++f) hp.cellFaces[pos[hp.owner[f]]++] = f;
        for (int f = 0; f < hp.nInternalFaces; ++f) hp.cellFaces[pos[hp.neighbour[f]]++] = f;
    }
    hp.cellCentres.resize((size_t)nLocal * 3);
    std::vector<int32_t> cellClass(nLocal);
    for (int k = 0; k < nz; ++k)
        for (int j = 0; j < ny; ++j)
            for (int i = 0; i < nx; ++i) {
                const int c = lid(i, j, k);
                const int g[3] = {b0[0] + i, b0[1] + j, b0[2] + k};
                for (int a = 0; a < 3; ++a)
                    hp.cellCentres[(size_t)c * 3 + a] = spec.lo[a] + (g[a] + 0.5) * h[a];
                cellClass[c] = classOf(g[0], g[1], g[2]);
            }

    // ---- B, intBasTrans, refFacAr (WENOBase.C:820-854) from the classes ----
    hp.B.resize((size_t)nLocal * K * K);
    parallelFor(nLocal, nThreads, [&](int c) {
        const std::vector<double>& B = cls[cellClass[c]].B;
        std::copy(B.begin(), B.end(), hp.B.begin() + (size_t)c * K * K);
    });
    hp.intBasTrans.assign((size_t)nF * 2 * cube, 0.0);
    hp.refFacAr.assign(nF, 0.0);
    parallelFor(nF, nThreads, [&](int f) {
        const FaceGen& fg = faces[f];
        const ClassData& co = cls[cellClass[fg.own]];
        if (!co.hasFace[fg.dir]) throw std::runtime_error("box_fast: class lacks a face direction");
        std::copy(co.I[fg.dir].begin(), co.I[fg.dir].end(), hp.intBasTrans.begin() + (size_t)f * 2 * cube);
        hp.refFacAr[f] = co.area[fg.dir];
        if (fg.nei >= 0) {
            const ClassData& cn = cls[cellClass[fg.nei]];
            const int od = fg.dir ^ 1;
            if (!cn.hasFace[od]) throw std::runtime_error("box_fast: class lacks a face direction");
            std::copy(cn.I[od].begin(), cn.I[od].end(),
                      hp.intBasTrans.begin() + ((size_t)f * 2 + 1) * cube);
        }
    });

    // ---- stencils: class offsets -> (processor-local id, owner rank) ----
    const bool parallel = nProcs > 1;
    const int R = ax[0].R;
    auto smallIjk = [&](int32_t g, int out[3]) {
        out[0] = g % mdim[0];
        out[1] = (g / mdim[0]) % mdim[1];
        out[2] = g / (mdim[0] * mdim[1]);
    };
    // big global coordinates of representative cell `g` seen from big cell `gc`
    auto toBig = [&](const int gc[3], const int cSmall[3], int32_t g, int out[3]) {
        int gs[3];
        smallIjk(g, gs);
        for (int a = 0; a < 3; ++a) out[a] = ax[a].apply(gc[a], cSmall[a], gs[a]);
    };
    hp.receiveProcList.assign(nProcs, -1);
    hp.sendProcList.assign(nProcs, -1);
    hp.receiveHaloSize.assign(nProcs, 0);
    hp.sendHaloCellIDList.assign(nProcs, {});
    hp.requestHaloCellID.assign(nProcs, {});
    // correctParallelRun (WENOBase.C:1030-1126): halo ids in order of first appearance
    // over the (uncut) central lists of the local cells
    std::vector<int32_t> haloId;  // global cell -> halo id on its owner's list
    const int64_t gStride[3] = {1, spec.n[0], (int64_t)spec.n[0] * spec.n[1]};
    if (parallel) {
        haloId.assign((size_t)spec.n[0] * spec.n[1] * spec.n[2], -1);
        for (int k = 0; k < nz; ++k)
            for (int j = 0; j < ny; ++j)
                for (int i = 0; i < nx; ++i) {
                    const int l[3] = {i, j, k};
                    bool near = false;
                    for (int a = 0; a < 3; ++a) {
                        if (nProc[a] == 1) continue;
                        if (ax[a].n == ax[a].m) near = true;
                        if (pc[a] > 0 && l[a] < R) near = true;
                        if (pc[a] + 1 < nProc[a] && l[a] >= bs[a] - R) near = true;
                    }
                    if (!near) continue;
                    const int gc[3] = {b0[0] + i, b0[1] + j, b0[2] + k};
                    const int cSmall[3] = {ax[0].clip(gc[0]), ax[1].clip(gc[1]), ax[2].clip(gc[2])};
                    const CellSetup& cs = cls[cellClass[lid(i, j, k)]].cs;
                    for (int32_t g : cs.central) {
                        int gb[3];
                        toBig(gc, cSmall, g, gb);
                        const int p = rankOf(gb[0], gb[1], gb[2]);
                        if (p == myProc) continue;
                        hp.receiveProcList[p] = p;
                        const int64_t gid = gb[0] * gStride[0] + gb[1] * gStride[1] + gb[2] * gStride[2];
                        if (haloId[gid] < 0) {
                            haloId[gid] = (int32_t)hp.requestHaloCellID[p].size();
                            hp.requestHaloCellID[p].push_back(procLocalId(gb[0], gb[1], gb[2]));
                        }
                    }
                }
    }
    for (int p = 0; p < nProcs; ++p) hp.receiveHaloSize[p] = (int32_t)hp.requestHaloCellID[p].size();

    // matrix pool: classes enter in order of first appearance over the local cells
    detail::MatrixPool pool;
    pool.plan = &hp;
    pool.nDvt = K;
    hp.matrixStart.assign(1, 0);
    hp.cellStencilStart.assign(nLocal + 1, 0);
    std::vector<int64_t> cellEntryStart(nLocal + 1, 0);
    for (int c = 0; c < nLocal; ++c) {
        ClassData& cd = cls[cellClass[c]];
        if (!cd.pooled) {
            cd.poolId.assign(cd.cs.globalID.size(), -1);
            for (size_t s = 0; s < cd.cs.globalID.size(); ++s)
                if (cd.cs.globalID[s][0] != CELL_DELETED) cd.poolId[s] = pool.similar(cd.cs.Ainv[s]);
            cd.pooled = true;
        }
        int64_t ne = 0;
        for (const auto& st : cd.cs.globalID) {
            ne += (int64_t)st.size();
            if (st[0] != CELL_DELETED) hp.nMatricesTotal++;
        }
        hp.cellStencilStart[c + 1] = hp.cellStencilStart[c] + (int64_t)cd.cs.globalID.size();
        cellEntryStart[c + 1] = cellEntryStart[c] + ne;
    }
    const int64_t nStencils = hp.cellStencilStart[nLocal], nEntries = cellEntryStart[nLocal];
    hp.stencilEntryStart.assign(nStencils + 1, 0);
    hp.stencilMatrix.assign(nStencils, -1);
    hp.stencilsID.assign(nEntries, 0);
    if (parallel) hp.cellToProcMap.assign(nEntries, CELL_LOCAL);
    hp.stencilEntryStart[nStencils] = nEntries;
    parallelFor(nLocal, nThreads, [&](int c) {
        const ClassData& cd = cls[cellClass[c]];
        const int l[3] = {c % nx, (c / nx) % ny, c / (nx * ny)};
        const int gc[3] = {b0[0] + l[0], b0[1] + l[1], b0[2] + l[2]};
        const int cSmall[3] = {ax[0].clip(gc[0]), ax[1].clip(gc[1]), ax[2].clip(gc[2])};
        int64_t e = cellEntryStart[c];
        int64_t si = hp.cellStencilStart[c];
        for (size_t s = 0; s < cd.cs.globalID.size(); ++s, ++si) {
            const auto& st = cd.cs.globalID[s];
            hp.stencilEntryStart[si] = e;
            if (st[0] == CELL_DELETED) {
                hp.stencilsID[e] = (s == 0 && cd.cs.empty) ? CELL_EMPTY : CELL_DELETED;
                if (parallel) hp.cellToProcMap[e] = CELL_DELETED;
                e++;
                continue;
            }
            hp.stencilMatrix[si] = cd.poolId[s];
            const int64_t e0 = e;
            for (int32_t g : st) {
                int gb[3];
                toBig(gc, cSmall, g, gb);
                const int p = parallel ? rankOf(gb[0], gb[1], gb[2]) : myProc;
                if (p == myProc) {
                    hp.stencilsID[e] = lid(gb[0] - b0[0], gb[1] - b0[1], gb[2] - b0[2]);
                } else {
                    hp.stencilsID[e] = haloId[gb[0] * gStride[0] + gb[1] * gStride[1] + gb[2] * gStride[2]];
                    hp.cellToProcMap[e] = p;
                }
                e++;
            }
            if (s == 0 && cd.cs.empty) hp.stencilsID[e0] = CELL_EMPTY;
        }
    });

    // patchNeighbourField addressing for the bounded limiter (WENOUpwindFit.C:130)
    {
        std::vector<int64_t> hoff(nProcs + 1, 0);
        for (int p = 0; p < nProcs; ++p) hoff[p + 1] = hoff[p] + hp.receiveHaloSize[p];
        const int nBnd = nF - hp.nInternalFaces;
        hp.patchNbrCell.assign(std::max(nBnd, 1), -1);
        for (const Patch& pt : hp.patches) {
            if (pt.type != PATCH_CYCLIC && pt.type != PATCH_PROCESSOR) continue;
            for (int i = 0; i < pt.size; ++i) {
                const int f = pt.start + i;
                const int c = hp.owner[f];
                if (pt.type == PATCH_CYCLIC) {
                    hp.patchNbrCell[f - hp.nInternalFaces] = hp.owner[hp.patches[pt.nbr].start + i];
                } else {
                    const int dir = faces[f].dir, a = dir / 2;
                    int gb[3] = {b0[0] + c % nx, b0[1] + (c / nx) % ny, b0[2] + c / (nx * ny)};
                    gb[a] += (dir & 1) ? 1 : -1;
                    const int p = rankOf(gb[0], gb[1], gb[2]);
                    const int32_t hid = haloId[gb[0] * gStride[0] + gb[1] * gStride[1] + gb[2] * gStride[2]];
                    if (hid < 0) throw std::runtime_error("box_fast: face neighbour missing from the halo");
                    hp.patchNbrCell[f - hp.nInternalFaces] = (int32_t)(nLocal + hoff[p] + hid);
                }
            }
        }
    }
    return hp;
}

}  // namespace wenoext
