// Stand-in for the parts of OpenFOAM's polyMesh/primitiveMesh the WENOBase
// setup reads. Geometry follows primitiveMesh's face-centre / cell-centre
// decomposition so that C(), cellVolumes() and Sf() agree with what the
// reference (libWENOEXT/WENOBase/WENOBase.C, geometryWENO.C) would be given.
#include "polymesh.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <stdexcept>

namespace wenoext {

double mag(Vec3 a) { return std::sqrt(dot(a, a)); }

void PolyMesh::faceTriangles(int f, std::vector<std::array<int32_t, 3>>& tris) const {
    tris.clear();
    const int n = faceSize(f);
    const int32_t* p = face(f);
    for (int i = 1; i + 1 < n; ++i) tris.push_back({p[0], p[i], p[i + 1]});
}

void PolyMesh::finalize() {
    const int nF = nFaces();
    // face centres and area vectors (primitiveMeshFaceCentresAndAreas.C algorithm)
    faceCentres.assign(nF, Vec3{});
    faceAreas.assign(nF, Vec3{});
    for (int f = 0; f < nF; ++f) {
        const int n = faceSize(f);
        const int32_t* p = face(f);
        if (n == 3) {
            faceCentres[f] = (1.0 / 3.0) * (points[p[0]] + points[p[1]] + points[p[2]]);
            faceAreas[f] = 0.5 * cross(points[p[1]] - points[p[0]], points[p[2]] - points[p[0]]);
            continue;
        }
        Vec3 est{};
        for (int i = 0; i < n; ++i) est = est + points[p[i]];
        est = (1.0 / n) * est;
        Vec3 sumN{}, sumAc{};
        double sumA = 0;
        for (int i = 0; i < n; ++i) {
            const Vec3& a = points[p[i]];
            const Vec3& b = points[p[(i + 1) % n]];
            Vec3 c = a + b + est;
            Vec3 nn = cross(b - a, est - a);
            double ar = mag(nn);
            sumN = sumN + nn;
            sumA += ar;
            sumAc = sumAc + ar * c;
        }
        faceCentres[f] = sumA > 0 ? (1.0 / 3.0) * ((1.0 / sumA) * sumAc) : est;
        faceAreas[f] = 0.5 * sumN;
    }
    // cells(): owner faces ascending, then neighbour faces ascending
    std::vector<int32_t> cnt(nCells, 0);
    for (int f = 0; f < nF; ++f) cnt[owner[f]]++;
    for (int f = 0; f < nInternalFaces; ++f) cnt[neighbour[f]]++;
    cellFaceStart.assign(nCells + 1, 0);
    for (int c = 0; c < nCells; ++c) cellFaceStart[c + 1] = cellFaceStart[c] + cnt[c];
    cellFaces.assign(cellFaceStart[nCells], 0);
    std::vector<int32_t> pos(cellFaceStart.begin(), cellFaceStart.end() - 1);
    for (int f = 0; f < nF; ++f) cellFaces[pos[owner[f]]++] = f;
    for (int f = 0; f < nInternalFaces; ++f) cellFaces[pos[neighbour[f]]++] = f;
    // cellCells(): per internal face ascending, appended to both cells
    std::fill(cnt.begin(), cnt.end(), 0);
    for (int f = 0; f < nInternalFaces; ++f) {
        cnt[owner[f]]++;
        cnt[neighbour[f]]++;
    }
    cellCellStart.assign(nCells + 1, 0);
    for (int c = 0; c < nCells; ++c) cellCellStart[c + 1] = cellCellStart[c] + cnt[c];
    cellCells.assign(cellCellStart[nCells], 0);
    pos.assign(cellCellStart.begin(), cellCellStart.end() - 1);
    for (int f = 0; f < nInternalFaces; ++f) {
        cellCells[pos[owner[f]]++] = neighbour[f];
        cellCells[pos[neighbour[f]]++] = owner[f];
    }
    // cell centres and volumes (primitiveMeshCellCentresAndVols.C algorithm)
    cellCentres.assign(nCells, Vec3{});
    cellVolumes.assign(nCells, 0.0);
    std::vector<Vec3> cEst(nCells, Vec3{});
    std::vector<int32_t> nCellFaces(nCells, 0);
    for (int f = 0; f < nF; ++f) {
        cEst[owner[f]] = cEst[owner[f]] + faceCentres[f];
        nCellFaces[owner[f]]++;
    }
    for (int f = 0; f < nInternalFaces; ++f) {
        cEst[neighbour[f]] = cEst[neighbour[f]] + faceCentres[f];
        nCellFaces[neighbour[f]]++;
    }
    for (int c = 0; c < nCells; ++c) cEst[c] = (1.0 / nCellFaces[c]) * cEst[c];
    auto addPyr = [&](int c, int f, double sgn) {
        double pyr3Vol = sgn * dot(faceAreas[f], faceCentres[f] - cEst[c]);
        Vec3 pc = 0.75 * faceCentres[f] + 0.25 * cEst[c];
        cellCentres[c] = cellCentres[c] + pyr3Vol * pc;
        cellVolumes[c] += pyr3Vol;
    };
    for (int f = 0; f < nF; ++f) addPyr(owner[f], f, 1.0);
    for (int f = 0; f < nInternalFaces; ++f) addPyr(neighbour[f], f, -1.0);
    for (int c = 0; c < nCells; ++c) {
        if (std::fabs(cellVolumes[c]) > 1e-300)
            cellCentres[c] = (1.0 / cellVolumes[c]) * cellCentres[c];
        else
            cellCentres[c] = cEst[c];
        cellVolumes[c] *= (1.0 / 3.0);
    }
}

namespace {
struct FaceRec {
    std::array<int32_t, 4> key;  // sorted point ids (tri: last = -1 -> INT_MAX)
    int32_t cell;
    int32_t lf;
};
}  // namespace

PolyMesh buildFromCells(const std::vector<Vec3>& points, const std::vector<CellShape>& cells,
                        const std::vector<Patch>& patchProto,
                        const std::function<int(const std::vector<int32_t>&)>& patchOf) {
    std::vector<FaceRec> recs;
    recs.reserve(cells.size() * 6);
    for (size_t c = 0; c < cells.size(); ++c)
        for (size_t lf = 0; lf < cells[c].faces.size(); ++lf) {
            const auto& fp = cells[c].faces[lf];
            if (fp.size() > 4) throw std::runtime_error("buildFromCells: faces with >4 points unsupported");
            FaceRec r;
            r.key = {INT32_MAX, INT32_MAX, INT32_MAX, INT32_MAX};
            for (size_t i = 0; i < fp.size(); ++i) r.key[i] = fp[i];
            std::sort(r.key.begin(), r.key.end());
            r.cell = (int32_t)c;
            r.lf = (int32_t)lf;
            recs.push_back(r);
        }
    std::sort(recs.begin(), recs.end(), [](const FaceRec& a, const FaceRec& b) {
        if (a.key != b.key) return a.key < b.key;
        return a.cell < b.cell;
    });
    struct Tmp {
        int32_t own, nei, lf, patch;
    };
    std::vector<Tmp> internal, boundary;
    for (size_t i = 0; i < recs.size();) {
        if (i + 1 < recs.size() && recs[i].key == recs[i + 1].key) {
            internal.push_back({recs[i].cell, recs[i + 1].cell, recs[i].lf, -1});
            i += 2;
        } else {
            int p = patchOf(cells[recs[i].cell].faces[recs[i].lf]);
            boundary.push_back({recs[i].cell, -1, recs[i].lf, p});
            i += 1;
        }
    }
    std::sort(internal.begin(), internal.end(), [](const Tmp& a, const Tmp& b) {
        return a.own != b.own ? a.own < b.own : a.nei < b.nei;
    });
    std::stable_sort(boundary.begin(), boundary.end(), [](const Tmp& a, const Tmp& b) {
        if (a.patch != b.patch) return a.patch < b.patch;
        if (a.own != b.own) return a.own < b.own;
        return a.lf < b.lf;
    });
    PolyMesh m;
    m.points = points;
    m.nCells = (int32_t)cells.size();
    m.nInternalFaces = (int32_t)internal.size();
    m.patches = patchProto;
    for (auto& p : m.patches) p.size = 0;
    m.faceStart.push_back(0);
    // approximate cell centres for orientation
    auto cellAvg = [&](int c) {
        Vec3 s{};
        int n = 0;
        for (const auto& f : cells[c].faces)
            for (int32_t p : f) {
                s = s + points[p];
                n++;
            }
        return (1.0 / n) * s;
    };
    auto emit = [&](const Tmp& t) {
        std::vector<int32_t> fp = cells[t.own].faces[t.lf];
        Vec3 fc{};
        for (int32_t p : fp) fc = fc + points[p];
        fc = (1.0 / fp.size()) * fc;
        Vec3 nrm{};
        for (size_t i = 0; i < fp.size(); ++i)
            nrm = nrm + cross(points[fp[i]] - fc, points[fp[(i + 1) % fp.size()]] - fc);
        if (dot(nrm, fc - cellAvg(t.own)) < 0) std::reverse(fp.begin() + 1, fp.end());
        for (int32_t p : fp) m.facePoints.push_back(p);
        m.faceStart.push_back((int32_t)m.facePoints.size());
        m.owner.push_back(t.own);
    };
    for (const auto& t : internal) {
        emit(t);
        m.neighbour.push_back(t.nei);
    }
    for (const auto& t : boundary) {
        emit(t);
        m.patches[t.patch].size++;
    }
    int32_t start = m.nInternalFaces;
    for (auto& p : m.patches) {
        p.start = start;
        start += p.size;
    }
    m.finalize();
    return m;
}

PolyMesh makeBoxMesh(const BoxSpec& s) {
    const int nx = s.n[0], ny = s.n[1], nz = s.n[2];
    std::vector<Vec3> pts((size_t)(nx + 1) * (ny + 1) * (nz + 1));
    auto pid = [&](int i, int j, int k) { return i + (nx + 1) * (j + (ny + 1) * k); };
    const int den[3] = {s.nDen[0] ? s.nDen[0] : nx, s.nDen[1] ? s.nDen[1] : ny,
                        s.nDen[2] ? s.nDen[2] : nz};
    for (int k = 0; k <= nz; ++k)
        for (int j = 0; j <= ny; ++j)
            for (int i = 0; i <= nx; ++i)
                pts[pid(i, j, k)] = {s.lo[0] + (s.hi[0] - s.lo[0]) * i / den[0],
                                     s.lo[1] + (s.hi[1] - s.lo[1]) * j / den[1],
                                     s.lo[2] + (s.hi[2] - s.lo[2]) * k / den[2]};
    // actual upper corner (differs from s.hi for a sub-box)
    const double hiAct[3] = {s.lo[0] + (s.hi[0] - s.lo[0]) * nx / den[0],
                             s.lo[1] + (s.hi[1] - s.lo[1]) * ny / den[1],
                             s.lo[2] + (s.hi[2] - s.lo[2]) * nz / den[2]};
    std::vector<CellShape> cells;
    cells.reserve((size_t)nx * ny * nz * (s.tetSplit ? 6 : 1));
    for (int k = 0; k < nz; ++k)
        for (int j = 0; j < ny; ++j)
            for (int i = 0; i < nx; ++i) {
                int32_t v[2][2][2];
                for (int a = 0; a < 2; ++a)
                    for (int b = 0; b < 2; ++b)
                        for (int c = 0; c < 2; ++c) v[a][b][c] = pid(i + a, j + b, k + c);
                if (!s.tetSplit) {
                    CellShape h;
                    h.faces = {{v[0][0][0], v[0][0][1], v[0][1][1], v[0][1][0]},
                               {v[1][0][0], v[1][1][0], v[1][1][1], v[1][0][1]},
                               {v[0][0][0], v[1][0][0], v[1][0][1], v[0][0][1]},
                               {v[0][1][0], v[0][1][1], v[1][1][1], v[1][1][0]},
                               {v[0][0][0], v[0][1][0], v[1][1][0], v[1][0][0]},
                               {v[0][0][1], v[1][0][1], v[1][1][1], v[0][1][1]}};
                    cells.push_back(std::move(h));
                } else {
                    // Kuhn split: one tet per permutation of the axes
                    static const int perm[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2},
                                                   {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
                    for (int t = 0; t < 6; ++t) {
                        int a[3] = {0, 0, 0};
                        int32_t q[4];
                        q[0] = v[0][0][0];
                        for (int e = 0; e < 3; ++e) {
                            a[perm[t][e]] = 1;
                            q[e + 1] = v[a[0]][a[1]][a[2]];
                        }
                        CellShape tet;
                        tet.faces = {{q[0], q[1], q[2]}, {q[0], q[1], q[3]},
                                     {q[0], q[2], q[3]}, {q[1], q[2], q[3]}};
                        cells.push_back(std::move(tet));
                    }
                }
            }
    static const char* names[6] = {"xlo", "xhi", "ylo", "yhi", "zlo", "zhi"};
    std::vector<Patch> proto(6);
    for (int sd = 0; sd < 6; ++sd) {
        proto[sd].name = names[sd];
        proto[sd].type = s.sideType[sd];
        proto[sd].nbr = (s.sideType[sd] == PATCH_CYCLIC) ? (sd ^ 1) : -1;
    }
    const double tol[3] = {1e-9 * (hiAct[0] - s.lo[0]), 1e-9 * (hiAct[1] - s.lo[1]),
                           1e-9 * (hiAct[2] - s.lo[2])};
    auto patchOf = [&](const std::vector<int32_t>& fp) -> int {
        for (int ax = 0; ax < 3; ++ax) {
            bool lo = true, hi = true;
            for (int32_t p : fp) {
                if (std::fabs(pts[p][ax] - s.lo[ax]) > tol[ax]) lo = false;
                if (std::fabs(pts[p][ax] - hiAct[ax]) > tol[ax]) hi = false;
            }
            if (lo) return 2 * ax;
            if (hi) return 2 * ax + 1;
        }
        throw std::runtime_error("makeBoxMesh: boundary face not on a box side");
    };
    PolyMesh m = buildFromCells(pts, cells, proto, patchOf);
    for (int ax = 0; ax < 3; ++ax) {
        bool empty = s.sideType[2 * ax] == PATCH_EMPTY && s.sideType[2 * ax + 1] == PATCH_EMPTY;
        m.solutionD[ax] = empty ? 0 : 1;
    }
    return m;
}

std::vector<int32_t> decomposeBox(const BoxSpec& s, const int32_t np[3]) {
    const int mult = s.tetSplit ? 6 : 1;
    std::vector<int32_t> proc((size_t)s.n[0] * s.n[1] * s.n[2] * mult);
    size_t c = 0;
    for (int k = 0; k < s.n[2]; ++k)
        for (int j = 0; j < s.n[1]; ++j)
            for (int i = 0; i < s.n[0]; ++i) {
                int pi = (int)((int64_t)i * np[0] / s.n[0]);
                int pj = (int)((int64_t)j * np[1] / s.n[1]);
                int pk = (int)((int64_t)k * np[2] / s.n[2]);
                for (int t = 0; t < mult; ++t) proc[c++] = pi + np[0] * (pj + np[1] * pk);
            }
    return proc;
}

LocalMesh extractLocalMesh(const PolyMesh& g, const std::vector<int32_t>& cellProc, int32_t me,
                           std::vector<int32_t>& g2l) {
    LocalMesh L;
    PolyMesh& m = L.mesh;
    g2l.assign(g.nCells, -1);
    for (int c = 0; c < g.nCells; ++c)
        if (cellProc[c] == me) {
            g2l[c] = (int32_t)L.localToGlobalCell.size();
            L.localToGlobalCell.push_back(c);
        }
    m.nCells = (int32_t)L.localToGlobalCell.size();
    std::vector<int32_t> pointMap(g.points.size(), -1);
    auto addFace = [&](int gf, bool flip) {
        const int n = g.faceSize(gf);
        const int32_t* p = g.face(gf);
        for (int i = 0; i < n; ++i) {
            int32_t gp = flip ? p[(n - i) % n] : p[i];
            if (pointMap[gp] < 0) {
                pointMap[gp] = (int32_t)m.points.size();
                m.points.push_back(g.points[gp]);
            }
            m.facePoints.push_back(pointMap[gp]);
        }
        m.faceStart.push_back((int32_t)m.facePoints.size());
        L.localToGlobalFace.push_back(gf);
    };
    m.faceStart.push_back(0);
    // internal faces (global order keeps upper-triangular order locally)
    std::map<int32_t, std::vector<std::pair<int32_t, bool>>> procFaces;  // nbr -> (gf, flip)
    for (int f = 0; f < g.nInternalFaces; ++f) {
        int o = g.owner[f], n = g.neighbour[f];
        bool lo = cellProc[o] == me, ln = cellProc[n] == me;
        if (lo && ln) {
            addFace(f, false);
            m.owner.push_back(g2l[o]);
            m.neighbour.push_back(g2l[n]);
        } else if (lo) {
            procFaces[cellProc[n]].push_back({f, false});
        } else if (ln) {
            procFaces[cellProc[o]].push_back({f, true});
        }
    }
    m.nInternalFaces = (int32_t)m.owner.size();
    for (const Patch& gp : g.patches) {
        Patch p = gp;
        p.start = (int32_t)m.owner.size();
        for (int i = 0; i < gp.size; ++i) {
            int f = gp.start + i;
            if (cellProc[g.owner[f]] == me) {
                addFace(f, false);
                m.owner.push_back(g2l[g.owner[f]]);
            }
        }
        p.size = (int32_t)m.owner.size() - p.start;
        m.patches.push_back(p);
    }
    for (auto& kv : procFaces) {
        Patch p;
        p.name = "procBoundary" + std::to_string(me) + "to" + std::to_string(kv.first);
        p.type = PATCH_PROCESSOR;
        p.nbr = kv.first;
        p.start = (int32_t)m.owner.size();
        for (auto& fr : kv.second) {
            addFace(fr.first, fr.second);
            int gc = fr.second ? g.neighbour[fr.first] : g.owner[fr.first];
            m.owner.push_back(g2l[gc]);
        }
        p.size = (int32_t)m.owner.size() - p.start;
        m.patches.push_back(p);
    }
    for (int ax = 0; ax < 3; ++ax) m.solutionD[ax] = g.solutionD[ax];
    m.finalize();
    return L;
}

}  // namespace wenoext
