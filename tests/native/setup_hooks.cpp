// Test-only C hooks into the host setup helpers (geometry_weno.cpp), so that the
// known answers of the reference's own unit tests can be checked from pytest:
//   tests/src/geometryWENO-BasicFunc-Test.C:50-145 (Jacobian, transformPoint),
//   :149-203 (gaussQuad), :206-478 (initIntegrals / transformIntegral),
//   tests/src/mathFunctionsWENO-Test.C:45-131 (det, eigenvalues, inverse).
// Built by wenoext_b200/build.py into tests/native/libsetup_hooks.so; not part
// of the product ABI (include/wenoext_b200.h).
#include <array>
#include <cstring>
#include <vector>

#include "setup/geometry_weno.h"
#include "setup/polymesh.h"

using namespace wenoext;

static Mat3 toMat(const double* a) {
    Mat3 M;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) M(i, j) = a[3 * i + j];
    return M;
}
static void fromMat(const Mat3& M, double* a) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) a[3 * i + j] = M(i, j);
}
static Vec3 toVec(const double* v) { return Vec3{v[0], v[1], v[2]}; }

extern "C" {

double hook_det3(const double* A) { return geom::det3(toMat(A)); }

void hook_inv3(const double* A, double* out) { fromMat(geom::inv3(toMat(A)), out); }

void hook_eigen_magnitudes(const double* A, double* out) { geom::eigenMagnitudes(toMat(A), out); }

void hook_transform_point(const double* Jinv, const double* xp, const double* x0, double* out) {
    const Vec3 r = geom::transformPoint(toMat(Jinv), toVec(xp), toVec(x0));
    out[0] = r.x;
    out[1] = r.y;
    out[2] = r.z;
}

double hook_gauss_quad(int n, int m, int l, const double* p0, const double* v0, const double* v1,
                       const double* v2) {
    return geom::gaussQuad(n, m, l, toVec(p0), toVec(v0), toVec(v1), toVec(v2));
}

// initIntegrals + transformIntegral of one cell of a hex box (the reference uses
// cell 33 of its 100x100x1 case). vol/trans: (polOrder+1)^3 doubles each.
int hook_cell_integrals(const int32_t* n, const double* lo, const double* hi, int cellI,
                        int polOrder, double* vol, double* trans, double* JInv, double* refPoint,
                        double* refDet) {
    BoxSpec spec;
    for (int a = 0; a < 3; ++a) {
        spec.n[a] = n[a];
        spec.lo[a] = lo[a];
        spec.hi[a] = hi[a];
    }
    PolyMesh mesh = makeBoxMesh(spec);
    if (cellI < 0 || cellI >= mesh.nCells) return 1;
    List3D V;
    V.resize(polOrder + 1);
    Mat3 J;
    Vec3 rp;
    double det = 0;
    geom::initIntegrals(mesh, cellI, polOrder, V, J, rp, det);
    const Vec3 tc = geom::transformPoint(J, mesh.cellCentres[cellI], rp);
    List3D T;
    T.resize(polOrder + 1);
    geom::transformIntegral(mesh, cellI, tc, polOrder, J, rp, det, T);
    std::memcpy(vol, V.d.data(), V.d.size() * sizeof(double));
    std::memcpy(trans, T.d.data(), T.d.size() * sizeof(double));
    fromMat(J, JInv);
    refPoint[0] = rp.x;
    refPoint[1] = rp.y;
    refPoint[2] = rp.z;
    *refDet = det;
    return 0;
}

// A box mesh (hexes, or 6 tets per hex) with the tables wenoext_ls_batch wants, and the host's
// initIntegrals / transformIntegral of (cell I, cell J) pairs on it: the reference data of
// tests/test_gpu_pinv.py::test_ls_rows_on_device_equal_host_rows.
struct HookMesh {
    PolyMesh mesh;
    std::vector<double> points, centres;
    std::vector<int64_t> triStart;
    std::vector<int32_t> triPoints;
};
void* hook_mesh_create(const int32_t* n, const double* lo, const double* hi, int tetSplit) {
    BoxSpec spec;
    for (int a = 0; a < 3; ++a) {
        spec.n[a] = n[a];
        spec.lo[a] = lo[a];
        spec.hi[a] = hi[a];
    }
    spec.tetSplit = tetSplit;
    HookMesh* h = new HookMesh;
    h->mesh = makeBoxMesh(spec);
    const PolyMesh& m = h->mesh;
    for (const Vec3& p : m.points) h->points.insert(h->points.end(), {p.x, p.y, p.z});
    h->triStart.assign(1, 0);
    std::vector<std::array<int32_t, 3>> tris;
    for (int c = 0; c < m.nCells; ++c) {
        geom::cellTriangleList(m, c, tris);
        for (const auto& t : tris) h->triPoints.insert(h->triPoints.end(), t.begin(), t.end());
        h->triStart.push_back((int64_t)h->triPoints.size() / 3);
        h->centres.insert(h->centres.end(), {m.cellCentres[c].x, m.cellCentres[c].y, m.cellCentres[c].z});
    }
    return h;
}
void hook_mesh_destroy(void* p) { delete static_cast<HookMesh*>(p); }
void hook_mesh_sizes(void* p, int32_t* nPoints, int32_t* nCells, int64_t* nTri) {
    const HookMesh* h = static_cast<HookMesh*>(p);
    *nPoints = (int32_t)h->mesh.points.size();
    *nCells = h->mesh.nCells;
    *nTri = (int64_t)h->triPoints.size() / 3;
}
void hook_mesh_copy(void* p, double* points, int64_t* triStart, int32_t* triPoints, double* centres, double* volumes) {
    const HookMesh* h = static_cast<HookMesh*>(p);
    std::memcpy(points, h->points.data(), h->points.size() * 8);
    std::memcpy(triStart, h->triStart.data(), h->triStart.size() * 8);
    std::memcpy(triPoints, h->triPoints.data(), h->triPoints.size() * 4);
    std::memcpy(centres, h->centres.data(), h->centres.size() * 8);
    std::memcpy(volumes, h->mesh.cellVolumes.data(), h->mesh.cellVolumes.size() * 8);
}
// frame of cell I (volI cube, JInv, refPoint, refDet, tI) and, for each cell J of the list, tJ and the
// transformed volume moments of J in that frame (cubes of (polOrder+1)^3)
int hook_pair_moments(void* p, int polOrder, int cellI, int nJ, const int32_t* cellsJ, double* volI, double* JInv,
                      double* refPoint, double* refDet, double* tI, double* tJ, double* mom) {
    const PolyMesh& m = static_cast<HookMesh*>(p)->mesh;
    if (cellI < 0 || cellI >= m.nCells) return 1;
    List3D V;
    Mat3 J;
    Vec3 rp;
    double det = 0;
    geom::initIntegrals(m, cellI, polOrder, V, J, rp, det);
    std::memcpy(volI, V.d.data(), V.d.size() * 8);
    fromMat(J, JInv);
    refPoint[0] = rp.x, refPoint[1] = rp.y, refPoint[2] = rp.z;
    *refDet = det;
    const Vec3 ti = geom::transformPoint(J, m.cellCentres[cellI], rp);
    tI[0] = ti.x, tI[1] = ti.y, tI[2] = ti.z;
    const size_t cube = V.d.size();
    for (int k = 0; k < nJ; ++k) {
        if (cellsJ[k] < 0 || cellsJ[k] >= m.nCells) return 1;
        const Vec3 tj = geom::transformPoint(J, m.cellCentres[cellsJ[k]], rp);
        tJ[3 * k] = tj.x, tJ[3 * k + 1] = tj.y, tJ[3 * k + 2] = tj.z;
        List3D T;
        geom::transformIntegral(m, cellsJ[k], tj, polOrder, J, rp, det, T);
        std::memcpy(mom + (size_t)k * cube, T.d.data(), cube * 8);
    }
    return 0;
}
}
