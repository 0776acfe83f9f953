// Test-only C hooks into the host setup helpers (geometry_weno.cpp), so that the
// known answers of the reference's own unit tests can be checked from pytest:
//   tests/src/geometryWENO-BasicFunc-Test.C:50-145 (Jacobian, transformPoint),
//   :149-203 (gaussQuad), :206-478 (initIntegrals / transformIntegral),
//   tests/src/mathFunctionsWENO-Test.C:45-131 (det, eigenvalues, inverse).
// Built by wenoext_b200/build.py into tests/native/libsetup_hooks.so; not part
// of the product ABI (include/wenoext_b200.h).
#include <cstring>

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
}
