// Host restatement of the geometry helpers the WENOBase setup uses
// (libWENOEXT/WENOBase/geometryWENO/geometryWENO.{H,C}, mathFunctionsWENO.H,
// List3D.H). Setup only: runs once on the host, never on the per-step path.
#pragma once
#include <array>
#include <vector>

#include "polymesh.h"

namespace wenoext {

struct Mat3 {
    double a[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    double& operator()(int i, int j) { return a[i][j]; }
    double operator()(int i, int j) const { return a[i][j]; }
};

// List3D<scalar> (List3D.H:226-256): dense cube, index i*Y*Z + j*Z + k.
struct List3D {
    int s = 0;
    std::vector<double> d;
    void resize(int n) {
        s = n;
        d.assign((size_t)n * n * n, 0.0);
    }
    void setZero() { std::fill(d.begin(), d.end(), 0.0); }
    double& operator()(int i, int j, int k) { return d[(size_t)i * s * s + j * s + k]; }
    double operator()(int i, int j, int k) const { return d[(size_t)i * s * s + j * s + k]; }
};

namespace geom {

// mathFunctionsWENO.H:105-140, :212-240, :142-210
double det2(double a, double d, double c, double b);
double det3(const Mat3& A);
Mat3 inv3(const Mat3& A);
void pivot(Mat3& A);
// |eigenvalues| of a real 3x3 (stands in for realEigenValues.H); cond() as
// geometryWENO.C:917-931
void eigenMagnitudes(const Mat3& A, double out[3]);
double cond(const Mat3& J);

// all triangles of a cell's faces, face after face, each face as a fan from its first point
// (polyMeshTetDecomposition::cellTetIndices stand-in)
void cellTriangleList(const PolyMesh& mesh, int cellI, std::vector<std::array<int32_t, 3>>& out);
// geometryWENO.C:327-343
Vec3 transformPoint(const Mat3& Jinv, const Vec3& xP, const Vec3& x0);
// geometryWENO.C:346-460 (scalar branch)
double gaussQuad(int n, int m, int l, const Vec3& xi0, const Vec3& v0, const Vec3& v1,
                 const Vec3& v2);

// geometryWENO.C:42-219
void initIntegrals(const PolyMesh& mesh, int cellI, int polOrder, List3D& volIntegrals,
                   Mat3& JInvI, Vec3& refPointI, double& refDetI);
// geometryWENO.C:222-325
void transformIntegral(const PolyMesh& mesh, int cellJ, const Vec3& transCenterJ, int polOrder,
                       const Mat3& JInvI, const Vec3& refPointI, double refDetI,
                       List3D& transVolMom);
// geometryWENO.C:464-586
void smoothIndIntegrals(const PolyMesh& mesh, int cellI, int polOrder, const Mat3& JInvI,
                        const Vec3& refPointI, List3D& out);
// geometryWENO.C:588-710; returns row-major nDvt x nDvt
std::vector<double> getB(const PolyMesh& mesh, int cellI, int polOrder, int nDvt,
                         const Mat3& JInvI, const Vec3& refPointI, const int dim[3]);
// geometryWENO.C:713-874, one cell's contribution. intBasTrans is
// [nFaces][2] cubes of (polOrder+1)^3; refFacAr [nFaces] must be zeroed first.
void surfIntTransCell(const PolyMesh& mesh, int cellI, int polOrder, const List3D& volIntegralsI,
                      const Mat3& JInvI, const Vec3& refPointI, double* intBasTrans,
                      double* refFacAr);

// one face of one cell; I (cube of (polOrder+1)^3) must be zeroed first
void surfIntTransFace(const PolyMesh& mesh, int cellI, int f, int polOrder,
                      const List3D& volIntegralsI, const Mat3& JInvI, const Vec3& refPointI,
                      double* I, double& faceArea);

}  // namespace geom
}  // namespace wenoext
