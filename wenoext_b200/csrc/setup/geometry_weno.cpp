// See geometry_weno.h. Each routine cites the reference lines it restates.
#include "geometry_weno.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <stdexcept>

namespace wenoext {
namespace geom {

namespace {
const double ROOTVSMALL = 1.0e-150;  // OpenFOAM doubleScalarROOTVSMALL
const double SMALL = 1.0e-15;        // OpenFOAM doubleScalarSMALL
const double GREAT = 1.0e+15;        // OpenFOAM doubleScalarGREAT

// geometryWENO.H:198-213: 13-point degree-7 rule on the unit triangle
const double gaussCoeff[13][3] = {
    {0.33333333333333, 0.33333333333333, -0.14957004446768},
    {0.26034596607904, 0.26034596607904, 0.17561525743321},
    {0.26034596607904, 0.47930806784192, 0.17561525743321},
    {0.47930806784192, 0.26034596607904, 0.17561525743321},
    {0.06513010290222, 0.06513010290222, 0.05334723560884},
    {0.06513010290222, 0.86973979419557, 0.05334723560884},
    {0.86973979419557, 0.06513010290222, 0.05334723560884},
    {0.31286549600487, 0.63844418856981, 0.07711376089026},
    {0.63844418856981, 0.04869031542532, 0.07711376089026},
    {0.04869031542532, 0.31286549600487, 0.07711376089026},
    {0.63844418856981, 0.31286549600487, 0.07711376089026},
    {0.31286549600487, 0.04869031542532, 0.07711376089026},
    {0.04869031542532, 0.63844418856981, 0.07711376089026}};

inline double intPow(double base, unsigned e) {
    double t = 1.0;
    for (unsigned i = 0; i < e; ++i) t *= base;
    return t;
}

Mat3 matmul(const Mat3& A, const Mat3& B) {
    Mat3 C;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += A(i, k) * B(k, j);
            C(i, j) = s;
        }
    return C;
}

// all triangles of a cell's faces (polyMeshTetDecomposition::cellTetIndices stand-in)
void cellTriangles(const PolyMesh& mesh, int cellI, std::vector<std::array<int32_t, 3>>& out) {
    out.clear();
    std::vector<std::array<int32_t, 3>> t;
    const int nf = mesh.cellNFaces(cellI);
    const int32_t* fl = mesh.cellFaceList(cellI);
    for (int i = 0; i < nf; ++i) {
        mesh.faceTriangles(fl[i], t);
        out.insert(out.end(), t.begin(), t.end());
    }
}

// Shared body of initIntegrals / transformIntegral / smoothIndIntegrals:
// divergence-theorem moments over the triangulated cell surface.
void surfaceMoments(const PolyMesh& mesh, int cell, int maxOrder, const Mat3& JInv,
                    const Vec3& refPoint, const Vec3& centre, List3D& out) {
    out.resize(maxOrder + 1);
    std::vector<std::array<int32_t, 3>> tris;
    cellTriangles(mesh, cell, tris);
    for (const auto& tri : tris) {
        Vec3 v0 = transformPoint(JInv, mesh.points[tri[0]], refPoint);
        Vec3 v1 = transformPoint(JInv, mesh.points[tri[1]], refPoint);
        Vec3 v2 = transformPoint(JInv, mesh.points[tri[2]], refPoint);
        Vec3 vn = cross(v1 - v0, v2 - v0);
        double area = 0.5 * mag(vn);
        if (area < ROOTVSMALL) continue;
        if (dot(vn, v0 - centre) < 0.0)
            vn = (-1.0 / mag(vn)) * vn;
        else
            vn = (1.0 / mag(vn)) * vn;
        for (int n = 0; n <= maxOrder; ++n)
            for (int m = 0; m <= maxOrder; ++m)
                for (int l = 0; l <= maxOrder; ++l) {
                    if (n + m + l > maxOrder) continue;
                    if (n > 0)
                        out(n, m, l) += 1.0 / (n + 1) * area * vn.x *
                                        gaussQuad(n + 1, m, l, centre, v0, v1, v2);
                    else if (m > 0)
                        out(n, m, l) += 1.0 / (m + 1) * area * vn.y *
                                        gaussQuad(n, m + 1, l, centre, v0, v1, v2);
                    else
                        out(n, m, l) += 1.0 / (l + 1) * area * vn.z *
                                        gaussQuad(n, m, l + 1, centre, v0, v1, v2);
                }
    }
}
}  // namespace

void cellTriangleList(const PolyMesh& mesh, int cellI, std::vector<std::array<int32_t, 3>>& out) {
    cellTriangles(mesh, cellI, out);
}

double det2(double a, double d, double c, double b) {
    double cb = c * b;
    double err = std::fma(-c, b, cb);
    double dop = std::fma(a, d, -cb);
    return dop + err;
}

double det3(const Mat3& A) {
    return det2(A(0, 0), det2(A(1, 1), A(2, 2), A(1, 2), A(2, 1)), A(0, 1),
                det2(A(1, 0), A(2, 2), A(1, 2), A(2, 0))) +
           A(0, 2) * det2(A(1, 0), A(2, 1), A(1, 1), A(2, 0));
}

Mat3 inv3(const Mat3& A) {
    Mat3 R;
    R(0, 0) = det2(A(1, 1), A(2, 2), A(2, 1), A(1, 2));
    R(1, 0) = det2(A(1, 2), A(2, 0), A(1, 0), A(2, 2));
    R(2, 0) = det2(A(1, 0), A(2, 1), A(2, 0), A(1, 1));
    const double Det = A(0, 0) * R(0, 0) + A(0, 1) * R(1, 0) + A(0, 2) * R(2, 0);
    R(0, 1) = det2(A(0, 2), A(2, 1), A(0, 1), A(2, 2));
    R(0, 2) = det2(A(0, 1), A(1, 2), A(0, 2), A(1, 1));
    R(1, 1) = det2(A(0, 0), A(2, 2), A(0, 2), A(2, 0));
    R(1, 2) = det2(A(1, 0), A(0, 2), A(0, 0), A(1, 2));
    R(2, 1) = det2(A(2, 0), A(0, 1), A(0, 0), A(2, 1));
    R(2, 2) = det2(A(0, 0), A(1, 1), A(1, 0), A(0, 1));
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R(i, j) /= Det;
    return R;
}

void pivot(Mat3& A) {
    // mathFunctionsWENO.H:142-210. Note the second search scans the ORIGINAL A
    // (not the eliminated PA), as the reference does.
    Mat3 I;
    I(0, 0) = I(1, 1) = I(2, 2) = 1.0;
    int maxRow = 0;
    double Amax = -1e15;
    for (int i = 0; i < 3; ++i)
        if (std::fabs(A(i, 0)) > Amax) {
            maxRow = i;
            Amax = std::fabs(A(i, 0));
        }
    Mat3 P1 = I;
    for (int j = 0; j < 3; ++j) std::swap(P1(0, j), P1(maxRow, j));
    Amax = -1e15;
    for (int i = 1; i < 3; ++i)
        if (std::fabs(A(i, 1)) > Amax) {
            maxRow = i;
            Amax = std::fabs(A(i, 1));
        }
    if (maxRow != 1) {
        Mat3 P2 = I;
        for (int j = 0; j < 3; ++j) std::swap(P2(1, j), P2(2, j));
        P1 = matmul(P2, P1);
    }
    A = matmul(P1, A);
}

void eigenMagnitudes(const Mat3& A, double out[3]) {
    // characteristic polynomial l^3 - a l^2 + b l - c
    const double a = A(0, 0) + A(1, 1) + A(2, 2);
    const double b = (A(0, 0) * A(1, 1) - A(0, 1) * A(1, 0)) +
                     (A(0, 0) * A(2, 2) - A(0, 2) * A(2, 0)) +
                     (A(1, 1) * A(2, 2) - A(1, 2) * A(2, 1));
    const double c = det3(A);
    // depressed cubic t^3 + p t + q, l = t + a/3
    const double p = b - a * a / 3.0;
    const double q = -2.0 * a * a * a / 27.0 + a * b / 3.0 - c;
    const double disc = q * q / 4.0 + p * p * p / 27.0;
    if (disc > 0) {
        double sq = std::sqrt(disc);
        double t = std::cbrt(-q / 2.0 + sq) + std::cbrt(-q / 2.0 - sq);
        double l1 = t + a / 3.0;
        // remaining pair: product = c / l1 (complex conjugates => |l|^2)
        double m2 = std::fabs(l1) > 0 ? std::fabs(c / l1) : 0.0;
        out[0] = std::fabs(l1);
        out[1] = out[2] = std::sqrt(m2);
    } else {
        double r = std::sqrt(std::max(0.0, -p / 3.0));
        double arg = r > 0 ? std::max(-1.0, std::min(1.0, -q / (2.0 * r * r * r))) : 0.0;
        double phi = std::acos(arg);
        for (int k = 0; k < 3; ++k)
            out[k] = std::fabs(2.0 * r * std::cos((phi + 2.0 * M_PI * k) / 3.0) + a / 3.0);
    }
}

double cond(const Mat3& J) {
    double s[3];
    eigenMagnitudes(J, s);
    double mn = std::min(s[0], std::min(s[1], s[2]));
    double mx = std::max(s[0], std::max(s[1], s[2]));
    if (mn < ROOTVSMALL) return GREAT;
    return mx / mn;
}

Vec3 transformPoint(const Mat3& Jinv, const Vec3& xP, const Vec3& x0) {
    const double v0 = xP.x - x0.x, v1 = xP.y - x0.y, v2 = xP.z - x0.z;
    return {Jinv(0, 0) * v0 + Jinv(0, 1) * v1 + Jinv(0, 2) * v2,
            Jinv(1, 0) * v0 + Jinv(1, 1) * v1 + Jinv(1, 2) * v2,
            Jinv(2, 0) * v0 + Jinv(2, 1) * v1 + Jinv(2, 2) * v2};
}

double gaussQuad(int n, int m, int l, const Vec3& xi0, const Vec3& v0, const Vec3& v1,
                 const Vec3& v2) {
    double sum = 0.0;
    for (int j = 0; j < 13; ++j) {
        const double g0 = gaussCoeff[j][0], g1 = gaussCoeff[j][1];
        double xi = v0.x * (1 - g0 - g1) + v1.x * g0 + v2.x * g1 - xi0.x;
        double eta = v0.y * (1 - g0 - g1) + v1.y * g0 + v2.y * g1 - xi0.y;
        double zeta = v0.z * (1 - g0 - g1) + v1.z * g0 + v2.z * g1 - xi0.z;
        sum += gaussCoeff[j][2] * intPow(xi, n) * intPow(eta, m) * intPow(zeta, l);
    }
    return sum;
}

void initIntegrals(const PolyMesh& mesh, int cellI, int polOrder, List3D& volIntegrals,
                   Mat3& JInvI, Vec3& refPointI, double& refDetI) {
    // cell point labels in order of first appearance over the cell's faces
    std::vector<int32_t> pLabels;
    const int nf = mesh.cellNFaces(cellI);
    const int32_t* fl = mesh.cellFaceList(cellI);
    for (int i = 0; i < nf; ++i) {
        const int n = mesh.faceSize(fl[i]);
        const int32_t* p = mesh.face(fl[i]);
        for (int k = 0; k < n; ++k)
            if (std::find(pLabels.begin(), pLabels.end(), p[k]) == pLabels.end())
                pLabels.push_back(p[k]);
    }
    // in-cell edge neighbours of a point (mesh.pointPoints restricted to the cell,
    // geometryWENO.C:66-77)
    auto edgeNbrs = [&](int32_t pt) {
        std::vector<int32_t> nb;
        for (int i = 0; i < nf; ++i) {
            const int n = mesh.faceSize(fl[i]);
            const int32_t* p = mesh.face(fl[i]);
            for (int k = 0; k < n; ++k)
                if (p[k] == pt) {
                    int32_t a = p[(k + 1) % n], b = p[(k + n - 1) % n];
                    if (std::find(nb.begin(), nb.end(), a) == nb.end()) nb.push_back(a);
                    if (std::find(nb.begin(), nb.end(), b) == nb.end()) nb.push_back(b);
                }
        }
        return nb;
    };
    auto jacobiOf = [&](const std::array<int32_t, 4>& fr) {
        Mat3 J;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) J(i, j) = mesh.points[fr[j + 1]][i] - mesh.points[fr[0]][i];
        return J;
    };
    std::array<int32_t, 4> frame{-1, -1, -1, -1};
    double bestCond = 1e10;
    for (size_t k = 0; k < pLabels.size(); ++k) {
        std::vector<int32_t> nb = edgeNbrs(pLabels[k]);
        if (nb.size() < 3) continue;
        // OpenFOAM's pointPoints ordering is not reproducible without OpenFOAM.
        // Order the three frame edges by dominant axis so the reference axes
        // follow x,y,z (needed for the 2-D dimList convention, WENOBase.C:1147).
        std::array<int32_t, 4> fr{pLabels[k], nb[0], nb[1], nb[2]};
        std::array<int32_t, 3> bestPerm{0, 1, 2};
        double bestScore = -1;
        std::array<int32_t, 3> perm{0, 1, 2};
        do {
            double score = 1;
            for (int ax = 0; ax < 3; ++ax)
                score *= std::fabs(mesh.points[nb[perm[ax]]][ax] - mesh.points[pLabels[k]][ax]);
            if (score > bestScore * (1 + 1e-12)) {
                bestScore = score;
                bestPerm = perm;
            }
        } while (std::next_permutation(perm.begin(), perm.end()));
        for (int ax = 0; ax < 3; ++ax) fr[ax + 1] = nb[bestPerm[ax]];
        double cnd = cond(jacobiOf(fr));
        if (cnd < bestCond) {
            bestCond = cnd;
            frame = fr;
        } else if (frame[0] < 0 && k + 1 == pLabels.size()) {
            frame = fr;
        }
    }
    if (frame[0] < 0) throw std::runtime_error("initIntegrals: no reference frame for cell");
    refPointI = mesh.points[frame[0]];
    Mat3 J = jacobiOf(frame);
    JInvI = inv3(J);  // calculateInverseJacobi, geometryWENO.C:969-986 (non-LAPACK branch)
    pivot(JInvI);
    refDetI = det3(JInvI);
    if (std::fabs(refDetI) < SMALL) refDetI = refDetI < 0 ? -SMALL : SMALL;

    const Vec3 refPointTrans = transformPoint(JInvI, mesh.cellCentres[cellI], refPointI);
    surfaceMoments(mesh, cellI, polOrder, JInvI, refPointI, refPointTrans, volIntegrals);
    const double scale = 1.0 / (std::fabs(refDetI) * mesh.cellVolumes[cellI]);
    for (double& v : volIntegrals.d) v *= scale;
}

void transformIntegral(const PolyMesh& mesh, int cellJ, const Vec3& transCenterJ, int polOrder,
                       const Mat3& JInvI, const Vec3& refPointI, double refDetI,
                       List3D& transVolMom) {
    surfaceMoments(mesh, cellJ, polOrder, JInvI, refPointI, transCenterJ, transVolMom);
    const double scale = 1.0 / (std::fabs(refDetI) * mesh.cellVolumes[cellJ]);
    for (double& v : transVolMom.d) v *= scale;
}

void smoothIndIntegrals(const PolyMesh& mesh, int cellI, int polOrder, const Mat3& JInvI,
                        const Vec3& refPointI, List3D& out) {
    const int maxOrder = 2 * polOrder - 2;
    const Vec3 transCenterI = transformPoint(JInvI, mesh.cellCentres[cellI], refPointI);
    surfaceMoments(mesh, cellI, maxOrder, JInvI, refPointI, transCenterI, out);
}

namespace {
double Fac(int x) {
    if (x <= 0) return 1;
    double f = 1;
    for (int i = 2; i <= x; ++i) f *= i;
    return f;
}
inline double pos0(int x) { return x >= 0 ? 1.0 : 0.0; }
}  // namespace

std::vector<double> getB(const PolyMesh& mesh, int cellI, int polOrder, int nDvt,
                         const Mat3& JInvI, const Vec3& refPointI, const int dim[3]) {
    std::vector<double> B((size_t)nDvt * nDvt, 0.0);
    List3D intB;
    smoothIndIntegrals(mesh, cellI, polOrder, JInvI, refPointI, intB);
    int p = 0, q = 0;
    for (int n1 = 0; n1 <= dim[0]; ++n1)
        for (int m1 = 0; m1 <= dim[1]; ++m1)
            for (int l1 = 0; l1 <= dim[2]; ++l1) {
                if (!((n1 + m1 + l1) <= polOrder && (n1 + m1 + l1) > 0)) continue;
                for (int n2 = 0; n2 <= dim[0]; ++n2)
                    for (int m2 = 0; m2 <= dim[1]; ++m2)
                        for (int l2 = 0; l2 <= dim[2]; ++l2) {
                            if (!((n2 + m2 + l2) <= polOrder && (n2 + m2 + l2) > 0)) continue;
                            for (int lambda = 1; lambda <= polOrder; ++lambda)
                                for (int alpha = 0; alpha <= lambda; ++alpha)
                                    for (int beta = 0; beta <= lambda - alpha; ++beta) {
                                        int gamma = lambda - alpha - beta;
                                        double K =
                                            (pos0(n1 - alpha) * pos0(n2 - alpha) * pos0(m1 - beta) *
                                             pos0(m2 - beta) * pos0(l1 - gamma) * pos0(l2 - gamma) *
                                             Fac(n1) * Fac(m1) * Fac(l1) * Fac(n2) * Fac(m2) *
                                             Fac(l2)) /
                                            (Fac(n1 - alpha) * Fac(n2 - alpha) * Fac(m1 - beta) *
                                             Fac(m2 - beta) * Fac(l1 - gamma) * Fac(l2 - gamma));
                                        if (K != 0)
                                            B[(size_t)p * nDvt + q] +=
                                                K * intB(n1 + n2 - 2 * alpha, m1 + m2 - 2 * beta,
                                                         l1 + l2 - 2 * gamma);
                                    }
                            q++;
                        }
                p++;
                q = 0;
            }
    return B;
}

// One face of one cell (geometryWENO.C:772-869): integrals of the basis over the
// face in the cell's reference space, and the face's reference-space area.
void surfIntTransFace(const PolyMesh& mesh, int cellI, int f, int polOrder, const List3D& volI,
                      const Mat3& JInvI, const Vec3& refPointI, double* I, double& faceArea) {
    const int s = polOrder + 1;
    const Vec3 refPointTrans = transformPoint(JInvI, mesh.cellCentres[cellI], refPointI);
    std::vector<std::array<int32_t, 3>> tris;
    mesh.faceTriangles(f, tris);
    double area = 0;
    faceArea = 0;
    for (const auto& tri : tris) {
        Vec3 v0 = transformPoint(JInvI, mesh.points[tri[0]], refPointI);
        Vec3 v1 = transformPoint(JInvI, mesh.points[tri[1]], refPointI);
        Vec3 v2 = transformPoint(JInvI, mesh.points[tri[2]], refPointI);
        Vec3 vn = cross(v1 - v0, v2 - v0);
        area = 0.5 * mag(vn);
        faceArea += area;
        for (int n = 0; n <= polOrder; ++n)
            for (int m = 0; m <= polOrder; ++m)
                for (int l = 0; l <= polOrder; ++l)
                    if (n + m + l <= polOrder)
                        I[(size_t)n * s * s + m * s + l] +=
                            area * gaussQuad(n, m, l, refPointTrans, v0, v1, v2);
    }
    // geometryWENO.C:853-869: subtracts with the LAST triangle's area (the
    // reference's `area` is overwritten per triangle, not accumulated)
    for (int n = 0; n <= polOrder; ++n)
        for (int m = 0; m <= polOrder; ++m)
            for (int l = 0; l <= polOrder; ++l)
                if (n + m + l <= polOrder) I[(size_t)n * s * s + m * s + l] -= area * volI(n, m, l);
}

void surfIntTransCell(const PolyMesh& mesh, int cellI, int polOrder, const List3D& volI,
                      const Mat3& JInvI, const Vec3& refPointI, double* intBasTrans,
                      double* refFacAr) {
    const int s = polOrder + 1;
    const size_t cube = (size_t)s * s * s;
    const int nf = mesh.cellNFaces(cellI);
    const int32_t* fl = mesh.cellFaceList(cellI);
    for (int i = 0; i < nf; ++i) {
        const int f = fl[i];
        int side = (f < mesh.nInternalFaces && cellI == mesh.neighbour[f]) ? 1 : 0;
        double area = 0;
        surfIntTransFace(mesh, cellI, f, polOrder, volI, JInvI, refPointI,
                         intBasTrans + ((size_t)f * 2 + side) * cube, area);
        if (side == 0) refFacAr[f] += area;
    }
}

}  // namespace geom
}  // namespace wenoext
