// Minimal polyhedral mesh in OpenFOAM's polyMesh conventions (host, setup only).
//
// The reference consumes an OpenFOAM fvMesh (points, faces, owner, neighbour,
// cells(), cellCells(), C(), cellVolumes(), boundary patches). OpenFOAM is not
// available, so this is the stand-in the WENOBase restatement (weno_setup.cpp)
// runs on. Conventions kept: internal faces first in upper-triangular order,
// owner < neighbour, boundary faces grouped per patch, face normal points from
// owner to neighbour.
#pragma once
#include <array>
#include <cstdint>
#include <functional>
#include <string>
#include <vector>

namespace wenoext {

struct Vec3 {
    double x = 0, y = 0, z = 0;
    double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(double s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
double mag(Vec3 a);

// Patch kinds the hot path distinguishes (WENOUpwindFit.C:433 coupled(),
// :478 processorFvPatch, :493 cyclicFvPatch; everything else gets zero).
enum PatchType : int32_t {
    PATCH_GENERIC = 0,   // wall / patch / inlet ...: not coupled, zero correction
    PATCH_PROCESSOR = 1, // coupled, neighbour rank in `nbr`
    PATCH_CYCLIC = 2,    // coupled, neighbour patch index in `nbr`
    PATCH_EMPTY = 3      // 2-D front/back: zero-size fields in OpenFOAM
};

struct Patch {
    std::string name;
    int32_t type = PATCH_GENERIC;
    int32_t start = 0;
    int32_t size = 0;
    int32_t nbr = -1;
};

struct PolyMesh {
    std::vector<Vec3> points;
    std::vector<int32_t> faceStart;  // CSR over faces -> facePoints
    std::vector<int32_t> facePoints;
    std::vector<int32_t> owner;      // [nFaces]
    std::vector<int32_t> neighbour;  // [nInternalFaces]
    std::vector<Patch> patches;
    int32_t nCells = 0;
    int32_t nInternalFaces = 0;
    // derived
    std::vector<int32_t> cellFaceStart, cellFaces;  // mesh.cells()
    std::vector<int32_t> cellCellStart, cellCells;  // mesh.cellCells()
    std::vector<Vec3> faceCentres, faceAreas, cellCentres;
    std::vector<double> cellVolumes;
    int32_t solutionD[3] = {1, 1, 1};  // mesh.solutionD(): -1 -> 0 for empty dirs

    int32_t nFaces() const { return (int32_t)owner.size(); }
    int32_t nSolutionD() const { return solutionD[0] + solutionD[1] + solutionD[2]; }
    int faceSize(int f) const { return faceStart[f + 1] - faceStart[f]; }
    const int32_t* face(int f) const { return &facePoints[faceStart[f]]; }
    int cellNFaces(int c) const { return cellFaceStart[c + 1] - cellFaceStart[c]; }
    const int32_t* cellFaceList(int c) const { return &cellFaces[cellFaceStart[c]]; }

    // Build cells(), cellCells(), centres, volumes from the primary arrays.
    void finalize();
    // Fan triangles (base = face point 0) of one face: the stand-in for
    // polyMeshTetDecomposition::faceTetIndices(...).faceTriIs(mesh).
    void faceTriangles(int f, std::vector<std::array<int32_t, 3>>& tris) const;
};

// Side specification of a box: which patch each of the 6 sides belongs to.
struct BoxSpec {
    int32_t n[3] = {1, 1, 1};
    double lo[3] = {0, 0, 0};
    double hi[3] = {1, 1, 1};
    // side order: xlo, xhi, ylo, yhi, zlo, zhi.
    // kind: PATCH_GENERIC, PATCH_EMPTY or PATCH_CYCLIC (cyclic must be set on
    // both sides of an axis).
    int32_t sideType[6] = {0, 0, 0, 0, 0, 0};
    // 0: hexes; 6: every hex split into 6 tets (Kuhn); used for irregular stencils
    int32_t tetSplit = 0;
    // If non-zero: the box is the first n[a] layers of a box with nDen[a] cells
    // between lo and hi (point i sits at lo + (hi-lo)*i/nDen, bit-identical to the
    // larger box). Used for the representative box of box_fast.cpp.
    int32_t nDen[3] = {0, 0, 0};
};

// blockMesh-like structured hex box, cells numbered i-fastest.
PolyMesh makeBoxMesh(const BoxSpec& spec);

// Generic builder: cells given as lists of faces (each a point list, outward
// orientation not required). `boundaryPatchOf(face points)` assigns boundary
// faces to a patch index.
struct CellShape {
    std::vector<std::vector<int32_t>> faces;
};
PolyMesh buildFromCells(const std::vector<Vec3>& points,
                        const std::vector<CellShape>& cells,
                        const std::vector<Patch>& patchProto,
                        const std::function<int(const std::vector<int32_t>&)>& patchOf);

// `decomposePar simple`-like block decomposition of a box: cell -> rank.
std::vector<int32_t> decomposeBox(const BoxSpec& spec, const int32_t nProc[3]);

// Rank-local mesh cut out of the global mesh (processor patches appended after
// the physical patches, ordered by neighbour rank; faces inside a processor
// patch ordered by global face index so both sides agree).
struct LocalMesh {
    PolyMesh mesh;
    std::vector<int32_t> localToGlobalCell;
    std::vector<int32_t> localToGlobalFace;
};
LocalMesh extractLocalMesh(const PolyMesh& global, const std::vector<int32_t>& cellProc,
                           int32_t myProc, std::vector<int32_t>& globalToLocalCell);

}  // namespace wenoext
