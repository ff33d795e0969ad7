// mesh.hpp -- host-side tetrahedral mesh container + the synthetic unit-cube generator.
//
// Stands where Feel++'s Mesh<Simplex<3>> (feeldiscr/mesh.hpp:162) and unitCube() (feelfilters/unitcube.cpp:38-52)
// stand for the assembly path: only what the engine consumes is kept -- point ids/coordinates (GeoND::G(),
// feelmesh/geond.hpp:626) and element connectivity in Feel++'s local vertex order.  gmsh is not available in this
// environment, so "gmsh-generated" unit-cube meshes are replaced by a deterministic Kuhn (6 tets per cell) split.
#pragma once
#include <cstdint>
#include <memory>
#include <vector>

namespace feelpp_b200 {

struct Mesh {
    std::vector<double> xyz;    // [nPts][3]
    std::vector<int32_t> conn;  // [nElt][4], point ids
    int64_t numPoints() const { return (int64_t)xyz.size() / 3; }
    int64_t numElements() const { return (int64_t)conn.size() / 4; }
    int64_t numGlobalElements() const { return numElements(); }
};
using mesh_ptrtype = std::shared_ptr<Mesh>;

// Unit cube [0,1]^3, m^3 cells x 6 tets (Kuhn/Freudenthal split: conforming, deterministic).
//   jitter  : every interior vertex is displaced by U(-jitter*h, jitter*h)^3 (std::mt19937_64, seed 20261104)
//   shuffle : Fisher-Yates permutation of the element order (seed 7), emulating an unstructured mesher's numbering
// Point ids are the lexicographic grid index; all elements are positively oriented.
mesh_ptrtype unitCube(int m, double jitter = 0.0, bool shuffle = false);

// boundaryfaces(mesh) (feelmesh/filters.hpp): the faces that belong to exactly one element, as (element, local face) pairs
// in element order; local face f is opposite to local vertex f (feelmesh/simplexordering.hpp:341-347).
void boundaryFaces(int64_t nPts, int64_t nElt, const int32_t* conn, std::vector<int32_t>& elt, std::vector<int32_t>& face);

// element range: elements(mesh) (feelmesh/filters.hpp:201) -- the only range kind on this path
struct ElementRange {
    mesh_ptrtype mesh;
};
inline ElementRange elements(mesh_ptrtype const& mesh) { return ElementRange{mesh}; }

} // namespace feelpp_b200
