// doftable.hpp -- host DofTable for continuous Lagrange Pk (k=1,2,3) on tetrahedra, scalar or vector.
//
// Mirrors what Pch<k>(mesh)/Pchv<k>(mesh) (feeldiscr/pch.hpp:85-90, pchv.hpp:43-48) leave behind for the assembly
// path: the local-to-global table returned by DofTable::localToGlobalIndices (feeldiscr/doftable.hpp:732-750).
// The numbering reproduces DofTable::build (doftable.hpp:1802-2070): dofs are created at first touch while
// walking the elements in mesh order, inside an element vertices -> edges -> faces -> volume
// (feeldiscr/doffromelement.hpp:492-530), edge dofs reversed when the edge runs from the higher to the lower point
// id (feelmesh/mesh3d.hpp:755-782); product spaces interleave components, global = ncomp*node + c
// (doftable.hpp:1289-1299).  Built with flat open-addressing tables instead of the reference's node-based maps.
#pragma once
#include <cstdint>
#include <memory>
#include <vector>
#include "mesh.hpp"

namespace feelpp_b200 {

inline int nLocalDofTet(int order) { return (order + 1) * (order + 2) * (order + 3) / 6; }

struct DofTable {
    int order = 1, ncomp = 1, n = 4, nd = 4;
    int64_t nDof = 0;                 // nLocalDofWithGhost == nLocalDofWithoutGhost on one process
    std::vector<int32_t> l2g;         // [nElt][nd], local index = n*c + i
    const int32_t* localToGlobalIndices(int64_t elt) const { return l2g.data() + elt * nd; }
    void build(Mesh const& mesh, int order, int ncomp);
};

// local entity numbering of the reference tetrahedron (feelmesh/simplexordering.hpp:283-291, 341-347)
extern const int kEdgeToPoint[6][2];
extern const int kFaceToPoint[4][3];

} // namespace feelpp_b200
