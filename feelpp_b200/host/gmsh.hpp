// gmsh.hpp -- reader of Gmsh .msh files (tetrahedra + marked boundary triangles), the data format on the caller side
// of the assembly path.  Stands where ImporterGmsh<MeshType>::readFromFile stands (feelfilters/importergmsh.hpp:983-1103:
// $MeshFormat / $PhysicalNames, then readFromFileVersion2 :1106-1420 or readFromFileVersion4): MSH 2.x and MSH 4.1, ASCII and
// binary.  As in the reference the node numbers of the file are the point ids ("we do not renumber anymore",
// :1262-1266); here they are compacted to 0..nPts-1 in ascending id order, which keeps every id comparison the DofTable
// makes (edge orientation, feelmesh/mesh3d.hpp:755-782).  The physical tag is the marker (elementary tag on request,
// M_use_elementary_region_as_physical_region :1268-1271).  Lines, points and higher-order elements are refused or
// skipped: the engine's path is order-1 tetrahedra.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace feelpp_b200 {

struct GmshMesh {
    double version = 0;
    bool binary = false;
    std::vector<double> xyz;           // [nPts][3]
    std::vector<int64_t> pointId;      // [nPts] node numbers of the file, ascending
    std::vector<int32_t> conn;         // [nElt][4] compact point indices, file order
    std::vector<int32_t> eltMarker;    // [nElt]
    std::vector<int32_t> eltPartition; // [nElt] 0-based partition tag of the file (MSH 2.x: tags numPartitions, partition, -ghosts;
                                       // importergmsh.hpp:1239-1242, 1337-1340), 0 when the file is not partitioned
    int32_t numPartitions = 1;         // 1 + the largest partition tag met
    std::vector<int32_t> tri;          // [nTri][3] compact point indices (marked faces), file order
    std::vector<int32_t> triMarker;    // [nTri]
    std::string physicalNames;         // "dim tag name\n" per entry
};

// throws std::runtime_error with the reason (bad section, unknown node, unsupported element type or version)
GmshMesh readGmsh(std::string const& path, bool elementaryAsPhysical = false);

} // namespace feelpp_b200
