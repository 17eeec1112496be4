// ref_stubs/libmesh/mesh_subdivision_support.h — STUB: subdivision surfaces are outside this path.
#ifndef REF_STUB_MESH_SUBDIVISION_SUPPORT_H
#define REF_STUB_MESH_SUBDIVISION_SUPPORT_H
#include "libmesh/elem.h"
namespace libMesh {
class Tri3Subdivision : public Elem {};
namespace MeshTools { namespace Subdivision { inline void find_one_ring(const Tri3Subdivision*, std::vector<const Node*>&) { libmesh_not_implemented(); } } }
}  // namespace libMesh
#endif
