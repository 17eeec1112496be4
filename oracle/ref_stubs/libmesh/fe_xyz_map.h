// ref_stubs/libmesh/fe_xyz_map.h — STUB: the XYZ family's map is outside this path; FEMap::build names the class.
#ifndef REF_STUB_FE_XYZ_MAP_H
#define REF_STUB_FE_XYZ_MAP_H
#include "libmesh/fe_map.h"
namespace libMesh { class FEXYZMap : public FEMap {}; }
#endif
