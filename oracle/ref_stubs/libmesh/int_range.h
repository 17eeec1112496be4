// ref_stubs/libmesh/int_range.h — STUB: index_range / make_range live in this directory's libmesh_common.h.
#include "libmesh/libmesh_common.h"
