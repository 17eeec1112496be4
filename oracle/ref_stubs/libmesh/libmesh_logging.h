// ref_stubs/libmesh/libmesh_logging.h — STUB: LOG_SCOPE is a no-op (this directory's libmesh_common.h).
#include "libmesh/libmesh_common.h"
