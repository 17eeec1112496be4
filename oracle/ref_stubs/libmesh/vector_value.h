// ref_stubs/libmesh/vector_value.h — STUB: RealVectorValue / RealGradient are this directory's Point.
#include "libmesh/point.h"
