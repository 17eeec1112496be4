// ref_stubs/libmesh/tensor_value.h — STUB: named by an include line of the reference's shape sources, nothing of it is used
// by the LAGRANGE functions that are called.
#ifndef REF_STUB_TENSOR_VALUE_H
#define REF_STUB_TENSOR_VALUE_H
#include "libmesh/point.h"
#endif
