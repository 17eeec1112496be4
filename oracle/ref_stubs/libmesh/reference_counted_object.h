// ref_stubs/libmesh/reference_counted_object.h — STUB: the leak-counting base class of include/base/reference_counted_object.h
// (debug bookkeeping only) as an empty base.
#ifndef REF_STUB_REFERENCE_COUNTED_OBJECT_H
#define REF_STUB_REFERENCE_COUNTED_OBJECT_H
namespace libMesh { template <typename T> class ReferenceCountedObject {}; }
#endif
