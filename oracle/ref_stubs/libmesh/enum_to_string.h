// ref_stubs/libmesh/enum_to_string.h — STUB: only used inside error messages.
#ifndef REF_STUB_ENUM_TO_STRING_H
#define REF_STUB_ENUM_TO_STRING_H
#include <string>
namespace libMesh { namespace Utility { template <typename T> inline std::string enum_to_string(const T v) { return std::to_string((int)v); } } }
#endif
