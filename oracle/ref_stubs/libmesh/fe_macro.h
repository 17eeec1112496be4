// ref_stubs/libmesh/fe_macro.h — STUB: no subdivision-surface maps are instantiated.
#define INSTANTIATE_SUBDIVISION_MAPS
