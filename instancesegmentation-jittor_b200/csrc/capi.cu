// capi.cu -- library identification entry points of libroipath_b200.so.
#include "common.cuh"

#if !defined(__CUDA_ARCH_LIST__)
#define __CUDA_ARCH_LIST__ 0
#endif

extern "C" {

int rp_version(void) { return RP_ABI_VERSION; }

// The library is built with -gencode arch=compute_100a,code=sm_100a only.
int rp_built_for_sm(void) { return 100; }

}  // extern "C"
