// capi.cu -- library identification entry points of libroipath_b200.so and the knob table.
#include <string.h>

#include <mutex>

#include "common.cuh"

#if !defined(__CUDA_ARCH_LIST__)
#define __CUDA_ARCH_LIST__ 0
#endif

namespace {
struct RpKnob {
  char name[40];
  int value;
};
std::mutex knob_mu;
RpKnob knobs[64];
int n_knobs = 0;
}  // namespace

int rp_env_int(const char* name, int dflt) {
  std::lock_guard<std::mutex> lock(knob_mu);
  for (int i = 0; i < n_knobs; ++i)
    if (strcmp(knobs[i].name, name) == 0) return knobs[i].value == INT32_MIN ? dflt : knobs[i].value;
  const char* s = getenv(name);
  const int v = (s && *s) ? atoi(s) : INT32_MIN;  // INT32_MIN marks "not set": the call site's default applies
  if (n_knobs < 64 && strlen(name) < sizeof(knobs[0].name)) {
    strcpy(knobs[n_knobs].name, name);
    knobs[n_knobs].value = v;
    ++n_knobs;
  }
  return v == INT32_MIN ? dflt : v;
}

namespace {
struct RpSmemEntry {
  const void* kernel;
  int device, bytes;
};
std::mutex smem_mu;
RpSmemEntry smem_set[128];
int n_smem = 0;
}  // namespace

int rp_smem_ensure(const void* kernel, int bytes) {
  if (bytes <= 0) return 0;  // (static + dynamic may exceed the 48 KB default even when dynamic alone does not)
  int dev = 0;
  (void)cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(smem_mu);
  RpSmemEntry* e = nullptr;
  for (int i = 0; i < n_smem; ++i)
    if (smem_set[i].kernel == kernel && smem_set[i].device == dev) e = &smem_set[i];
  if (e && e->bytes >= bytes) return 0;
  const cudaError_t rc = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (rc != cudaSuccess) {
    (void)cudaGetLastError();
    return (int)rc;
  }
  if (!e && n_smem < 128) {
    e = &smem_set[n_smem++];
    e->kernel = kernel;
    e->device = dev;
  }
  if (e) e->bytes = bytes;
  return 0;
}

extern "C" {

int rp_version(void) { return RP_ABI_VERSION; }

// The library is built with -gencode arch=compute_100a,code=sm_100a only.
int rp_built_for_sm(void) { return 100; }

void rp_reload_env(void) {
  std::lock_guard<std::mutex> lock(knob_mu);
  n_knobs = 0;
}

}  // extern "C"
