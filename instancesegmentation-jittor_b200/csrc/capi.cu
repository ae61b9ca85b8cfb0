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

extern "C" {

int rp_version(void) { return RP_ABI_VERSION; }

// The library is built with -gencode arch=compute_100a,code=sm_100a only.
int rp_built_for_sm(void) { return 100; }

void rp_reload_env(void) {
  std::lock_guard<std::mutex> lock(knob_mu);
  n_knobs = 0;
}

}  // extern "C"
