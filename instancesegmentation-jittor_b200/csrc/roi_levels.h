// roi_levels.h -- the per-level feature-map table shared by the ROIAlign / Pooler translation units.
#pragma once
#include "common.cuh"

struct RtLevels {
  const float* feat[RP_MAX_LEVELS];
  float* grad[RP_MAX_LEVELS];
  int H[RP_MAX_LEVELS];
  int W[RP_MAX_LEVELS];
  float scale[RP_MAX_LEVELS];
  int L;
};

// roi_fwd_tc.cu: tensor-core forward (TMA tiles -> SIMT y-pass -> tcgen05 x-pass).  Returns
// RP_OK when it launched, 1 when the configuration is not eligible (the caller launches the SIMT
// kernel instead), or an error code.
int rx_forward(const RtLevels& lv, int N, int C, const float* rois, const int* levels, int K, int PH,
               int PW, float sampling_ratio, float* out, cudaStream_t st);

// roi_nhwc.cu: channels-last forward (features [N,H,W,C]); output [K,C,PH,PW].
int rn_forward_nhwc(const RtLevels& lv, int N, int C, const float* rois, const int* levels, int K, int PH,
                    int PW, float sampling_ratio, float* out, cudaStream_t st);
