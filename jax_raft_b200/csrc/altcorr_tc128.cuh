// The M = 128 / A-in-TMEM form of the tensor-core alt_corr lookup (altcorr_tc128.cu), launched by
// raftk_altcorr_lookup_tc (altcorr_tc.cu) after it has validated the arguments.
#pragma once

namespace raftk {

int altcorr_tc128_launch(const float* fmap1, const float* const* f2_levels, const void* f1_split, const void* const* f2_split,
                         int num_levels, const float* coords, float* out, int B, int nq, int h, int w, int D, int radius,
                         void* stream);

// 0: choose by problem size, 1: the M = 64 kernel (altcorr_tc.cu), 2: the M = 128 kernel; raftk_set_altcorr_mode
int& altcorr_mode_flag();

}  // namespace raftk
