// The split schedule's kernels (pp_hands.cu, MODE = PP_SPLIT_*) and their launcher, as a translation unit of their own so
// that they compile beside the whole-step kernels.
#define PP_HANDS_SPLIT_TU 1
#include "pp_hands.cu"
