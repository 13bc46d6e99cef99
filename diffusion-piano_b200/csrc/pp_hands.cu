// placeholder - replaced below
#include "pp_env.cuh"
void pp_launch_hands(const HandsArgs&, cudaStream_t) {}
size_t pp_hands_smem_bytes() { return 0; }
