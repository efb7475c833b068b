// hex_mma_prof.cu -- assembly variant 21: hex_mma.cu with per-phase clock accounting (diagnostic; prints to stderr)
#define FEMX_MMA_PROF 1
#define FEMX_MMA_KERNEL hex_mma_prof_kernel
#define FEMX_MMA_LAUNCH launch_hex_mma_prof
#include "hex_mma.cu"
