// hex_sym4.cu -- assembly variant 7: the symmetric pair-block kernel of hex_sym2.cu with FOUR threads per cell (78 / 72
// accumulators, 255 registers: 5 FMA per loaded double, so the FMA phase runs at the FP64 pipe's peak instead of being
// bound by operand wavefronts) on a half-size tile (7 x 3 node positions = 8 x 4 cells, 128 threads, <= 113 KB shared
// memory) with TWO CTAs per SM, so that the FP64-bound phases of one CTA overlap the LSU-bound gather of the other.
#define FEMX_SYM_NX 7
#define FEMX_SYM_NY 3
#define FEMX_SYM_ELP 32
#define FEMX_SYM_CTAS 2
#define FEMX_SYM_TPC 4
#define FEMX_SYM_KERNEL hex_sym4_kernel
#define FEMX_SYM_LAUNCH launch_hex_sym4
#include "hex_sym2.cu"
