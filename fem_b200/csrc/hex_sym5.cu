// hex_sym5.cu -- assembly variant 8: hex_sym2.cu with FOUR threads per cell (78 / 72 accumulators, 255 registers) on the
// full 8 x 6 tile, one CTA of 8 warps per SM: half the operand wavefronts of variant 5 in the FMA phase, half the warps.
#define FEMX_SYM_NX 8
#define FEMX_SYM_NY 6
#define FEMX_SYM_ELP 64
#define FEMX_SYM_CTAS 1
#define FEMX_SYM_TPC 4
#define FEMX_SYM_KERNEL hex_sym5_kernel
#define FEMX_SYM_LAUNCH launch_hex_sym5
#include "hex_sym2.cu"
