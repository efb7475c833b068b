// hex_sym6.cu -- assembly variant 12: hex_sym2.cu (full 8 x 6 tile, 16 warps, eight threads per cell) with the gather
// lanes permuted inside each 27-value run: the 15 values that sum 2..4 cell contributions occupy the lower half-warp,
// the 12 single-contribution values the upper one, so that the 2nd..4th shared loads of a value touch one half-warp
// (one wavefront instead of two; 54 instead of 75 wavefronts per node and step), and the three f sums are reduced
// with 8 instead of 15 shuffles.
#define FEMX_SYM_NX 8
#define FEMX_SYM_NY 6
#define FEMX_SYM_ELP 64
#define FEMX_SYM_CTAS 1
#define FEMX_SYM_TPC 8
#define FEMX_SYM_GPERM 1
#define FEMX_SYM_IPU 1
#define FEMX_SYM_KERNEL hex_sym6_kernel
#define FEMX_SYM_LAUNCH launch_hex_sym6
#include "hex_sym2.cu"
