// hex_sym3.cu -- assembly variant 6: the symmetric pair-block kernel of hex_sym2.cu with a half-size tile
// (7 x 3 node positions = 8 x 4 cells, 256 threads, <= 113 KB shared memory) so that TWO CTAs share an SM.  The two
// CTAs drift out of phase, which overlaps the FP64-bound phases (geometry, FMA, G' -> K) of one with the
// LSU-bound gather of the other -- at the price of more halo cells per owned node (1.52 instead of 1.31).
#define FEMX_SYM_NX 7
#define FEMX_SYM_NY 3
#define FEMX_SYM_ELP 32
#define FEMX_SYM_CTAS 2
#define FEMX_SYM_TPC 8
#define FEMX_SYM_KERNEL hex_sym3_kernel
#define FEMX_SYM_LAUNCH launch_hex_sym3
#include "hex_sym2.cu"
