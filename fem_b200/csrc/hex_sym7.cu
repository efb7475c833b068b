// hex_sym7.cu -- assembly variant 13 (diagnostic): variant 12 (hex_sym6.cu) with per-phase clock64 accounting
// (FEMX_SYM_PROF in hex_sym2.cu): thread 0 of every CTA accumulates the cycles between the barriers of a layer step and the
// launcher prints the averages to stderr.  Results are those of variant 12; the timing is not a bench number.
#define FEMX_SYM_NX 8
#define FEMX_SYM_NY 6
#define FEMX_SYM_ELP 64
#define FEMX_SYM_CTAS 1
#define FEMX_SYM_TPC 8
#define FEMX_SYM_GPERM 1
#define FEMX_SYM_PROF 1
#define FEMX_SYM_IPU 1
#define FEMX_SYM_KERNEL hex_sym7_kernel
#define FEMX_SYM_LAUNCH launch_hex_sym7
#include "hex_sym2.cu"
