// Blocked column-block scan kernels with 9 elements per thread; see fsq_bscan.inc.
#define FSQ_E 9
#include "fsq_bscan.inc"
