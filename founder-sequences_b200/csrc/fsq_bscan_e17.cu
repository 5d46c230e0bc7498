// Blocked column-block scan kernels with 17 elements per thread; see fsq_bscan.inc.
#define FSQ_E 17
#include "fsq_bscan.inc"
