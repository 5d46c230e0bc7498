// Blocked column-block scan kernels with 33 elements per thread; see fsq_bscan.inc.
#define FSQ_E 33
#include "fsq_bscan.inc"
