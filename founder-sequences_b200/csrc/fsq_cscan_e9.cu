// Column-block scan, chunked step, 9 elements per thread.
#define FSQ_E 9
#include "fsq_cscan.inc"
