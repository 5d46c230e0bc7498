// Column-block scan, chunked step, 17 elements per thread.
#define FSQ_E 17
#include "fsq_cscan.inc"
