// Column-block scan, chunked step, 29 elements per thread.
#define FSQ_E 29
#include "fsq_cscan.inc"
