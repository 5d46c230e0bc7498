// Column-block scan, chunked step, 31 elements per thread.
#define FSQ_E 31
#include "fsq_cscan.inc"
