// Column-block scan, chunked step, 13 elements per thread.
#define FSQ_E 13
#include "fsq_cscan.inc"
