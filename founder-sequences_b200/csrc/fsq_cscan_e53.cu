// Column-block scan, chunked step, 53 elements per thread.
#define FSQ_E 53
#include "fsq_cscan.inc"
