// Column-block scan, chunked step, 21 elements per thread.
#define FSQ_E 21
#include "fsq_cscan.inc"
