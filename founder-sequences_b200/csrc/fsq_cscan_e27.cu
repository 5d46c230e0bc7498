// Column-block scan, chunked step, 27 elements per thread.
#define FSQ_E 27
#include "fsq_cscan.inc"
