// Column-block scan, chunked step, 25 elements per thread.
#define FSQ_E 25
#include "fsq_cscan.inc"
