// Column-block scan, chunked step, 33 elements per thread.
#define FSQ_E 33
#include "fsq_cscan.inc"
