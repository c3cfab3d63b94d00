#include "common.cuh"
extern "C" int ccn_version(void) { return 100; }
