#include "common.cuh"
unsigned long long g_ccn_kernel_launches = 0;
extern "C" int ccn_version(void) { return 100; }
extern "C" int64_t ccn_kernel_launches(void) { return (int64_t)g_ccn_kernel_launches; }
