// capi.cu -- version entry point of libadkf_b200.so (the other extern "C" symbols live beside their kernels).
#include "../../include/adkf_b200.h"

extern "C" int adkf_version(void) { return 100; }  // 0.1.0 -> major*10000 + minor*100 + patch
