// cc.cuh -- connected components kernels (placeholder)
#pragma once
#include <stdint.h>
namespace ffsb { struct SpotAcc { unsigned long long v[8]; }; }
