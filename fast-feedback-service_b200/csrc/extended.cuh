// extended.cuh -- dispersion_extended kernels (placeholder, filled in below)
#pragma once
#include "threshold.cuh"
