// Stand-in for scitbx/array_family/ref_reductions.h: baseline.cpp uses nothing from it.
#pragma once
#include "accessors/c_grid.h"
