#pragma once
#include "accessors/c_grid.h"
