// Stand-in for dials/error.h (see ../README.md).  Not DIALS code.
#pragma once
#include <stdexcept>
#include <string>
#define DIALS_ASSERT(cond)                                                                       \
    do {                                                                                         \
        if (!(cond)) throw std::runtime_error(std::string("DIALS_ASSERT(" #cond ") failed"));   \
    } while (0)
