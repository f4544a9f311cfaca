// Stand-in for scitbx/array_family/tiny_types.h (see ../../README.md).  Not cctbx code.
#pragma once
#include <cstddef>

namespace scitbx { namespace af {
template <typename T, std::size_t N>
class tiny {
    T v_[N] = {};

  public:
    tiny() = default;
    tiny(T a, T b) {
        static_assert(N == 2, "two-element constructor");
        v_[0] = a, v_[1] = b;
    }
    T& operator[](std::size_t i) { return v_[i]; }
    const T& operator[](std::size_t i) const { return v_[i]; }
    static constexpr std::size_t size() { return N; }
    bool all_gt(T x) const {
        for (std::size_t i = 0; i < N; ++i)
            if (!(v_[i] > x)) return false;
        return true;
    }
    template <class Other>
    bool all_eq(const Other& o) const {
        for (std::size_t i = 0; i < N; ++i)
            if (!((long long)v_[i] == (long long)o[i])) return false;
        return true;
    }
};
typedef tiny<int, 2> int2;
}}  // namespace scitbx::af
