// Stand-in for scitbx/array_family/accessors/c_grid.h (see ../../../README.md).  Not cctbx code.
#pragma once
#include <cstddef>
#include <vector>

#include "../tiny_types.h"

namespace scitbx { namespace af {
struct trivial_accessor {
    std::size_t n = 0;
    trivial_accessor() = default;
    trivial_accessor(std::size_t n_) : n(n_) {}
    std::size_t size_1d() const { return n; }
};

template <std::size_t Nd>
class c_grid {
    std::size_t d_[Nd] = {};

  public:
    c_grid() = default;
    c_grid(std::size_t n0, std::size_t n1) {
        static_assert(Nd == 2, "two-dimensional grid");
        d_[0] = n0, d_[1] = n1;
    }
    template <typename T>
    c_grid(const tiny<T, Nd>& t) {
        for (std::size_t i = 0; i < Nd; ++i) d_[i] = (std::size_t)t[i];
    }
    std::size_t operator[](std::size_t i) const { return d_[i]; }
    std::size_t size_1d() const {
        std::size_t n = 1;
        for (std::size_t i = 0; i < Nd; ++i) n *= d_[i];
        return n;
    }
    std::size_t operator()(std::size_t j, std::size_t i) const { return j * d_[1] + i; }
    template <class Other>
    bool all_eq(const Other& o) const {
        for (std::size_t i = 0; i < Nd; ++i)
            if (!((long long)d_[i] == (long long)o[i])) return false;
        return true;
    }
};

template <typename T, typename Accessor = trivial_accessor>
class const_ref {
    const T* p_ = nullptr;
    Accessor a_;

  public:
    const_ref() = default;
    const_ref(const T* p, const Accessor& a) : p_(p), a_(a) {}
    const Accessor& accessor() const { return a_; }
    std::size_t size() const { return a_.size_1d(); }
    const T& operator[](std::size_t k) const { return p_[k]; }
    const T& operator()(std::size_t j, std::size_t i) const { return p_[a_(j, i)]; }
    const T* begin() const { return p_; }
    const T& front() const { return p_[0]; }
};

// A writable view IS a read-only view (scitbx derives ref from const_ref the same way), so that a
// ref binds to `const const_ref<T, A>&` parameters with T still deduced.
template <typename T, typename Accessor = trivial_accessor>
class ref : public const_ref<T, Accessor> {
    typedef const_ref<T, Accessor> base;

  public:
    ref() = default;
    ref(T* p, const Accessor& a) : base(p, a) {}
    T& operator[](std::size_t k) const { return const_cast<T&>(base::operator[](k)); }
    T& operator()(std::size_t j, std::size_t i) const { return const_cast<T&>(base::operator()(j, i)); }
    T* begin() const { return const_cast<T*>(base::begin()); }
    T& front() const { return const_cast<T&>(base::front()); }
};

template <typename T, typename Accessor = trivial_accessor>
class versa {
    std::vector<T> v_;
    Accessor a_;

  public:
    versa(const Accessor& a, const T& init) : v_(a.size_1d(), init), a_(a) {}
    T& operator[](std::size_t k) { return v_[k]; }
    const T& operator[](std::size_t k) const { return v_[k]; }
    af::ref<T, Accessor> ref() { return af::ref<T, Accessor>(v_.data(), a_); }
    const Accessor& accessor() const { return a_; }
};
}}  // namespace scitbx::af
