/* No-op stand-in for {fmt}: enough surface for the reference headers to compile. */
#pragma once
#include <cstddef>
#include <string>
namespace fmt {
enum class terminal_color {
    black, red, green, yellow, blue, magenta, cyan, white,
    bright_black, bright_red, bright_green, bright_yellow, bright_blue, bright_magenta, bright_cyan, bright_white
};
enum class emphasis { bold = 1, faint = 2, italic = 4, underline = 8 };
struct text_style {
    constexpr text_style() {}
};
constexpr text_style fg(terminal_color) { return {}; }
constexpr text_style bg(terminal_color) { return {}; }
constexpr text_style operator|(text_style, emphasis) { return {}; }
constexpr text_style operator|(text_style, text_style) { return {}; }
template <class T> const T& runtime(const T& s) { return s; }
template <class T> const T& styled(const T& v, text_style) { return v; }
/* format(): returns the format string itself (arguments dropped) so error texts stay readable. */
template <class... A> std::string format(const char* f, A&&...) { return f; }
template <class... A> std::string format(const std::string& f, A&&...) { return f; }
template <class... A> std::string format(text_style, const char* f, A&&...) { return f; }
template <class... A> void print(A&&...) {}
template <class... A> void println(A&&...) {}
template <class... A> std::size_t formatted_size(A&&...) { return 0; }
struct ostream_stub {
    template <class... A> void print(A&&...) {}
    void close() {}
};
template <class... A> ostream_stub output_file(A&&...) { return {}; }
}  // namespace fmt
