/* No-op stand-in for spdlog: a logger whose calls compile to nothing. */
#pragma once
#include <cstddef>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
namespace spdlog {
namespace level {
enum level_enum { trace, debug, info, warn, err, critical, off };
inline level_enum from_str(const std::string&) { return info; }
}  // namespace level
struct spdlog_ex : std::runtime_error {
    using std::runtime_error::runtime_error;
};
namespace sinks {
struct sink {
    virtual ~sink() = default;
    void set_pattern(const std::string&) {}
};
struct stdout_color_sink_mt : sink {};
struct stdout_sink_mt : sink {};
struct rotating_file_sink_mt : sink {
    template <class... A> rotating_file_sink_mt(A&&...) {}
};
}  // namespace sinks
using sink_ptr = std::shared_ptr<sinks::sink>;
class logger {
  public:
    virtual ~logger() = default;
    void set_level(level::level_enum) {}
    void flush() {}
    template <class... A> void trace(A&&...) {}
    template <class... A> void debug(A&&...) {}
    template <class... A> void info(A&&...) {}
    template <class... A> void warn(A&&...) {}
    template <class... A> void error(A&&...) {}
    template <class... A> void critical(A&&...) {}
};
struct thread_pool_stub {};
enum class async_overflow_policy { block, overrun_oldest };
class async_logger : public logger {
  public:
    template <class... A> async_logger(A&&...) {}
};
inline void init_thread_pool(std::size_t, std::size_t) {}
inline std::shared_ptr<thread_pool_stub> thread_pool() { return {}; }
inline void register_logger(std::shared_ptr<logger>) {}
inline void shutdown() {}
}  // namespace spdlog
