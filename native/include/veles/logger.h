// veles_b200 native runtime - logging (reference: libVeles/inc/veles/logger.h, a per-class logger with a domain and a
// colour on top of EINA's log domains).  Here: no dependency; every Logger has a domain string and a colour, levels
// are critical < error < warning < info < debug, the threshold is global with per-domain overrides, and records go to
// stderr (coloured when it is a terminal) or to a sink callback an embedding application installs.
//
// Environment (read once, on first use):
//   VELES_LOG_LEVEL=debug|info|warning|error|critical|0..4      global threshold (default: warning)
//   VELES_LOG_LEVELS=WorkflowLoader:debug,Workflow:info         per-domain thresholds
#pragma once
#include <cstdarg>
#include <string>

namespace veles {

enum class LogLevel : int { kCritical = 0, kError = 1, kWarning = 2, kInfo = 3, kDebug = 4 };

// sink(level, domain, message, user): called under the logger's lock, one record at a time
typedef void (*LogSink)(int level, const char* domain, const char* message, void* user);

class Logger {
 public:
  explicit Logger(const std::string& domain = "default", const std::string& color = "") noexcept;

  const std::string& domain() const noexcept { return domain_; }
  void set_domain(const std::string& d) noexcept { domain_ = d; }
  const std::string& color() const noexcept { return color_; }
  void set_color(const std::string& c) noexcept { color_ = c; }

  bool enabled(LogLevel level) const noexcept;
  void log(LogLevel level, const char* fmt, ...) const noexcept __attribute__((format(printf, 3, 4)));
  void vlog(LogLevel level, const char* fmt, va_list ap) const noexcept;

  // process-wide configuration
  static void SetLevel(LogLevel level) noexcept;
  static LogLevel Level() noexcept;
  static void SetDomainLevel(const std::string& domain, LogLevel level) noexcept;
  static void ClearDomainLevels() noexcept;
  static void SetSink(LogSink sink, void* user) noexcept;   // nullptr: back to stderr
  static bool ParseLevel(const std::string& text, LogLevel* out) noexcept;
  static const char* LevelName(LogLevel level) noexcept;
  // "veles::WorkflowLoader" -> "WorkflowLoader": domain names for loggers derived from a class name
  static std::string Demangle(const char* symbol) noexcept;

 private:
  std::string domain_;
  std::string color_;
};

// ANSI colour prefixes for Logger(color)
extern const char* const kLogColorCyan;
extern const char* const kLogColorGreen;
extern const char* const kLogColorYellow;
extern const char* const kLogColorMagenta;

}  // namespace veles

#define VELES_LOG_AT(logger, lvl, ...)                                                  \
  do {                                                                                  \
    if ((logger).enabled(lvl)) (logger).log(lvl, __VA_ARGS__);                          \
  } while (0)
#define VELES_DBG(logger, ...) VELES_LOG_AT(logger, ::veles::LogLevel::kDebug, __VA_ARGS__)
#define VELES_INF(logger, ...) VELES_LOG_AT(logger, ::veles::LogLevel::kInfo, __VA_ARGS__)
#define VELES_WRN(logger, ...) VELES_LOG_AT(logger, ::veles::LogLevel::kWarning, __VA_ARGS__)
#define VELES_ERR(logger, ...) VELES_LOG_AT(logger, ::veles::LogLevel::kError, __VA_ARGS__)
#define VELES_CRT(logger, ...) VELES_LOG_AT(logger, ::veles::LogLevel::kCritical, __VA_ARGS__)
