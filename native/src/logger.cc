// veles_b200 native runtime - logging implementation (see veles/logger.h).
#include "veles/logger.h"

#include <cxxabi.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace veles {

const char* const kLogColorCyan = "\033[36m";
const char* const kLogColorGreen = "\033[32m";
const char* const kLogColorYellow = "\033[33m";
const char* const kLogColorMagenta = "\033[35m";

namespace {

struct State {
  std::mutex mutex;
  LogLevel level = LogLevel::kWarning;
  std::map<std::string, LogLevel> domains;
  LogSink sink = nullptr;
  void* sink_user = nullptr;
  bool tty = false;
  std::chrono::steady_clock::time_point start = std::chrono::steady_clock::now();
};

State& state() {
  static State* s = [] {
    State* st = new State();     // leaked on purpose: loggers of static objects may outlive main()
    st->tty = isatty(fileno(stderr)) != 0;
    if (const char* env = getenv("VELES_LOG_LEVEL")) {
      LogLevel l;
      if (Logger::ParseLevel(env, &l)) st->level = l;
    }
    if (const char* env = getenv("VELES_LOG_LEVELS")) {
      std::string all(env);
      size_t pos = 0;
      while (pos <= all.size()) {
        size_t end = all.find(',', pos);
        if (end == std::string::npos) end = all.size();
        const std::string item = all.substr(pos, end - pos);
        const size_t colon = item.rfind(':');
        LogLevel l;
        if (colon != std::string::npos && Logger::ParseLevel(item.substr(colon + 1), &l))
          st->domains[item.substr(0, colon)] = l;
        pos = end + 1;
      }
    }
    return st;
  }();
  return *s;
}

const char* level_color(LogLevel l) {
  switch (l) {
    case LogLevel::kCritical: return "\033[1;31m";
    case LogLevel::kError: return "\033[31m";
    case LogLevel::kWarning: return "\033[33m";
    case LogLevel::kInfo: return "\033[32m";
    default: return "\033[37m";
  }
}

}  // namespace

Logger::Logger(const std::string& domain, const std::string& color) noexcept : domain_(domain), color_(color) {}

bool Logger::enabled(LogLevel level) const noexcept {
  State& s = state();
  if (s.domains.empty()) return static_cast<int>(level) <= static_cast<int>(s.level);
  std::lock_guard<std::mutex> lock(s.mutex);
  auto it = s.domains.find(domain_);
  const LogLevel limit = it != s.domains.end() ? it->second : s.level;
  return static_cast<int>(level) <= static_cast<int>(limit);
}

void Logger::log(LogLevel level, const char* fmt, ...) const noexcept {
  va_list ap;
  va_start(ap, fmt);
  vlog(level, fmt, ap);
  va_end(ap);
}

void Logger::vlog(LogLevel level, const char* fmt, va_list ap) const noexcept {
  if (!enabled(level)) return;
  char stack_buf[512];
  std::vector<char> heap;
  char* msg = stack_buf;
  va_list ap2;
  va_copy(ap2, ap);
  const int n = vsnprintf(stack_buf, sizeof(stack_buf), fmt, ap);
  if (n >= static_cast<int>(sizeof(stack_buf))) {
    heap.resize(static_cast<size_t>(n) + 1);
    vsnprintf(heap.data(), heap.size(), fmt, ap2);
    msg = heap.data();
  } else if (n < 0) {
    snprintf(stack_buf, sizeof(stack_buf), "(unformattable log record: %s)", fmt);
  }
  va_end(ap2);
  State& s = state();
  std::lock_guard<std::mutex> lock(s.mutex);
  if (s.sink != nullptr) {
    s.sink(static_cast<int>(level), domain_.c_str(), msg, s.sink_user);
    return;
  }
  const double t = std::chrono::duration<double>(std::chrono::steady_clock::now() - s.start).count();
  if (s.tty)
    fprintf(stderr, "%s%-8s\033[0m %9.3f %s%s\033[0m: %s\n", level_color(level), LevelName(level), t,
            color_.empty() ? "" : color_.c_str(), domain_.c_str(), msg);
  else
    fprintf(stderr, "%-8s %9.3f %s: %s\n", LevelName(level), t, domain_.c_str(), msg);
}

void Logger::SetLevel(LogLevel level) noexcept {
  State& s = state();
  std::lock_guard<std::mutex> lock(s.mutex);
  s.level = level;
}

LogLevel Logger::Level() noexcept { return state().level; }

void Logger::SetDomainLevel(const std::string& domain, LogLevel level) noexcept {
  State& s = state();
  std::lock_guard<std::mutex> lock(s.mutex);
  s.domains[domain] = level;
}

void Logger::ClearDomainLevels() noexcept {
  State& s = state();
  std::lock_guard<std::mutex> lock(s.mutex);
  s.domains.clear();
}

void Logger::SetSink(LogSink sink, void* user) noexcept {
  State& s = state();
  std::lock_guard<std::mutex> lock(s.mutex);
  s.sink = sink;
  s.sink_user = user;
}

bool Logger::ParseLevel(const std::string& text, LogLevel* out) noexcept {
  std::string t(text);
  std::transform(t.begin(), t.end(), t.begin(), [](unsigned char c) { return static_cast<char>(tolower(c)); });
  static const struct { const char* name; LogLevel level; } names[] = {
      {"critical", LogLevel::kCritical}, {"crit", LogLevel::kCritical}, {"error", LogLevel::kError},
      {"err", LogLevel::kError}, {"warning", LogLevel::kWarning}, {"warn", LogLevel::kWarning},
      {"info", LogLevel::kInfo}, {"debug", LogLevel::kDebug}, {"dbg", LogLevel::kDebug}};
  for (const auto& e : names)
    if (t == e.name) {
      *out = e.level;
      return true;
    }
  if (t.size() == 1 && t[0] >= '0' && t[0] <= '4') {
    *out = static_cast<LogLevel>(t[0] - '0');
    return true;
  }
  return false;
}

const char* Logger::LevelName(LogLevel level) noexcept {
  switch (level) {
    case LogLevel::kCritical: return "CRITICAL";
    case LogLevel::kError: return "ERROR";
    case LogLevel::kWarning: return "WARNING";
    case LogLevel::kInfo: return "INFO";
    default: return "DEBUG";
  }
}

std::string Logger::Demangle(const char* symbol) noexcept {
  int status = 0;
  char* d = abi::__cxa_demangle(symbol, nullptr, nullptr, &status);
  std::string name = (status == 0 && d != nullptr) ? d : symbol;
  free(d);
  const size_t pos = name.rfind("::");
  return pos == std::string::npos ? name : name.substr(pos + 2);
}

}  // namespace veles
