// Stand-in for <mrpt/system/CTimeLogger.h>: section timers are no-ops in the oracle/_ref build
#pragma once
#include <string>
namespace mrpt::system {
class CTimeLogger {
 public:
  CTimeLogger(bool = true, const std::string& = "") {}
  void enter(const char*) {}
  void leave(const char*) {}
  void enable(bool = true) {}
  void clear() {}
};
struct CTimeLoggerEntry {
  CTimeLoggerEntry(CTimeLogger&, const char*) {}
  void stop() {}
};
}  // namespace mrpt::system
