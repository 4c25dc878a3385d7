// bench_common.h — utilities shared by the native benchmark binaries.
// Own restatement of what the reference's driver links from the SDK:
//   * flag parser  : `--key=value` / `--flag`; every argument must start with '-' (cmd_arg_reader.cpp:44-64,118-151;
//                    cutCheckCmdLineFlag / cutGetCmdLineArgumenti / ...str, cutil.cpp:1134-1286)
//   * stopwatch    : host wall clock averaged over sessions (stopwatch_linux.h:86-164) PLUS CUDA-event timing
//   * log + QA     : console + log file + optional master CSV (shrUtils.cpp:173-260), and the
//                    "[name] starting..." / PASSED|FAILED|WAIVED sentinel with exit code (shrQATest.h:83-229)
#pragma once
#include <cuda_runtime.h>
#include <strings.h>

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "b2r.h"

namespace b2rbench {

// ---------------------------------------------------------------- flags
class Flags {
 public:
  // Returns false (and fills error()) if an argument does not start with '-'.
  bool parse(int argc, char** argv) {
    for (int i = 1; i < argc; ++i) {
      const char* a = argv[i];
      if (a[0] != '-') {
        err_ = std::string("Invalid command line argument: ") + a;
        return false;
      }
      while (*a == '-') ++a;
      const char* eq = strchr(a, '=');
      if (eq) kv_[std::string(a, eq - a)] = std::string(eq + 1);
      else kv_[a] = "";
    }
    return true;
  }
  bool has(const std::string& k) const { return kv_.count(k) != 0; }
  std::string str(const std::string& k, const std::string& dflt = "") const {
    auto it = kv_.find(k);
    return it == kv_.end() ? dflt : it->second;
  }
  long long i64(const std::string& k, long long dflt) const {
    auto it = kv_.find(k);
    if (it == kv_.end() || it->second.empty()) return dflt;
    return strtoll(it->second.c_str(), nullptr, 0);
  }
  double f64(const std::string& k, double dflt) const {
    auto it = kv_.find(k);
    if (it == kv_.end() || it->second.empty()) return dflt;
    return strtod(it->second.c_str(), nullptr);
  }
  const std::string& error() const { return err_; }

 private:
  std::map<std::string, std::string> kv_;
  std::string err_;
};

// ---------------------------------------------------------------- timers
class HostTimer {  // average over start/stop sessions, in milliseconds
 public:
  void start() { t0_ = std::chrono::steady_clock::now(); }
  void stop() {
    total_ms_ += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    ++sessions_;
  }
  double average_ms() const { return sessions_ ? total_ms_ / sessions_ : 0.0; }
  double total_ms() const { return total_ms_; }
  void reset() { total_ms_ = 0; sessions_ = 0; }

 private:
  std::chrono::steady_clock::time_point t0_;
  double total_ms_ = 0;
  int sessions_ = 0;
};

class EventTimer {  // device time between two events on a stream
 public:
  EventTimer() { cudaEventCreate(&a_); cudaEventCreate(&b_); }
  ~EventTimer() { cudaEventDestroy(a_); cudaEventDestroy(b_); }
  void start(cudaStream_t s) { cudaEventRecord(a_, s); }
  void stop(cudaStream_t s) { cudaEventRecord(b_, s); }
  double ms() {
    cudaEventSynchronize(b_);
    float m = 0;
    cudaEventElapsedTime(&m, a_, b_);
    return (double)m;
  }

 private:
  cudaEvent_t a_, b_;
};

// ---------------------------------------------------------------- log + QA sentinel
class Log {
 public:
  void open(const char* path) { if (path && *path) f_ = fopen(path, "a"); }
  void open_master(const char* path) { if (path && *path) m_ = fopen(path, "a"); }
  ~Log() { if (f_) fclose(f_); if (m_) fclose(m_); }
  void print(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vprint(false, fmt, ap);
    va_end(ap);
  }
  void print_master(const char* fmt, ...) {  // console + log file + master CSV (LOGBOTH | MASTER)
    va_list ap;
    va_start(ap, fmt);
    vprint(true, fmt, ap);
    va_end(ap);
  }

 private:
  void vprint(bool master, const char* fmt, va_list ap) {
    char buf[2048];
    vsnprintf(buf, sizeof(buf), fmt, ap);
    fputs(buf, stdout);
    fflush(stdout);
    if (f_) { fputs(buf, f_); fflush(f_); }
    if (master && m_) { fputs(buf, m_); fflush(m_); }
  }
  FILE* f_ = nullptr;
  FILE* m_ = nullptr;
};

enum QAStatus { QA_FAILED = 0, QA_PASSED = 1, QA_WAIVED = 2 };

struct QA {
  std::string name, cmdline;
  bool qatest = false;
  void start(int argc, char** argv, const Flags& f) {
    name = argv[0];
    size_t slash = name.find_last_of('/');
    if (slash != std::string::npos) name = name.substr(slash + 1);
    for (int i = 0; i < argc; ++i) cmdline += std::string(i ? " " : "") + argv[i];
    qatest = f.has("qatest");
    if (qatest) fprintf(stderr, "&&&& RUNNING %s\n", cmdline.c_str());
    else printf("[%s] starting...\n", name.c_str());
    fflush(stdout);
  }
  int finish(QAStatus s) {  // returns the process exit code
    const char* w = s == QA_PASSED ? "PASSED" : (s == QA_WAIVED ? "WAIVED" : "FAILED");
    if (qatest) fprintf(stderr, "&&&& %s %s\n", w, cmdline.c_str());
    else printf("[%s] test results...\n%s\n", name.c_str(), w);
    fflush(stdout);
    return s == QA_FAILED ? 1 : 0;
  }
};

#define BENCH_CUDA(expr)                                                                              \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess) {                                                                          \
      fprintf(stderr, "%s:%d: %s -> %s\n", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));        \
      exit(2);                                                                                        \
    }                                                                                                 \
  } while (0)
#define BENCH_B2R(expr)                                                                               \
  do {                                                                                                \
    b2r_status_t _s = (expr);                                                                         \
    if (_s != B2R_OK) {                                                                               \
      fprintf(stderr, "%s:%d: %s -> %s (%s)\n", __FILE__, __LINE__, #expr, b2r_status_string(_s),     \
              b2r_last_error_string());                                                               \
      exit(2);                                                                                        \
    }                                                                                                 \
  } while (0)

inline bool parse_dtype(const std::string& s, b2r_dtype_t* d) {
  if (!strcasecmp(s.c_str(), "int") || !strcasecmp(s.c_str(), "int32") || !strcasecmp(s.c_str(), "i32")) *d = B2R_INT32;
  else if (!strcasecmp(s.c_str(), "float") || !strcasecmp(s.c_str(), "fp32") || !strcasecmp(s.c_str(), "f32")) *d = B2R_FLOAT32;
  else if (!strcasecmp(s.c_str(), "double") || !strcasecmp(s.c_str(), "fp64") || !strcasecmp(s.c_str(), "f64")) *d = B2R_FLOAT64;
  else if (!strcasecmp(s.c_str(), "bf16") || !strcasecmp(s.c_str(), "bfloat16")) *d = B2R_BFLOAT16;
  else return false;
  return true;
}
inline bool parse_op(const std::string& s, b2r_op_t* o) {
  if (!strcasecmp(s.c_str(), "sum")) *o = B2R_SUM;
  else if (!strcasecmp(s.c_str(), "min")) *o = B2R_MIN;
  else if (!strcasecmp(s.c_str(), "max")) *o = B2R_MAX;
  else return false;
  return true;
}
inline int parse_variant(const std::string& s) {
  if (s == "1" || s == "ldg128") return B2R_VARIANT_LDG128;
  if (s == "2" || s == "ldg256") return B2R_VARIANT_LDG256;
  if (s == "3" || s == "tma") return B2R_VARIANT_TMA;
  return B2R_VARIANT_AUTO;
}
inline const char* variant_name(int v) {
  switch (v) {
    case B2R_VARIANT_LDG128: return "ldg128";
    case B2R_VARIANT_LDG256: return "ldg256";
    case B2R_VARIANT_TMA: return "tma";
  }
  return "auto";
}

// Rotating input buffers: enough copies of an n-byte input that consecutive timed launches never find their data
// in the 126 MB L2. Returns the number of copies.
inline int rotating_copies(size_t bytes, size_t min_total = (size_t)512 << 20, int max_copies = 64) {
  if (bytes == 0) return 1;
  size_t c = (min_total + bytes - 1) / bytes;
  if (c < 2) c = 2;
  if (c > (size_t)max_copies) c = max_copies;
  return (int)c;
}

}  // namespace b2rbench
