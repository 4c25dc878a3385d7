// bench_common.h — utilities shared by the native benchmark binaries.
// Own restatement of what the reference's driver links from the SDK:
//   * flag parser  : `--key=value` / `--flag`; every argument must start with '-' (cmd_arg_reader.cpp:44-64,118-151;
//                    cutCheckCmdLineFlag / cutGetCmdLineArgumenti / ...str, cutil.cpp:1134-1286)
//   * stopwatch    : host wall clock averaged over sessions (stopwatch_linux.h:86-164) PLUS CUDA-event timing
//   * log + QA     : console + log file + optional master CSV (shrUtils.cpp:173-260), and the
//                    "[name] starting..." / PASSED|FAILED|WAIVED sentinel with exit code (shrQATest.h:83-229)
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvml.h>
#include <strings.h>

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <thread>
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

// ---------------------------------------------------------------- clocks + NVLink byte counters (NVML through dlopen)
// Every number a harness prints is accompanied by a clocks record taken DURING its timed regions (B200_PROFILING.md:
// a number measured under hw_slowdown / thermal slowdown, or with SM clocks stuck low for no reason, is rejected). A
// background thread polls SM clock + clock-event reasons of every visible GPU about once per millisecond; the harness
// brackets its timed loops with ClockLog::Window, and print() emits one line per GPU:
//   # clocks gpu0: sm_mhz median 1965 min 1965 (max 1965), samples 412, reasons [none]
// libnvidia-ml.so.1 is loaded with dlopen so the binaries still start on a machine without the driver library.
// The same handle reads the per-GPU NVLink data counters (NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX/RX, KiB, all links): the
// measured bytes a collective really moved over NVLink, to set against its algorithmic bytes — what ncu cannot give for a
// kernel that waits on its peers (a replayed launch would deadlock).
class ClockLog {
 public:
  static ClockLog& get() {
    static ClockLog c;
    return c;
  }
  struct Window {  // RAII: samples taken while at least one Window is alive count
    Window() { ClockLog::get().active_.fetch_add(1); }
    ~Window() { ClockLog::get().active_.fetch_sub(1); }
  };
  bool ok() const { return ok_; }
  void start(int ndev) {
    if (started_ || !load()) return;
    ndev_ = std::min(ndev, 16);
    for (int i = 0; i < ndev_; ++i) {
      // CUDA ordinal -> NVML handle by PCI bus id (CUDA_VISIBLE_DEVICES may renumber)
      char bus[32] = {0};
      if (cudaDeviceGetPCIBusId(bus, sizeof(bus), i) != cudaSuccess || by_bus_(bus, &dev_[i]) != NVML_SUCCESS) {
        if (by_index_((unsigned)i, &dev_[i]) != NVML_SUCCESS) { ndev_ = i; break; }
      }
      unsigned mx = 0;
      if (max_clock_(dev_[i], NVML_CLOCK_SM, &mx) == NVML_SUCCESS) max_mhz_[i] = mx;
    }
    started_ = true;
    stop_ = false;
    th_ = std::thread([this] { run(); });
  }
  void stop() {
    if (!started_) return;
    stop_ = true;
    if (th_.joinable()) th_.join();
    started_ = false;
  }
  ~ClockLog() { stop(); }
  // One line per GPU with the samples taken inside windows so far. Returns false if any bad reason was seen.
  bool print(FILE* f = stdout) {
    if (!ok_) {
      fprintf(f, "# clocks: unavailable (%s)\n", err_.c_str());
      return true;
    }
    std::lock_guard<std::mutex> lk(mu_);
    bool good = true;
    for (int i = 0; i < ndev_; ++i) {
      std::vector<unsigned> v = mhz_[i];
      if (v.empty()) {
        fprintf(f, "# clocks gpu%d: no samples inside a timed region (max %u MHz)\n", i, max_mhz_[i]);
        continue;
      }
      std::sort(v.begin(), v.end());
      const unsigned long long r = reasons_[i];
      std::string names;
      auto add = [&](unsigned long long bit, const char* n) { if (r & bit) names += std::string(names.empty() ? "" : ",") + n; };
      add(nvmlClocksEventReasonSwPowerCap, "sw_power_cap");
      add(nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown");
      add(nvmlClocksEventReasonSwThermalSlowdown, "sw_thermal_slowdown");
      add(nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown");
      add(nvmlClocksThrottleReasonHwPowerBrakeSlowdown, "hw_power_brake_slowdown");
      add(nvmlClocksEventReasonApplicationsClocksSetting, "applications_clocks_setting");
      const bool bad = r & (nvmlClocksThrottleReasonHwSlowdown | nvmlClocksEventReasonSwThermalSlowdown | nvmlClocksThrottleReasonHwThermalSlowdown);
      good = good && !bad;
      fprintf(f, "# clocks gpu%d: sm_mhz median %u min %u (max %u), samples %zu, reasons [%s]%s\n", i, v[v.size() / 2], v.front(), max_mhz_[i],
              v.size(), names.empty() ? "none" : names.c_str(), bad ? "  REJECT: slowdown seen during a timed region" : "");
    }
    return good;
  }
  // Cumulative NVLink payload bytes (all links) sent / received by device i; false if the counters cannot be read.
  bool nvlink_bytes(int i, unsigned long long* tx, unsigned long long* rx) {
    if (!ok_ || i >= ndev_ || !field_values_) return false;
    nvmlFieldValue_t fv[2];
    memset(fv, 0, sizeof(fv));
    fv[0].fieldId = NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX;
    fv[1].fieldId = NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX;
    fv[0].scopeId = fv[1].scopeId = 0xffffffffu;  // all links
    if (field_values_(dev_[i], 2, fv) != NVML_SUCCESS || fv[0].nvmlReturn != NVML_SUCCESS || fv[1].nvmlReturn != NVML_SUCCESS) return false;
    *tx = fv[0].value.ullVal * 1024ull;
    *rx = fv[1].value.ullVal * 1024ull;
    return true;
  }

 private:
  bool load() {
    void* h = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!h) { err_ = "libnvidia-ml.so.1 not found"; return false; }
    auto sym = [&](const char* n) { return dlsym(h, n); };
    auto init = (nvmlReturn_t(*)())sym("nvmlInit_v2");
    by_index_ = (nvmlReturn_t(*)(unsigned, nvmlDevice_t*))sym("nvmlDeviceGetHandleByIndex_v2");
    by_bus_ = (nvmlReturn_t(*)(const char*, nvmlDevice_t*))sym("nvmlDeviceGetHandleByPciBusId_v2");
    clock_ = (nvmlReturn_t(*)(nvmlDevice_t, nvmlClockType_t, unsigned*))sym("nvmlDeviceGetClockInfo");
    max_clock_ = (nvmlReturn_t(*)(nvmlDevice_t, nvmlClockType_t, unsigned*))sym("nvmlDeviceGetMaxClockInfo");
    reasons_fn_ = (nvmlReturn_t(*)(nvmlDevice_t, unsigned long long*))sym("nvmlDeviceGetCurrentClocksEventReasons");
    if (!reasons_fn_) reasons_fn_ = (nvmlReturn_t(*)(nvmlDevice_t, unsigned long long*))sym("nvmlDeviceGetCurrentClocksThrottleReasons");
    field_values_ = (nvmlReturn_t(*)(nvmlDevice_t, int, nvmlFieldValue_t*))sym("nvmlDeviceGetFieldValues");
    if (!init || !by_index_ || !by_bus_ || !clock_ || !max_clock_ || !reasons_fn_) { err_ = "NVML symbols missing"; return false; }
    const nvmlReturn_t r = init();
    if (r != NVML_SUCCESS) { err_ = "nvmlInit_v2 failed with " + std::to_string((int)r); return false; }
    ok_ = true;
    return true;
  }
  void run() {
    while (!stop_) {
      if (active_.load() > 0) {
        for (int i = 0; i < ndev_; ++i) {
          unsigned mhz = 0;
          unsigned long long r = 0;
          if (clock_(dev_[i], NVML_CLOCK_SM, &mhz) != NVML_SUCCESS) continue;
          reasons_fn_(dev_[i], &r);
          std::lock_guard<std::mutex> lk(mu_);
          mhz_[i].push_back(mhz);
          reasons_[i] |= r;
        }
      }
      std::this_thread::sleep_for(std::chrono::microseconds(active_.load() > 0 ? 500 : 200));
    }
  }
  bool ok_ = false, started_ = false;
  std::atomic<bool> stop_{false};
  std::atomic<int> active_{0};
  std::string err_;
  int ndev_ = 0;
  nvmlDevice_t dev_[16];
  unsigned max_mhz_[16] = {0};
  std::vector<unsigned> mhz_[16];
  unsigned long long reasons_[16] = {0};
  std::mutex mu_;
  std::thread th_;
  nvmlReturn_t (*by_index_)(unsigned, nvmlDevice_t*) = nullptr;
  nvmlReturn_t (*by_bus_)(const char*, nvmlDevice_t*) = nullptr;
  nvmlReturn_t (*clock_)(nvmlDevice_t, nvmlClockType_t, unsigned*) = nullptr;
  nvmlReturn_t (*max_clock_)(nvmlDevice_t, nvmlClockType_t, unsigned*) = nullptr;
  nvmlReturn_t (*reasons_fn_)(nvmlDevice_t, unsigned long long*) = nullptr;
  nvmlReturn_t (*field_values_)(nvmlDevice_t, int, nvmlFieldValue_t*) = nullptr;
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
