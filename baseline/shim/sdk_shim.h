// sdk_shim.h — a ~150-line stand-in for the parts of the 2012 GPU Computing SDK utility libraries (cutil, shrUtils,
// shrQATest) that the reference's UNMODIFIED reduction sample calls, so that reduction.cpp / reduction_kernel.cu from
// /root/reference compile with CUDA 12.9 for sm_100 (the vendored libcutil/libshrutil are 2012 x86_64 objects and
// libshrutil drags in rendercheckGL — SURVEY §2.3 C5). Written from the call sites, not copied from the SDK:
// symbols needed = what `grep -o '\b(cut|shr|cutil)[A-Za-z_]*' reduction.cpp` lists.
// Behaviour kept where it affects the measurement: the stopwatch is a gettimeofday wall clock averaged over
// start/stop sessions (stopwatch_linux.h:86-164); errors print and exit(-1); the QA protocol prints PASSED/FAILED.
#pragma once
#include <cuda_runtime.h>
#include <strings.h>
#include <sys/time.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

enum CUTBoolean { CUTFalse = 0, CUTTrue = 1 };
enum eQAstatus { QA_FAILED = 0, QA_PASSED = 1, QA_WAIVED = 2 };
enum { LOGCONSOLE = 1, LOGFILE = 2, LOGBOTH = 3, MASTER = 32 };

// ---- command line: "--key=value" / "--flag" (one or two leading dashes)
inline const char* shim_find_arg(int argc, const char** argv, const char* name, bool* found) {
  const size_t n = strlen(name);
  *found = false;
  for (int i = 1; i < argc; ++i) {
    const char* a = argv[i];
    while (*a == '-') ++a;
    if (!strncasecmp(a, name, n) && (a[n] == '\0' || a[n] == '=')) {
      *found = true;
      return a[n] == '=' ? a + n + 1 : nullptr;
    }
  }
  return nullptr;
}
inline CUTBoolean cutCheckCmdLineFlag(int argc, const char** argv, const char* name) {
  bool f;
  shim_find_arg(argc, argv, name, &f);
  return f ? CUTTrue : CUTFalse;
}
inline CUTBoolean cutGetCmdLineArgumenti(int argc, const char** argv, const char* name, int* val) {
  bool f;
  const char* v = shim_find_arg(argc, argv, name, &f);
  if (!f || !v) return CUTFalse;
  *val = atoi(v);
  return CUTTrue;
}
inline CUTBoolean cutGetCmdLineArgumentstr(int argc, const char** argv, const char* name, char** val) {
  bool f;
  const char* v = shim_find_arg(argc, argv, name, &f);
  if (!f || !v) { *val = nullptr; return CUTFalse; }
  *val = strdup(v);
  return CUTTrue;
}

// ---- stopwatch registry
struct ShimTimer { double total_ms = 0; int sessions = 0; timeval t0; bool running = false; };
inline std::vector<ShimTimer>& shim_timers() { static std::vector<ShimTimer> t; return t; }
inline CUTBoolean cutCreateTimer(unsigned int* id) { shim_timers().push_back(ShimTimer()); *id = (unsigned)shim_timers().size(); return CUTTrue; }
inline CUTBoolean cutDeleteTimer(unsigned int) { return CUTTrue; }
inline CUTBoolean cutStartTimer(unsigned int id) { ShimTimer& t = shim_timers()[id - 1]; gettimeofday(&t.t0, nullptr); t.running = true; return CUTTrue; }
inline CUTBoolean cutStopTimer(unsigned int id) {
  ShimTimer& t = shim_timers()[id - 1];
  timeval t1;
  gettimeofday(&t1, nullptr);
  t.total_ms += 1000.0 * (t1.tv_sec - t.t0.tv_sec) + 0.001 * (t1.tv_usec - t.t0.tv_usec);
  t.sessions++;
  t.running = false;
  return CUTTrue;
}
inline CUTBoolean cutResetTimer(unsigned int id) { shim_timers()[id - 1] = ShimTimer(); return CUTTrue; }
inline float cutGetAverageTimerValue(unsigned int id) { ShimTimer& t = shim_timers()[id - 1]; return t.sessions ? (float)(t.total_ms / t.sessions) : 0.0f; }

// ---- error handling
#define cutilCheckError(x) do { if ((x) != CUTTrue) { fprintf(stderr, "CUTIL error at %s:%d\n", __FILE__, __LINE__); exit(-1); } } while (0)
#define cutilSafeCallNoSync(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s(%d) : CUDA Runtime API error %d: %s.\n", __FILE__, __LINE__, (int)e_, cudaGetErrorString(e_)); exit(-1); } } while (0)
#define cutilSafeCall(x) cutilSafeCallNoSync(x)
#define cutilCheckMsg(msg) do { cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { fprintf(stderr, "%s(%d) : cutilCheckMsg() CUTIL CUDA error : %s : %s.\n", __FILE__, __LINE__, msg, cudaGetErrorString(e_)); exit(-1); } } while (0)
inline void cutilDeviceSynchronize() { cutilSafeCallNoSync(cudaDeviceSynchronize()); }
inline void cutilDeviceReset() { cudaDeviceReset(); }
inline void cutilDeviceInit(int argc, char** argv) {
  int n = 0, dev = 0;
  cutilSafeCallNoSync(cudaGetDeviceCount(&n));
  if (n == 0) { fprintf(stderr, "cutil error: no devices supporting CUDA.\n"); exit(-1); }
  cutGetCmdLineArgumenti(argc, (const char**)argv, "device", &dev);
  if (dev < 0) dev = 0;
  if (dev > n - 1) { fprintf(stderr, "cutilDeviceInit (Device=%d) invalid GPU device.  %d GPU device(s) detected.\n", dev, n); exit(-1); }
  cutilSafeCallNoSync(cudaSetDevice(dev));
}

// ---- logging + QA protocol
inline FILE*& shim_logfile() { static FILE* f = nullptr; return f; }
inline void shrSetLogFileName(const char* name) { if (shim_logfile()) fclose(shim_logfile()); shim_logfile() = fopen(name, "w"); }
inline int shrLogEx(int mode, int, const char* fmt, ...) {
  char buf[4096];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
  if (mode & LOGCONSOLE) { fputs(buf, stdout); fflush(stdout); }
  if ((mode & LOGFILE) && shim_logfile()) { fputs(buf, shim_logfile()); fflush(shim_logfile()); }
  if (mode & MASTER) { FILE* m = fopen("SdkMasterLog.csv", "a"); if (m) { fputs(buf, m); fclose(m); } }
  return 0;
}
inline int shrLog(const char* fmt, ...) {
  char buf[4096];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
  return shrLogEx(LOGBOTH, 0, "%s", buf);
}
inline std::string& shim_exe() { static std::string s; return s; }
inline void shrQAStart(int, char** argv) {
  std::string s = argv[0];
  size_t p = s.find_last_of('/');
  shim_exe() = p == std::string::npos ? s : s.substr(p + 1);
  printf("[%s] starting...\n", shim_exe().c_str());
  fflush(stdout);
}
inline void shrQAFinishExit(int, const char**, int status) {
  const char* w = status == QA_PASSED ? "PASSED" : (status == QA_WAIVED ? "WAIVED" : "FAILED");
  printf("[%s] test results...\n%s\n", shim_exe().c_str(), w);
  fflush(stdout);
  exit(status == QA_FAILED ? EXIT_FAILURE : EXIT_SUCCESS);  // the SDK's 3-second exit delay (--prompt) is not reproduced
}
