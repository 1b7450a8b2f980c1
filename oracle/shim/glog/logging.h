// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// Minimal stand-in for <glog/logging.h> so that the reference's own sources
// (read in place from /root/reference, never copied) can be compiled by
// oracle/Makefile without running the reference's cmake build (which is what
// generates the real glog headers).  Only the macro surface the reference
// actually uses is provided: LOG/VLOG/CHECK*/DCHECK* + the FLAGS_* globals
// that src/main/pbrt.cpp touches.  Logging below FATAL is discarded unless
// FLAGS_minloglevel admits it and FLAGS_logtostderr is set, which keeps the
// per-sample LOG(INFO) in src/portals/aaportal.cpp:155 from costing anything
// but the stream formatting the reference itself pays for.
#ifndef ORACLE_SHIM_GLOG_LOGGING_H
#define ORACLE_SHIM_GLOG_LOGGING_H

#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

extern int FLAGS_v;
extern int FLAGS_minloglevel;
extern int FLAGS_stderrthreshold;
extern bool FLAGS_logtostderr;
extern std::string FLAGS_log_dir;

namespace google {

enum { GLOG_INFO = 0, GLOG_WARNING = 1, GLOG_ERROR = 2, GLOG_FATAL = 3 };

inline void InitGoogleLogging(const char *) {}

// Swallows everything streamed into it.
struct NullStream {
    template <typename T>
    NullStream &operator<<(const T &) { return *this; }
    NullStream &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};

class LogMessage {
  public:
    LogMessage(const char *file, int line, int severity)
        : severity_(severity) {
        static const char tag[4] = {'I', 'W', 'E', 'F'};
        os_ << tag[severity] << " " << file << ":" << line << "] ";
    }
    ~LogMessage() {
        if (severity_ >= FLAGS_minloglevel &&
            (severity_ >= FLAGS_stderrthreshold || FLAGS_logtostderr ||
             severity_ == GLOG_FATAL)) {
            os_ << "\n";
            std::string s = os_.str();
            fwrite(s.data(), 1, s.size(), stderr);
        }
        if (severity_ == GLOG_FATAL) {
            fflush(stderr);
            abort();
        }
    }
    std::ostream &stream() { return os_; }

  private:
    int severity_;
    std::ostringstream os_;
};

// Lets "cond ? (void)0 : Voidify() & stream" type-check.
struct Voidify {
    void operator&(std::ostream &) {}
};

}  // namespace google

#define SHIM_SEV_INFO ::google::GLOG_INFO
#define SHIM_SEV_WARNING ::google::GLOG_WARNING
#define SHIM_SEV_ERROR ::google::GLOG_ERROR
#define SHIM_SEV_FATAL ::google::GLOG_FATAL

#define LOG(sev) \
    ::google::LogMessage(__FILE__, __LINE__, SHIM_SEV_##sev).stream()

#define VLOG_IS_ON(n) (FLAGS_v >= (n))
#define VLOG(n)                   \
    !VLOG_IS_ON(n) ? (void)0      \
                   : ::google::Voidify() & LOG(INFO)

#define CHECK(cond)                                \
    (cond) ? (void)0                               \
           : ::google::Voidify() &                 \
                 LOG(FATAL) << "Check failed: " #cond " "

#define SHIM_CHECK_OP(a, b, op)                                           \
    ((a)op(b)) ? (void)0                                                  \
               : ::google::Voidify() & LOG(FATAL)                         \
                     << "Check failed: " #a " " #op " " #b " (" << (a)     \
                     << " vs. " << (b) << ") "

#define CHECK_EQ(a, b) SHIM_CHECK_OP(a, b, ==)
#define CHECK_NE(a, b) SHIM_CHECK_OP(a, b, !=)
#define CHECK_LT(a, b) SHIM_CHECK_OP(a, b, <)
#define CHECK_LE(a, b) SHIM_CHECK_OP(a, b, <=)
#define CHECK_GT(a, b) SHIM_CHECK_OP(a, b, >)
#define CHECK_GE(a, b) SHIM_CHECK_OP(a, b, >=)
#define CHECK_NOTNULL(p) (p)

#ifdef NDEBUG
#define DCHECK(cond) \
    true ? (void)0 : ::google::Voidify() & LOG(FATAL) << !(cond)
#define SHIM_DCHECK_OP(a, b, op) \
    true ? (void)0 : ::google::Voidify() & LOG(FATAL) << ((a)op(b))
#define DCHECK_EQ(a, b) SHIM_DCHECK_OP(a, b, ==)
#define DCHECK_NE(a, b) SHIM_DCHECK_OP(a, b, !=)
#define DCHECK_LT(a, b) SHIM_DCHECK_OP(a, b, <)
#define DCHECK_LE(a, b) SHIM_DCHECK_OP(a, b, <=)
#define DCHECK_GT(a, b) SHIM_DCHECK_OP(a, b, >)
#define DCHECK_GE(a, b) SHIM_DCHECK_OP(a, b, >=)
#else
#define DCHECK(cond) CHECK(cond)
#define DCHECK_EQ(a, b) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) CHECK_NE(a, b)
#define DCHECK_LT(a, b) CHECK_LT(a, b)
#define DCHECK_LE(a, b) CHECK_LE(a, b)
#define DCHECK_GT(a, b) CHECK_GT(a, b)
#define DCHECK_GE(a, b) CHECK_GE(a, b)
#endif

#endif  // ORACLE_SHIM_GLOG_LOGGING_H
