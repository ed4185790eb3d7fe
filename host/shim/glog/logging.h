// host/shim/glog/logging.h -- build shim (not part of the GPU core).
//
// Minimal stand-in for <glog/logging.h> so that the reference's own source
// files (compiled in place from /root/reference by host/Makefile) build
// without glog's cmake-generated headers.  It implements only the macro
// surface the reference uses: LOG(sev), VLOG(n), CHECK*/DCHECK*, the FLAGS_*
// globals touched by src/main/pbrt.cpp:78-132 and google::InitGoogleLogging.
// It carries no path arithmetic.
#ifndef SPTX_GLOG_SHIM_H
#define SPTX_GLOG_SHIM_H

#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

namespace sptx_glog_shim {

enum Severity { INFO = 0, WARNING = 1, ERROR = 2, FATAL = 3 };

extern int min_log_level;  // defined in host/shim/shim_stubs.cpp

class Message {
  public:
    Message(const char *file, int line, int sev) : sev_(sev) {
        static const char *names[] = {"I", "W", "E", "F"};
        os_ << names[sev] << " " << file << ":" << line << "] ";
    }
    ~Message() {
        if (sev_ >= min_log_level || sev_ == FATAL) {
            os_ << "\n";
            std::cerr << os_.str();
        }
        if (sev_ == FATAL) std::abort();
    }
    std::ostream &stream() { return os_; }

  private:
    std::ostringstream os_;
    int sev_;
};

// Swallows a stream expression: lets `cond ? (void)0 : Voidify() & stream`
// type-check, the classic glog trick.
struct Voidify {
    void operator&(std::ostream &) {}
};

}  // namespace sptx_glog_shim

namespace google {
inline void InitGoogleLogging(const char *) {}
}  // namespace google

extern int FLAGS_stderrthreshold;
extern int FLAGS_minloglevel;
extern int FLAGS_v;
extern bool FLAGS_logtostderr;
extern std::string FLAGS_log_dir;

#define SPTX_LOG_SEV_INFO ::sptx_glog_shim::INFO
#define SPTX_LOG_SEV_WARNING ::sptx_glog_shim::WARNING
#define SPTX_LOG_SEV_ERROR ::sptx_glog_shim::ERROR
#define SPTX_LOG_SEV_FATAL ::sptx_glog_shim::FATAL

#define LOG(sev) \
    ::sptx_glog_shim::Message(__FILE__, __LINE__, SPTX_LOG_SEV_##sev).stream()

#define VLOG(n)                  \
    (FLAGS_v < (n)) ? (void)0    \
                    : ::sptx_glog_shim::Voidify() & LOG(INFO)

#define CHECK(cond)                                   \
    (cond) ? (void)0                                  \
           : ::sptx_glog_shim::Voidify() &            \
                 LOG(FATAL) << "Check failed: " #cond " "

#define SPTX_CHECK_OP(a, op, b)                                      \
    ((a)op(b)) ? (void)0                                             \
               : ::sptx_glog_shim::Voidify() &                       \
                     LOG(FATAL) << "Check failed: " #a " " #op " " #b \
                                << " (" << (a) << " vs. " << (b) << ") "

#define CHECK_EQ(a, b) SPTX_CHECK_OP(a, ==, b)
#define CHECK_NE(a, b) SPTX_CHECK_OP(a, !=, b)
#define CHECK_LT(a, b) SPTX_CHECK_OP(a, <, b)
#define CHECK_LE(a, b) SPTX_CHECK_OP(a, <=, b)
#define CHECK_GT(a, b) SPTX_CHECK_OP(a, >, b)
#define CHECK_GE(a, b) SPTX_CHECK_OP(a, >=, b)
#define CHECK_NEAR(a, b, eps) \
    CHECK(((a) - (b)) <= (eps) && ((b) - (a)) <= (eps))
#define CHECK_NOTNULL(p) (p)

// The reference is built with NDEBUG (CMakeLists.txt:64): DCHECKs compile
// but never evaluate, exactly as in glog.
#ifdef NDEBUG
#define DCHECK(cond) \
    while (false) CHECK(cond)
#define DCHECK_EQ(a, b) \
    while (false) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) \
    while (false) CHECK_NE(a, b)
#define DCHECK_LT(a, b) \
    while (false) CHECK_LT(a, b)
#define DCHECK_LE(a, b) \
    while (false) CHECK_LE(a, b)
#define DCHECK_GT(a, b) \
    while (false) CHECK_GT(a, b)
#define DCHECK_GE(a, b) \
    while (false) CHECK_GE(a, b)
#else
#define DCHECK(cond) CHECK(cond)
#define DCHECK_EQ(a, b) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) CHECK_NE(a, b)
#define DCHECK_LT(a, b) CHECK_LT(a, b)
#define DCHECK_LE(a, b) CHECK_LE(a, b)
#define DCHECK_GT(a, b) CHECK_GT(a, b)
#define DCHECK_GE(a, b) CHECK_GE(a, b)
#endif

#endif  // SPTX_GLOG_SHIM_H
