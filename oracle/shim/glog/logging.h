// TEST INFRASTRUCTURE ONLY (oracle build): glog is absent; the hot path uses
// none of it, math_util.h merely includes it and geometry.h has DCHECK_NE in
// templates that the scan never instantiates.
#ifndef UMRS_ORACLE_GLOG_SHIM
#define UMRS_ORACLE_GLOG_SHIM
#include <iostream>
struct UmrsNullLog { template <typename T> UmrsNullLog& operator<<(const T&) { return *this; } };
#define DCHECK_NE(a, b) UmrsNullLog()
#define DCHECK_EQ(a, b) UmrsNullLog()
#define DCHECK_GT(a, b) UmrsNullLog()
#define DCHECK_GE(a, b) UmrsNullLog()
#define DCHECK_LT(a, b) UmrsNullLog()
#define DCHECK_LE(a, b) UmrsNullLog()
#define DCHECK(a) UmrsNullLog()
#define CHECK(a) UmrsNullLog()
#define CHECK_NE(a, b) UmrsNullLog()
#define CHECK_EQ(a, b) UmrsNullLog()
#define CHECK_GT(a, b) UmrsNullLog()
#define CHECK_GE(a, b) UmrsNullLog()
#define CHECK_LT(a, b) UmrsNullLog()
#define CHECK_LE(a, b) UmrsNullLog()
#define CHECK_NOTNULL(a) (a)
#define LOG(x) UmrsNullLog()
#endif
