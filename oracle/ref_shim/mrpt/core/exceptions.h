// Stand-in for <mrpt/core/exceptions.h>: the assertion / exception macros libopenbeam uses (oracle/_ref build only).
#pragma once
// the standard headers libopenbeam gets transitively through MRPT / Eigen
#include <algorithm>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <optional>
#include <set>
#include <vector>
#include <cmath>
#include <sstream>
#include <stdexcept>
#include <string>
#define OB_SHIM_THROW(msg_)                                                     \
  do {                                                                          \
    std::ostringstream ob_shim_s_;                                              \
    ob_shim_s_ << msg_ << " [" << __FILE__ << ":" << __LINE__ << "]";           \
    throw std::runtime_error(ob_shim_s_.str());                                 \
  } while (0)
#define THROW_EXCEPTION(msg_) OB_SHIM_THROW(msg_)
#define THROW_EXCEPTION_FMT(fmt_, ...) OB_SHIM_THROW(mrpt::format(fmt_, __VA_ARGS__))
#define ASSERT_(cond_)                                     \
  do {                                                     \
    if (!(cond_)) OB_SHIM_THROW("Assert condition failed: " #cond_); \
  } while (0)
#define ASSERTMSG_(cond_, msg_)                            \
  do {                                                     \
    if (!(cond_)) OB_SHIM_THROW(msg_);                     \
  } while (0)
#define ASSERT_EQUAL_(a_, b_) ASSERT_((a_) == (b_))
#define ASSERT_GE_(a_, b_) ASSERT_((a_) >= (b_))
#define ASSERT_GT_(a_, b_) ASSERT_((a_) > (b_))
#define ASSERT_LE_(a_, b_) ASSERT_((a_) <= (b_))
#define ASSERT_LT_(a_, b_) ASSERT_((a_) < (b_))
#define ASSERT_NEAR_(a_, b_, eps_) ASSERT_(std::abs((a_) - (b_)) <= (eps_))
#define MRPT_TODO(x_)
#define MRPT_START
#define MRPT_END
#define MRPT_UNUSED_PARAM(x_) (void)(x_)
