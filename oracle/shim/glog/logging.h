// TEST INFRASTRUCTURE (oracle/): stand-in for <glog/logging.h>; ivox3d.h only writes LOG(ERROR) << "..." on an unknown stencil.
#pragma once
#include <iostream>
namespace flivo_shim { struct NullLog { template <class T> NullLog& operator<<(const T&) { return *this; } }; }
#define LOG(severity) ::flivo_shim::NullLog()
