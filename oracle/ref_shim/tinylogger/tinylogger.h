// ORACLE BUILD STUB (test infrastructure): dependencies/tinylogger ships without its header in the reference mount.
#pragma once
#include <iostream>
namespace tlog { struct Null { template <class T> Null& operator<<(const T&) { return *this; } };
inline Null info() { return {}; } inline Null warning() { return {}; } inline Null error() { return {}; } inline Null success() { return {}; } inline Null debug() { return {}; } }
