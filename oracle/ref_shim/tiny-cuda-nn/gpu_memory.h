// ORACLE BUILD STUB (test infrastructure)
#pragma once
#include <vector>
namespace tcnn { template <typename T> class GPUMemory { public: T* data() const { return nullptr; } size_t size() const { return 0; } }; }
