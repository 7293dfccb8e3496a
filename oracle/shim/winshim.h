// Test-infrastructure shim (NOT product code): the two MSVC-only calls at MTPC/MTPC.cpp:19-20.
#pragma once
#include <cstdio>
static inline char* _itoa(int v, char* buf, int) { std::sprintf(buf, "%d", v); return buf; }
static inline int fopen_s(FILE** f, const char* name, const char* mode) { *f = std::fopen(name, mode); return *f ? 0 : 1; }
