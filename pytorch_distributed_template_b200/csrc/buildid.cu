// Identity of this build: a hash of every source and header that went into the library, passed in by
// build.py.  The Python binding compares it with the hash of the sources it finds next to itself, so a
// library left over from an older checkout (same ABI number, different kernel bodies) is rebuilt or refused
// instead of silently used.
#include "../../include/pdt_b200.h"

#ifndef PDT_BUILD_ID
#define PDT_BUILD_ID "unknown"
#endif

extern "C" const char* pdt_build_id(void) { return PDT_BUILD_ID; }
