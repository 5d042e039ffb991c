/* Hand-written stand-in for the header the reference's CMake configure step
 * would generate from src/lib/jp2/grk_config_private.h.cmake.in.
 * Test infrastructure only (used to compile oracle/_ref). */
#pragma once
#define GROK_HAVE_INTTYPES_H 1
#define GRK_PACKAGE_VERSION "9.2.0"
#define GROK_HAVE_FSEEKO 1
#define GROK_HAVE_MALLOC_H
#define GROK_HAVE_ALIGNED_ALLOC
#define GROK_HAVE_MEMALIGN
#define GROK_HAVE_POSIX_MEMALIGN
