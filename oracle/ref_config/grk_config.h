/* Hand-written stand-in for the header the reference's CMake configure step
 * would generate from src/lib/jp2/grk_config.h.cmake.in (CMakeLists.txt:9,18-22).
 * Test infrastructure only (used to compile oracle/_ref). */
#pragma once
#define GRK_VERSION_MAJOR 9
#define GRK_VERSION_MINOR 2
#define GRK_VERSION_BUILD 0
#define GROK_PLUGIN_NAME "grokj2k_plugin"
#define AVX2_FOUND ""
#define AVX_FOUND ""
#define SSE4_1_FOUND ""
#define SSE3_FOUND ""
