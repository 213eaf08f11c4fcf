// TEST INFRASTRUCTURE (oracle build only). Stand-in for the header CMake would
// generate from /root/reference/CMakeConfig/templates/yafaray_config.h.cmake:29-30
// with the reference's default option values (CMakeLists.txt: YAF_MIN_RAY_DIST
// 0.00005, YAF_SHADOW_BIAS 0.0005).
#pragma once
#ifndef Y_CONFIG_H
#define Y_CONFIG_H
#include <string>
#define MIN_RAYDIST 0.00005
#define YAF_SHADOW_BIAS 0.0005
#define HAVE_UNISTD_H 1
#define YAFARAY_BUILD_VERSION std::string("oracle-ref")
#define YAFARAY_BUILD_SEARCH_PLUGIN_DIR std::string("")
#define YAFARAY_BUILD_ARCHITECTURE std::string("64bit")
#define YAFARAY_BUILD_COMPILER std::string("GCC")
#define YAFARAY_BUILD_OS std::string("Linux")
#define YAFARAY_BUILD_PLATFORM std::string("")
#endif
