// The reference's tests include <catch2/catch_all.hpp> from a CMake-built Catch2 (CMakeLists.txt:88-97).  Without CMake the
// vendored single-file distribution ext/Catch2/extras/catch_amalgamated.{hpp,cpp} is compiled instead (oracle/Makefile, -I extras).
#pragma once
#include "catch_amalgamated.hpp"
