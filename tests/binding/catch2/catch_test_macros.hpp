#pragma once
#include "catch_amalgamated.hpp"
