// Test infrastructure, see catch_all.hpp (Catch::Session lives there).
#pragma once
#include <catch2/catch_all.hpp>
