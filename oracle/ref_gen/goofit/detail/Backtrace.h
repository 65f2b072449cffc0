#pragma once
#define Backtrace_FOUND 1
#include <execinfo.h>
