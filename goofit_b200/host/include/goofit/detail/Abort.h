// detail/Abort.h — include/goofit/detail/Abort.h of the reference: GooFit::abort(file, line, reason, pdf) prints the reason and the
// PDF's parameters and throws (src/goofit/Abort.cpp:10-42). Declared with the rest of the host API in detail/Core.h.
#pragma once
#include <goofit/detail/Core.h>
