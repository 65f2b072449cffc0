// Source-compatibility header: the reference's <goofit/FitManager.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Fit.h>
