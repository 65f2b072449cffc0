// Source-compatibility header: the reference's <goofit/BinnedDataSet.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Core.h>
