// Source-compatibility header: the reference's <goofit/DataSet.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Core.h>
