// Source-compatibility header: the reference's <goofit/PDFs/physics/detail/NormEvents_4Body_Base.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Physics4.h>
