// Source-compatibility header: the reference's <goofit/PDFs/physics/Lineshapes.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Physics4.h>
