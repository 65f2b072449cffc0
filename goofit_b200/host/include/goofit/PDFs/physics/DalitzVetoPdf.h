// Source-compatibility header: the reference's <goofit/PDFs/physics/DalitzVetoPdf.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Physics3.h>
