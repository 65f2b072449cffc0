// Source-compatibility header: the reference's pre-2.2 <goofit/PDFs/physics/ThreeGaussResolution_Aux.h> include path (there it forwards to the header without _Aux).
#pragma once
#include <goofit/detail/Physics4.h>
