// Source-compatibility header: the reference's <goofit/PDFs/utilities/VariableBinTransform1DPdf.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Pdfs.h>
