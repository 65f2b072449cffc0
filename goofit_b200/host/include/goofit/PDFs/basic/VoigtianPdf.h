// Source-compatibility header: the reference's <goofit/PDFs/basic/VoigtianPdf.h> include path, served by the B200 host API.
#pragma once
#include <goofit/detail/Pdfs.h>
