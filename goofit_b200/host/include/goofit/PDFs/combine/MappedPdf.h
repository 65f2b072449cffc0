// MappedPdf (reference include/goofit/PDFs/combine/MappedPdf.h): see goofit/detail/Pdfs.h
#pragma once
#include <goofit/detail/Pdfs.h>
