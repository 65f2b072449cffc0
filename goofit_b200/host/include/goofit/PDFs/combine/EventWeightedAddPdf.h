// EventWeightedAddPdf (reference include/goofit/PDFs/combine/EventWeightedAddPdf.h): see goofit/detail/Pdfs.h
#pragma once
#include <goofit/detail/Pdfs.h>
