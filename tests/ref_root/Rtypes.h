// tests/ref_root stand-in (see TH1.h): ROOT's basic typedefs
#pragma once
typedef int Int_t;
typedef unsigned int UInt_t;
typedef double Double_t;
typedef float Float_t;
typedef bool Bool_t;
typedef long Long_t;
typedef char Char_t;
