// tests/ref_root stand-in (see TH1.h)
#pragma once
#define ROOT_VERSION(a, b, c) (((a) << 16) + ((b) << 8) + (c))
#define ROOT_VERSION_CODE ROOT_VERSION(6, 30, 0)
#define ROOT_RELEASE "6.30/00-standin"
