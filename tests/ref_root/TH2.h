#pragma once
#include "TH2F.h"
