#pragma once
#include <mcbooster/EvaluateArray.h>
