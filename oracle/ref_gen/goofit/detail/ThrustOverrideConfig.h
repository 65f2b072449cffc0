#pragma once
#define OR_GROUPSIZE 128
#define OR_GRAINSIZE 7
