#pragma once
#define GOOFIT_TAG "v2.4.0"
#define GOOFIT_GIT_VERSION "7af21b66"
