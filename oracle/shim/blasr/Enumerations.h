// TEST INFRASTRUCTURE: blasr_libcpp stub header (submodule missing from the reference).
#pragma once
#include "Types.h"
