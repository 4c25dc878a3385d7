// cutil_inline.h — forwards to the single-file shim (see sdk_shim.h).
#pragma once
#include "sdk_shim.h"
