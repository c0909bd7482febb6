// compat/optix_stubs.h — intentionally empty: see compat/optix.h (not OptiX).
#pragma once
#include "optix.h"
