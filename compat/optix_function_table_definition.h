// compat/optix_function_table_definition.h — intentionally empty: see compat/optix.h (not OptiX).
#pragma once
#include "optix.h"
