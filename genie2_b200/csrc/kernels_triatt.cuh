// K8 triangle attention (modules/triangular_attention.py:67-130, primitives.py:222-281).
// Off in the base configuration (genie/config.py:62); specialised for 4 heads x 32.
#pragma once
#include "common.cuh"

#define TA_H 4
#define TA_C 32

static inline void triatt_set_attributes() {}
