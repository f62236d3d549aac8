// TEST INFRASTRUCTURE (oracle reference build): stand-in for the absent <sutil/vec_math.h>; see ref_shim_base.h
#pragma once
#include "../ref_shim_base.h"
