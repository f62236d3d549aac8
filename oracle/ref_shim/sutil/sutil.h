// TEST INFRASTRUCTURE (oracle reference build): stand-in for the absent <sutil/sutil.h>; see ref_shim_base.h
#pragma once
#include "../ref_shim_base.h"
