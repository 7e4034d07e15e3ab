/* empty stand-in: the reference device code needs nothing from this header (see cuda_on_host.h) */
#pragma once
#include "cuda_on_host.h"
