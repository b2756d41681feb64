// narrowBand.h -- forwards to the libmrb shim of the reference API (see mesh_reconstruction_compat.h).
#pragma once
#include "mesh_reconstruction_compat.h"
