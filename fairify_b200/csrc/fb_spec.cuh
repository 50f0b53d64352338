// fb_spec.cuh -- all compile-time-specialised kernels (the instantiation units include only the part they build).
#pragma once
#include "fb_spec_fwd.cuh"
#include "fb_spec_profile.cuh"
#include "fb_spec_ibp.cuh"
