// Constants shared by the symmetric-memory host code and the collective kernels.
#pragma once
#include "../../include/rtdc_b200.h"

// Signal pad at the end of every rank's allocation: [RTDC_SYMM_MAX_BLOCKS][RTDC_SYMM_MAX_WORLD] u32
// arrival counters, written by peers, read by the owner.
#define RTDC_SYMM_PAD_BYTES (RTDC_SYMM_MAX_BLOCKS * RTDC_SYMM_MAX_WORLD * 4)
