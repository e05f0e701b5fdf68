// Large path (C5, BAL-style): multi-kernel LM with a block-sparse reduced camera system and PCG.
#pragma once
#include "ba_math.cuh"
#include "../../include/b2ba.h"

namespace b2 {

struct LargeState {
  bool ready = false;
};

}  // namespace b2
