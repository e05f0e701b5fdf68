// Large path (C5) kernels -- see ba_large.cuh.
#include "ba_large.cuh"
