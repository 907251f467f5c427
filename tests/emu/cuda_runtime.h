/* tests/emu/cuda_runtime.h — TEST INFRASTRUCTURE: what the product's `#include <cuda_runtime.h>` finds when
 * tests/emu compiles gym_eldorado_b200/csrc with g++ (see cuda_emu.h). */
#include "cuda_emu.h"
