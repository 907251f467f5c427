/* emu_lib.cpp — TEST INFRASTRUCTURE: builds the package's CUDA sources against the
 * lock-step warp emulator (cuda_emu.h) into tests/emu/libcog_emu.so, which exports the
 * same C ABI as the real library.  Loaded only by tests (`-m "not gpu"`). */
#include "../../gym_eldorado_b200/csrc/cog_capi.cu"

namespace cog {
uint4 cog_smem[4096]; /* 64 KiB of "shared memory" */
}
