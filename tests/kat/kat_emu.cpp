/* tests/kat/kat_emu.cpp — TEST INFRASTRUCTURE: prng_kat.cu compiled for the host under the lock-step
 * warp emulator (tests/emu/cuda_emu.h), so that the checker itself is exercised on CPU-only machines
 * (sub-ranges of the engine states; the exhaustive run needs the GPU). */
#define COG_EMU 1
#include "prng_kat.cu"
