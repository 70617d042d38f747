// oracle/shim/fnl_capi.cpp - TEST INFRASTRUCTURE ONLY.
// The reference's vendored FastNoiseLite (source/soillib/external/FastNoiseLite.h),
// compiled where it lies, configured as op/noise.hpp:16-32 does.  Used by
// oracle/make_golden.py to pin the oracle's and the product's own OpenSimplex2
// restatements.
#pragma GCC diagnostic ignored "-Waggressive-loop-optimizations"
#include <soillib/external/FastNoiseLite.h>
extern "C" void fnl_noise(float* out, int nx, int ny, float frequency, int octaves, float gain, float lacunarity,
                          float seed, float extx, float exty) {
  FastNoiseLite src;
  src.SetNoiseType(FastNoiseLite::NoiseType_OpenSimplex2);
  src.SetFractalType(FastNoiseLite::FractalType_FBm);
  src.SetFrequency(frequency);
  src.SetFractalOctaves(octaves);
  src.SetFractalGain(gain);
  src.SetFractalLacunarity(lacunarity);
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) out[(size_t)x * ny + y] = src.GetNoise(x / extx, y / exty, seed);
}
