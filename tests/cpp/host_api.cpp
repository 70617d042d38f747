// Exercises include/soilb200.hpp the way the reference's bindings would call soil:: (model.cpp:157-292):
// allocating flow ops, in-place erosion ops, the fused step, and the exception mapping.
// Usage: host_api OUTDIR   -> writes direction.i32, steepest.i32, accumulate.f32, normal.f32, height.f32 (raw,
// row-major) for tests/test_gpu_cpp_host.py to compare with the Python surface, prints "OK" on success.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>

#include "soilb200.hpp"

template <class T>
static void dump(const std::string& path, const std::vector<T>& v) {
  std::ofstream f(path, std::ios::binary);
  f.write(reinterpret_cast<const char*>(v.data()), static_cast<std::streamsize>(v.size() * sizeof(T)));
}

#define REQUIRE(cond)                                                        \
  do {                                                                       \
    if (!(cond)) {                                                           \
      std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
      return 1;                                                              \
    }                                                                        \
  } while (0)

int main(int argc, char** argv) {
  const std::string out = argc > 1 ? argv[1] : ".";
  const int nx = 96, ny = 128;
  soil::noise_param_t np;
  np.seed = 3.0f;
  np.ext = {float(nx), float(ny)};
  const auto h = soil::noise(soil::shape{nx, ny}, np);

  // flow graph: allocating ops return new device tensors
  const auto dir = soil::direction(h, soil::D8);
  const auto rec = soil::steepest(h, soil::D8);
  const auto rain = soil::tensor_t<float>::from_host(std::vector<float>(std::size_t(nx) * ny, 1.0f), soil::shape{nx, ny});
  const auto acc = soil::accumulate(rec, rain, soil::D8);
  const auto nrm = soil::op::normal(h, {20.0f / nx, 20.0f / ny, 4.0f});
  const auto rw0 = soil::random_weighted(h, soil::D8, 0, 0, 10.0f);
  const auto rw1 = soil::random_weighted(h, soil::D8, 0, 1, 10.0f);  // continues the cached streams

  const auto mf = soil::multiflow(h, rain, soil::D8, 3, 10.0f).to_host();
  double mean_out = 0.0;

  // gradient family: a ramp has a constant gradient and a zero Laplacian inside
  {
    std::vector<float> ramp(std::size_t(nx) * ny);
    for (int x = 0; x < nx; ++x)
      for (int y = 0; y < ny; ++y) ramp[std::size_t(x) * ny + y] = 0.5f * x + 0.25f * y;
    const auto rd = soil::tensor_t<float>::from_host(ramp, soil::shape{nx, ny});
    for (float v : soil::negslope(rd, {1.0f, 1.0f}).to_host()) REQUIRE(v >= 0.0f);
    const auto gr = soil::gradient(rd, {1.0f, 1.0f}).to_host();
    for (std::size_t n = 0; n < ramp.size(); ++n) REQUIRE(gr[2 * n] == 0.5f && gr[2 * n + 1] == 0.25f);
    const auto r1 = soil::tensor_t<float>::from_host(ramp, soil::shape{nx, ny, 1});
    const auto lap = soil::laplacian(r1, {1.0f, 1.0f}).to_host();
    REQUIRE(lap[std::size_t(5) * ny + 7] == 0.0f);
    const auto sl = soil::slope(rd, soil::steepest(rd, soil::D8), {1.0f, 1.0f}).to_host();
    REQUIRE(sl[0] == 0.0f && sl[std::size_t(5) * ny + 7] < 0.0f);
    auto blurred = soil::gaussian_blur(r1, 2.0f);
    REQUIRE(blurred.data() == r1.data());
  }

  // albedo generators and the Monte-Carlo estimator: shapes, finiteness, the trivial cases
  {
    const std::size_t N = std::size_t(nx) * ny;
    const auto lay = soil::tensor_t<float>::from_host(std::vector<float>(2 * N, 0.0f), soil::shape{nx, ny, 2});
    const auto bed = soil::tensor_t<float>::from_host(std::vector<float>(3 * N, 0.25f), soil::shape{nx, ny, 3});
    const auto sed = soil::tensor_t<float>::from_host(std::vector<float>(3 * N, 0.75f), soil::shape{nx, ny, 3});
    soil::tensor_t<float> alb(soil::shape{nx, ny, 3});
    soil::albedo_layer(alb, bed, sed, lay, 10.0f, {0.0f, 0.0f, 0.0f});
    for (float v : alb.to_host()) REQUIRE(v == 0.25f);               // no sediment: the bedrock colour
    soil::albedo_discharge(alb, rain, {0.0f, 0.0f, 1.0f}, 0.0f, 1.0f);  // zero extinction: untouched
    for (float v : alb.to_host()) REQUIRE(v == 0.25f);
    soil::albedo_stratum(alb, rain, lay, {1.0f, 1.0f, 1.0f}, soil::param_t{}, {1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, 0.0f, 1.0f);
    const auto st = alb.to_host();
    REQUIRE(st[0] == 1.0f && st[1] == 0.0f);                         // depth 0 -> stratum A
    std::vector<float> fl(2 * N, 0.0f);
    for (std::size_t n = 0; n < N; ++n) fl[2 * n + 1] = 2.0f;         // uniform flow along +y
    soil::tensor_t<soil::rng> states(soil::shape{4096});
    soil::seed(states, 1, 0);
    const auto flux = soil::solve_uniform(soil::tensor_t<float>::from_host(fl, soil::shape{nx, ny, 2}),
                                          soil::tensor_t<float>::from_host(std::vector<float>(N, 1.0f), soil::shape{nx, ny, 1}),
                                          soil::tensor_t<float>::from_host(std::vector<float>(N, 0.0f), soil::shape{nx, ny}),
                                          states, {1.0f, 1.0f}, 4096).to_host();
    for (int x = 0; x < nx; ++x) REQUIRE(flux[std::size_t(x) * ny] == 0.5f);   // first column: the local term only
    for (float v : flux) REQUIRE(std::isfinite(v) && v >= 0.5f);
  }

  const auto hd = dir.to_host();
  const auto hr = rec.to_host();
  const auto ha = acc.to_host();
  static const int sx[8] = {-1, 0, 0, 1, -1, -1, 1, 1}, sy[8] = {0, -1, 1, 0, -1, 1, -1, 1};  // graph.hpp:35-44
  double outflow = 0.0;
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const std::size_t n = std::size_t(x) * ny + y;
      if (hd[n] < 0) {
        REQUIRE(hr[n] == -1);
        outflow += ha[n];  // roots collect everything upstream
      } else {
        REQUIRE(hr[n] == (x + sx[hd[n]]) * ny + (y + sy[hd[n]]));
      }
    }
  REQUIRE(outflow == double(nx) * ny);  // every unit of rain leaves through exactly one root
  for (double v : mf) mean_out = v > mean_out ? v : mean_out;
  REQUIRE(mean_out >= 1.0 && mean_out <= double(nx) * ny);
  REQUIRE(rw0.to_host() != rw1.to_host());

  // error mapping: argument errors -> std::invalid_argument
  bool threw = false;
  try {
    soil::direction(h, static_cast<soil::edge_t>(5));
  } catch (const std::invalid_argument&) {
    threw = true;
  }
  REQUIRE(threw);
  threw = false;
  try {
    soil::accumulate(rec, soil::tensor_t<float>(soil::shape{nx, ny + 4}), soil::D8);
  } catch (const std::invalid_argument&) {
    threw = true;
  }
  REQUIRE(threw);

  // the fused step
  std::vector<float> layers(std::size_t(nx) * ny * 2, 0.0f);
  const auto hh = h.to_host();
  for (std::size_t n = 0; n < hh.size(); ++n) layers[2 * n] = hh[n];
  soil::param_t param;  // erosion.hpp defaults
  param.maxage = 128;
  param.timeStep = 10.0f;
  soil::model_t model(soil::tensor_t<float>::from_host(layers, soil::shape{nx, ny, 2}), rain,
                      soil::tensor_t<float>::from_host(std::vector<float>(hh.size(), 0.0f), soil::shape{nx, ny}), {}, {},
                      512, {20.0f / nx, 20.0f / ny, 4.0f});
  for (int i = 0; i < 3; ++i) model.step(param);
  const auto sums = model.mass_sums();
  REQUIRE(std::isfinite(sums[0]) && std::isfinite(sums[1]) && sums[1] >= 0.0);
  const auto l = model.layers().to_host();
  const auto hm = model.height.to_host();
  double sb = 0.0, ss = 0.0;
  for (std::size_t n = 0; n < hm.size(); ++n) {
    REQUIRE(hm[n] == l[2 * n] + l[2 * n + 1]);  // height is the merged map of the current layers
    sb += l[2 * n];
    ss += l[2 * n + 1];
  }
  REQUIRE(std::fabs(sb - sums[0]) <= 1e-6 * std::fabs(sb) + 1e-9 && std::fabs(ss - sums[1]) <= 1e-6 * std::fabs(ss) + 1e-9);

  dump(out + "/height.f32", hh);
  dump(out + "/direction.i32", hd);
  dump(out + "/steepest.i32", hr);
  dump(out + "/accumulate.f32", ha);
  dump(out + "/normal.f32", nrm.to_host());
  std::cout << "OK" << std::endl;
  return 0;
}
