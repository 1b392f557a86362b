// Host-side use of the drop-in boundary from C++, written the way src/main/splat.cpp drives
// IpuSplatter (:53-66 construct, :259-265 per-frame update + execute, :131-144 read-back).
// Prints one line: "ok <D> <sum of BGR bytes> <sum of tile counts>", or "no-device: <message>"
// (exit code 3) when there is no GPU — there is no CPU fallback to hide behind.
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

#include "gsplat_b200.hpp"

int main(int argc, char **argv) {
  const std::size_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 20000;
  const int W = 640, H = 360;
  std::vector<splat::Gaussian3DRecord> g(n);
  if (gsplat_synth_scene(0xB2000000ull, n, reinterpret_cast<float *>(g.data())) != GSPLAT_OK) return 2;
  // matrices arrive as 16 floats in glm order on stdin-free argv-free form: identity view pulled
  // back along -z, simple frustum — values only need to be the same as the Python twin's.
  float view[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, -9.0f, 1};
  const float nr = 4.0f, fr = 40.0f, hw = 4.0f * W / H, hh = 4.0f;
  float proj[16] = {2 * nr / (2 * hw), 0, 0, 0, 0, 2 * nr / (2 * hh), 0, 0, 0, 0, -(fr + nr) / (fr - nr), -1,
                    0, 0, -(2 * fr * nr) / (fr - nr), 0};
  try {
    splat::B200Splatter s(g, W, H);
    s.updateModelView(view);
    s.updateProjection(proj);
    s.updateFocalLengths(0.6981317f, 1.0f);
    s.execute();
    std::vector<std::uint8_t> img;
    std::vector<std::uint32_t> counts;
    s.getFrameBuffer(img);
    s.getIPUHistogram(counts);
    const unsigned long long px = std::accumulate(img.begin(), img.end(), 0ull);
    const unsigned long long cs = std::accumulate(counts.begin(), counts.end(), 0ull);
    std::printf("ok %llu %llu %llu\n", (unsigned long long)s.stats().num_duplicates, px, cs);
  } catch (const std::runtime_error &e) {
    std::printf("no-device: %s\n", e.what());
    return 3;
  }
  return 0;
}
