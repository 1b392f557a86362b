// A C++ host driving SEVERAL GPUs of one process through the group entry points of the C ABI
// (gsplat_create_group, include/gsplat_b200.h) — the reference's host fixes numIpus = 1
// (src/main/splat.cpp:41-46).  Renders a short orbit of a synthetic scene with the frame stream,
//   group_smoke <mode: views|bands> <devices, e.g. 0,1> <gaussians> <cameras.bin>
// cameras.bin (written by the test, so both hosts use bit-identical matrices): int32 frame count,
// 16 floats projection, then 16 floats model-view per frame, all in glm memory order.
// and prints one line per frame: "frame <k> <sum of BGR bytes> <sum of tile counts> <bands...>",
// then "ok".  The Python test renders the same frames on one context and compares the sums.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <numeric>
#include <string>
#include <vector>

#include "gsplat_b200.h"

int main(int argc, char **argv) {
  if (argc < 5) { std::fprintf(stderr, "usage: group_smoke views|bands dev,dev,... gaussians cameras.bin\n"); return 2; }
  const int mode = std::strcmp(argv[1], "bands") == 0 ? GSPLAT_GROUP_BANDS : GSPLAT_GROUP_VIEWS;
  std::vector<int> devices;
  for (char *tok = std::strtok(argv[2], ","); tok; tok = std::strtok(nullptr, ",")) devices.push_back(std::atoi(tok));
  const std::size_t n = std::strtoull(argv[3], nullptr, 10);
  int frames = 0;
  float proj[16];
  std::vector<float> views;
  {
    std::FILE *fh = std::fopen(argv[4], "rb");
    if (!fh || std::fread(&frames, 4, 1, fh) != 1 || frames < 1 || frames > 4096 || std::fread(proj, 4, 16, fh) != 16) return 2;
    views.resize((std::size_t)frames * 16);
    if (std::fread(views.data(), 4, views.size(), fh) != views.size()) return 2;
    std::fclose(fh);
  }
  const int W = 1280, H = 720;
  std::vector<float> g(n * 16);
  if (gsplat_synth_scene(0xB2000000ull, n, g.data()) != GSPLAT_OK) return 2;
  gsplat_config cfg;
  gsplat_config_default(&cfg);
  cfg.width = W; cfg.height = H;
  gsplat_group *grp = nullptr;
  int rc = gsplat_create_group(&cfg, devices.data(), (int)devices.size(), mode, &grp);
  if (rc) { std::printf("create failed (%d): %s\n", rc, gsplat_group_last_error(nullptr)); return rc == GSPLAT_E_CUDA ? 3 : 4; }
  auto die = [&](const char *what) { std::printf("%s failed: %s\n", what, gsplat_group_last_error(grp)); gsplat_destroy_group(grp); return 5; };
  if (gsplat_group_upload_scene(grp, g.data(), n)) return die("upload");
  const int depth = std::min(gsplat_group_max_outstanding(grp), 4);
  std::deque<std::pair<int, uint64_t>> inflight;
  auto deliver = [&]() -> bool {
    const auto [k, ticket] = inflight.front();
    inflight.pop_front();
    const uint8_t *img = nullptr;
    const uint32_t *counts = nullptr;
    if (gsplat_group_acquire_frame(grp, ticket, &img, &counts)) return false;
    const unsigned long long px = std::accumulate(img, img + (std::size_t)W * H * 3, 0ull);
    const unsigned long long cs = std::accumulate(counts, counts + (W / 32) * (H / 20), 0ull);
    std::printf("frame %d %llu %llu", k, px, cs);
    int cuts[17];
    const int nc = mode == GSPLAT_GROUP_BANDS ? gsplat_group_get_bands(grp, cuts, 17) : 0;
    for (int i = 0; i < nc; ++i) std::printf(" %d", cuts[i]);
    std::printf("\n");
    return true;
  };
  for (int k = 0; k < frames; ++k) {
    if (gsplat_group_set_camera(grp, &views[(std::size_t)k * 16], proj, 0.6981317f, 1.0f)) return die("set_camera");
    uint64_t t = 0;
    if (gsplat_group_render_streamed(grp, &t)) return die("render");
    inflight.emplace_back(k, t);
    if ((int)inflight.size() > depth - 1 && !deliver()) return die("acquire");
  }
  while (!inflight.empty())
    if (!deliver()) return die("acquire");
  gsplat_destroy_group(grp);
  std::printf("ok\n");
  return 0;
}
