// stream_loopback — B200StreamServer without a GPU: a stand-in render loop that follows
// src/main/splat.cpp:245-322 (consumeState, render, sendHistogram + sendPreviewImage, until `stop`) and
// paints the received UI state INTO the frame it sends, so the client side can check that every state
// packet was ingested:  pixel (0,0) = (X, Y, Z) as bytes, pixel (1,0) = (lambda1*10, lambda2*10, gamma*10),
// pixel (2,0) = (env_rotation, env_rotation_2, fov in degrees);  histogram[i] = i * 3 + 1.
//   stream_loopback <port> [width height]
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "../../apps/stream_server.hpp"

static std::uint8_t byte_of(float v) { return (std::uint8_t)std::lround(std::fmin(std::fmax(v, 0.f), 255.f)); }

int main(int argc, char **argv) {
  const int port = argc > 1 ? std::atoi(argv[1]) : 5000;
  const int W = argc > 3 ? std::atoi(argv[2]) : 64, H = argc > 3 ? std::atoi(argv[3]) : 40;
  b200ui::B200StreamServer ui(port);
  ui.start();
  if (!ui.connected()) { std::printf("no client\n"); return 3; }
  ui.initialiseVideoStream(W, H);
  ui.updateFov(0.6981317f);
  std::vector<std::uint8_t> img((std::size_t)W * H * 3);
  std::vector<std::uint32_t> hist((std::size_t)(W / 32) * (H / 20));
  for (std::size_t i = 0; i < hist.size(); ++i) hist[i] = (std::uint32_t)i * 3 + 1;
  int frames = 0;
  b200ui::B200StreamServer::State state;
  do {
    state = ui.consumeState();
    for (std::size_t i = 0; i < img.size(); ++i) img[i] = (std::uint8_t)(i * 7 + frames);
    img[0] = byte_of(state.X); img[1] = byte_of(state.Y); img[2] = byte_of(state.Z);
    img[3] = byte_of(state.lambda1 * 10); img[4] = byte_of(state.lambda2 * 10); img[5] = byte_of(state.gamma * 10);
    img[6] = byte_of(state.envRotationDegrees); img[7] = byte_of(state.envRotationDegrees2);
    img[8] = byte_of(state.fov * 180.f / (float)M_PI);
    ui.sendHistogram(hist);
    ui.sendPreviewImage(img.data(), W, H, (std::size_t)W * 3);
    ++frames;
    std::this_thread::sleep_for(std::chrono::milliseconds(10));
  } while (!state.stop && frames < 5000);
  std::printf("ok frames %d stop %d\n", frames, (int)state.stop);
  return state.stop ? 0 : 4;
}
