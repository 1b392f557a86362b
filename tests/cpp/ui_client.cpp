// ui_client — the `remote_render_ui` side of the wire, for tests: connects to a B200StreamServer (or to
// the reference's InterfaceServer) with packetcomms' own TcpSocket / PacketMuxer / PacketDemuxer, does
// the `ready` handshake (Sync.cpp:31-50, syncWithServer), sends the state packets it is given, waits for
// `frames` render_preview + tile_histogram pairs, writes the last frame (raw BGR8) and histogram to
// files, then sends `stop`.
//   ui_client <port> <frames> <out.bgr> <out.hist> [name=value ...]     (float state packets; fov in degrees)
#include <PacketComms.h>
#include <PacketSerialisation.h>
#include <network/TcpSocket.h>
#include <cereal/types/string.hpp>
#include <cereal/types/vector.hpp>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../apps/stream_server.hpp"

int main(int argc, char **argv) {
  if (argc < 5) return 2;
  const int port = std::atoi(argv[1]), frames = std::atoi(argv[2]);
  TcpSocket sock;
  bool ok = false;
  for (int attempt = 0; attempt < 400 && !ok; ++attempt) {   // the server may still be loading its scene
    ok = sock.Connect("127.0.0.1", port);
    if (!ok) std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
  if (!ok) { std::printf("connect failed\n"); return 3; }
  sock.setBlocking(false);
  PacketDemuxer rx(sock, b200ui::packetTypes());
  PacketMuxer tx(sock, b200ui::packetTypes());
  std::mutex mu;
  b200ui::PreviewFrame last;
  std::vector<std::uint32_t> hist;
  std::atomic<int> got_frames{0}, got_hist{0};
  std::atomic<float> fov_echo{-1.f};
  auto s1 = rx.subscribe("render_preview", [&](const ComPacket::ConstSharedPacket &p) {
    b200ui::PreviewFrame f;
    deserialise(p, f);
    std::lock_guard<std::mutex> lock(mu);
    last = std::move(f);
    ++got_frames;
  });
  auto s2 = rx.subscribe("tile_histogram", [&](const ComPacket::ConstSharedPacket &p) {
    std::vector<std::uint32_t> h;
    deserialise(p, h);
    std::lock_guard<std::mutex> lock(mu);
    hist = std::move(h);
    ++got_hist;
  });
  auto s3 = rx.subscribe("fov", [&](const ComPacket::ConstSharedPacket &p) { float v; deserialise(p, v); fov_echo = v; });
  syncWithServer(tx, rx, "ready");
  // the server echoes its fov right after the handshake (InterfaceServer::updateFov, splat.cpp:176-181):
  // wait for it so that a `fov` sent from here is not overwritten by that initial value
  for (int waited = 0; fov_echo < 0.f && waited < 2000; waited += 5) std::this_thread::sleep_for(std::chrono::milliseconds(5));
  for (int i = 5; i < argc; ++i) {   // state packets, as the UI's sliders send them
    const char *eq = std::strchr(argv[i], '=');
    if (!eq) continue;
    serialise(tx, std::string(argv[i], eq - argv[i]), (float)std::atof(eq + 1));
  }
  // the frames rendered BEFORE the state arrived are skipped: wait for `frames` more after a settle time
  std::this_thread::sleep_for(std::chrono::milliseconds(300));
  const int base = got_frames;
  for (int waited = 0; got_frames < base + frames && waited < 20000 && rx.ok(); waited += 5)
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
  const bool complete = got_frames >= base + frames && got_hist > 0;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (std::FILE *f = std::fopen(argv[3], "wb")) { std::fwrite(last.data.data(), 1, last.data.size(), f); std::fclose(f); }
    if (std::FILE *f = std::fopen(argv[4], "wb")) { std::fwrite(hist.data(), 4, hist.size(), f); std::fclose(f); }
    std::printf("%s frames %d hist %d width %d height %d format %u fov_echo %.6f\n", complete ? "ok" : "incomplete", (int)got_frames,
                (int)got_hist, last.width, last.height, last.format, (double)fov_echo.load());
  }
  serialise(tx, "stop", true);
  std::this_thread::sleep_for(std::chrono::milliseconds(200));   // let the muxer flush
  return complete ? 0 : 4;
}
